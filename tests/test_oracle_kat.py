"""CPU tests (no GPU): pin the ORACLE (oracle/oracle.cpp) against the reference's own known-answer
vectors, committed under tests/golden/ by tests/golden/make_golden.py, plus the
analytic checks that fill the gap where the reference has no radar test."""
import ctypes as C
import json
import math
import os

import numpy as np
import pytest

from beifong_b200 import _native as N
from beifong_b200 import configs, meshes
from beifong_b200.scene import FlatScene
from oracle_lib import OracleScene, lib, fvec, EMBREE

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_kats.json")))
f32 = C.c_float


def mesh_scene(pos, faces, extra_rects=()):
    """A scene holding one mesh (+ rectangles): enough for ray queries through the oracle."""
    s = FlatScene(spectrum_width=1)
    b = s.add_diffuse(0.5, two_sided=False)
    if pos is not None:
        s.add_mesh(pos, faces, b)
    ids = [s.add_rect(m, b) for m in extra_rects]
    return s, ids


# ---- sampler: TEA + PCG32 (include/mitsuba/core/random.h:74-116; enoki/random.h is absent) -----------------
def test_tea_float32_reference_kat():
    for a, b, r, v in GOLD["tea_float32"]["cases"]:      # src/libcore/tests/test_random.py:9-16
        assert lib().orc_tea_float32(a, b, r) == np.float32(v)


def test_tea_float64_reference_kat():
    for a, b, r, v in GOLD["tea_float64"]["cases"]:      # src/libcore/tests/test_random.py:22-29
        assert lib().orc_tea_float64(a, b, r) == v


def test_pcg32_public_demo_vector():
    """External KAT (the reference pins `independent` only against Enoki's own PCG32, test_independent.py:29-34):
    pcg32-demo of the public PCG reference, seed (42, 54): first six 32-bit outputs."""
    st, inc = C.c_uint64(), C.c_uint64()
    lib().orc_pcg32_seed(42, 54, C.byref(st), C.byref(inc))
    got = [lib().orc_pcg32_next(C.byref(st), C.byref(inc)) for _ in range(6)]
    assert got == [0xa15c02b7, 0x7b47f409, 0xba1d3330, 0x83d2f293, 0xbfa4784b, 0xcbed606e]


def test_pcg32_default_stream_and_float():
    # default-constructed PCG32 (state 0x853c49e6748fea9b, stream 0xda3e39cb94b95bdb) == seed(default, default)
    st, inc = C.c_uint64(), C.c_uint64()
    lib().orc_pcg32_seed(0x853c49e6748fea9b, 0xda3e39cb94b95bdb, C.byref(st), C.byref(inc))
    assert inc.value == ((0xda3e39cb94b95bdb << 1) | 1) & 0xFFFFFFFFFFFFFFFF
    s2, i2 = C.c_uint64(st.value), C.c_uint64(inc.value)
    u = lib().orc_pcg32_next(C.byref(st), C.byref(inc))
    f = lib().orc_pcg32_next_float(C.byref(s2), C.byref(i2))
    expect = np.array([(u >> 9) | 0x3f800000], np.uint32).view(np.float32)[0] - np.float32(1.0)
    assert f == expect and 0.0 <= f < 1.0


# ---- warps (include/mitsuba/core/warp.h:54-90,325-349) -----------------------------------------------------
def test_concentric_disk_reference_kat():                   # src/libcore/tests/test_warp.py:68-75
    out = (f32 * 2)()
    lib().orc_square_to_uniform_disk_concentric(0.0, 0.0, out)
    assert np.allclose(out[:], [-1 / math.sqrt(2), -1 / math.sqrt(2)])
    lib().orc_square_to_uniform_disk_concentric(0.5, 0.5, out)
    assert np.allclose(out[:], [0, 0])


def test_cosine_hemisphere_reference_kat():                 # src/libcore/tests/test_warp.py:142-149
    out = (f32 * 3)()
    lib().orc_square_to_cosine_hemisphere(0.5, 0.5, out)
    assert np.allclose(out[:], [0, 0, 1])
    lib().orc_square_to_cosine_hemisphere(0.5, 0.0, out)
    assert np.allclose(out[:], [0, -1, 0], atol=1e-7)


def test_coordinate_system_orthonormal():
    rng = np.random.default_rng(5)
    for _ in range(200):
        n = rng.normal(size=3); n /= np.linalg.norm(n)
        s, t = (f32 * 3)(), (f32 * 3)()
        lib().orc_coordinate_system(fvec(*n), s, t)
        s, t = np.array(s[:]), np.array(t[:])
        assert abs(s @ n) < 1e-6 and abs(t @ n) < 1e-6 and abs(s @ t) < 1e-6
        assert np.allclose(np.cross(s, t), n, atol=1e-6)


# ---- "Jacob functions" (include/mitsuba/core/math.h:68-131): analytic anchors ------------------------------
def test_sinc_tri_rect_wchirp():
    L = lib()
    assert L.orc_sinc(0.0) == 1.0 and abs(L.orc_sinc(math.pi)) < 1e-6
    assert L.orc_tri(0.0) == 1.0 and L.orc_tri(0.25) == 0.5 and L.orc_tri(0.5) == 0.0 and L.orc_tri(-0.75) == 0.0
    assert L.orc_rect(0.49) == 1.0 and L.orc_rect(0.5) == 0.0 and L.orc_rect(-0.2) == 1.0
    for T, A in ((200e-6, 1.0), (1e-3, 2.0)):             # wchirp(0, 0, T, A) = 2 A^2 T
        assert L.orc_wchirp(0.0, 0.0, T, A) == pytest.approx(2 * A * A * T, rel=1e-6)
    assert L.orc_wchirp(0.6 * 200e-6, 0.0, 200e-6, 1.0) == 0.0


def test_fmodulo_is_floor_mod():
    rng = np.random.default_rng(1)
    for a, b in zip(rng.uniform(-20, 20, 200), rng.uniform(0.1, 3, 200)):
        got = lib().orc_fmodulo(a, b)
        assert 0 <= got <= b * (1 + 1e-6)
        assert abs(((got - a) / b) - round((got - a) / b)) < 1e-3


# ---- geometry: Möller-Trumbore / rectangle conventions ------------------------------------------------------
def test_rectangle_obj_reference_kat():
    """src/librender/tests/test_mesh.py:257-298: t = 10, prim 0 uv (.35, .3), prim 1 uv (.3, .35)."""
    g = GOLD["rectangle_obj"]
    s, _ = mesh_scene(np.array(g["positions"], np.float32), np.array(g["faces"], np.uint32))
    for accel in (OracleScene.BRUTE, OracleScene.BVH):
        orc = OracleScene(s.export(), accel=accel)
        o = np.array([[-0.3, -0.3, -10.0], [0.3, 0.3, -10.0]], np.float32)
        d = np.array([[0, 0, 1.0], [0, 0, 1.0]], np.float32)
        shape, prim, t, u, v = orc.trace_first_hits(o, d, mint=np.zeros(2, np.float32))
        assert np.allclose(t, 10) and list(prim) == [0, 1]
        assert np.allclose([u[0], v[0]], [0.35, 0.3]) and np.allclose([u[1], v[1]], [0.3, 0.35])
        rc, t0, prim0, si = orc.surface_interaction([-0.3, -0.3, -10.0], [0, 0, 1.0])
        assert rc == 0 and prim0 == 0 and t0 == pytest.approx(10)
        assert np.allclose(si[0:3], [-0.3, -0.3, 0.0], atol=1e-6)       # si.p


def test_rectangle_shape_reference_kat():
    """src/shapes/tests/test_rectangle.py:37-62: 2 x 0.5 rectangle, 15 diagonal rays, 7 hit; ray_test agrees."""
    to_world = np.diag([2.0, 0.5, 1.0, 1.0]).astype(np.float32)
    s, _ = mesh_scene(None, None, [to_world])
    orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
    a = np.linspace(-1, 1, 15).astype(np.float32)
    o = np.stack([a, a, np.full(15, 5, np.float32)], 1)
    d = np.tile(np.array([[0, 0, -1.0]], np.float32), (15, 1))
    shape, prim, t, u, v = orc.trace_first_hits(o, d)
    occ = orc.trace_occluded(o, d)
    hit = shape != N.BF_NONE
    assert np.array_equal(hit, np.abs(a) <= 0.5) and hit.sum() == 7
    assert np.array_equal(occ.astype(bool), hit)
    assert np.allclose(t[hit], 5)


def test_rectangle_surface_area_reference_kat():             # src/shapes/tests/test_rectangle.py:94-126
    cases = [(np.diag([2.0, 2.0, 1.0, 1.0]), 16.0),
             (np.array([[1, 0, 0, 0], [0, 1, 0, 0], [1, 0, 1, 0], [0, 0, 0, 1.0]]), 4 * math.sqrt(2.0)),
             (np.array([[1, 1, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1.0]]), 4.0)]
    for m, area in cases:
        s, ids = mesh_scene(None, None, [m.astype(np.float32)])
        orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
        assert lib().orc_rect_area(C.c_void_p(orc._h), ids[0]) == pytest.approx(area, rel=1e-6)


def test_stairs_reference_kat():
    """src/librender/tests/test_kdtrees.py:25-57: t = 2 - floor(y n)/n, accel == naive, ray_test == hit."""
    n_steps = 20
    p, f = meshes.stairs(n_steps)
    s, _ = mesh_scene(p, f)
    blob = s.export()
    n = 128
    xs, ys = np.meshgrid(np.arange(n - 1) / (n - 1), np.arange(n - 1) / (n - 1), indexing="ij")
    o = np.stack([xs.ravel(), ys.ravel(), np.full(xs.size, 2.0)], 1).astype(np.float32)
    d = np.tile(np.array([[0, 0, -1.0]], np.float32), (len(o), 1))
    mint, maxt = np.zeros(len(o), np.float32), np.full(len(o), 100, np.float32)
    res = {}
    accels = [OracleScene.BRUTE, OracleScene.BVH] + ([OracleScene.EMBREE_ACCEL] if os.path.exists(EMBREE) else [])
    for accel in accels:
        orc = OracleScene(blob, accel=accel)
        res[accel] = orc.trace_first_hits(o, d, mint, maxt)
        assert orc.trace_occluded(o, d, mint, maxt).all()
    t = res[OracleScene.BRUTE][2]
    expected = 2.0 - np.floor(o[:, 1].astype(np.float64) * n_steps) / n_steps
    # a ray exactly on a riser edge may take either step (the reference compares with atol on the naive result)
    on_edge = np.isclose((o[:, 1].astype(np.float64) * n_steps) % 1.0, 0, atol=1e-5)
    assert np.allclose(t[~on_edge], expected[~on_edge], atol=1e-6)
    assert np.array_equal(res[OracleScene.BVH][2], t)                  # kd-tree == naive exactly (compare_results atol 0)
    if OracleScene.EMBREE_ACCEL in res:
        assert np.allclose(res[OracleScene.EMBREE_ACCEL][2], t, atol=1e-5)


def test_bunny_lowres_reference_kat():
    """src/librender/tests/test_kdtrees.py:60-91: 100 x 100 rays along +z from the bbox floor; accel == brute force."""
    g = GOLD["bunny_lowres_ply"]
    p = np.array(g["positions"], np.float32).reshape(-1, 3)
    f = np.array(g["faces"], np.uint32).reshape(-1, 3)
    assert p.shape == (453, 3) and f.shape == (948, 3)
    s, _ = mesh_scene(p, f)
    blob = s.export()
    lo, hi = p.min(0), p.max(0)
    n = 100
    k = np.arange(n) / (n - 1)
    xs, ys = np.meshgrid(lo[0] * (1 - k) + hi[0] * k, lo[1] * (1 - k) + hi[1] * k, indexing="ij")
    o = np.stack([xs.ravel(), ys.ravel(), np.full(xs.size, lo[2])], 1).astype(np.float32)
    d = np.tile(np.array([[0, 0, 1.0]], np.float32), (len(o), 1))
    mint, maxt = np.zeros(len(o), np.float32), np.full(len(o), 100, np.float32)
    brute = OracleScene(blob, accel=OracleScene.BRUTE)
    bvh = OracleScene(blob, accel=OracleScene.BVH)
    a, b = brute.trace_first_hits(o, d, mint, maxt), bvh.trace_first_hits(o, d, mint, maxt)
    for k in range(5):
        assert np.array_equal(a[k].view(np.uint32), b[k].view(np.uint32))
    hit = a[0] != N.BF_NONE
    assert 2000 < hit.sum() < 9000
    assert np.array_equal(brute.trace_occluded(o, d, mint, maxt).astype(bool), hit)
    if os.path.exists(EMBREE):
        e = OracleScene(blob, accel=OracleScene.EMBREE_ACCEL).trace_first_hits(o, d, mint, maxt)
        # bunny_lowres holds coincident duplicate faces (e.g. 530/371, 696/699): equal t, either id may win
        assert np.array_equal(e[0] != N.BF_NONE, hit)
        assert np.allclose(e[2][hit], a[2][hit], rtol=1e-5)
        same = (e[0] == a[0]) & (e[1] == a[1])
        assert same.mean() >= 0.99


# ---- BSDFs --------------------------------------------------------------------------------------------------
def _bsdf_scene():
    s = FlatScene(spectrum_width=1)
    diffuse = s.add_diffuse(0.5, two_sided=False)          # `diffuse` default reflectance 0.5
    two = s.add_diffuse(0.5, two_sided=True)
    eta, k = s.add_uniform(0.0), s.add_uniform(1.0)
    rc = s.add_bsdf(N.BF_BSDF_ROUGHCONDUCTOR, two_sided=True, eta=eta, k=k, alpha_u=0.1, alpha_v=0.1)
    return s, diffuse, two, rc


def test_diffuse_eval_pdf_reference_kat():
    """src/bsdfs/tests/test_diffuse.py:16-39: pdf = cos/pi, eval = 0.5 cos/pi for wi = +z."""
    s, diffuse, two, _ = _bsdf_scene()
    orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
    h = C.c_void_p(orc._h)
    wl = fvec(8.5e6, 0, 0, 0)
    for i in range(20):
        theta = i / 19.0 * (math.pi / 2)
        wo = fvec(math.sin(theta), 0.0, math.cos(theta))
        out = (f32 * 4)()
        lib().orc_bsdf_eval(h, diffuse, fvec(0, 0, 1), wo, wl, out)
        assert np.isclose(lib().orc_bsdf_pdf(h, diffuse, fvec(0, 0, 1), wo, wl), wo[2] / math.pi, atol=1e-7)
        assert np.isclose(out[0], 0.5 * wo[2] / math.pi, atol=1e-7)
        # twosided (src/bsdfs/tests/test_twosided.py:43-90): the back side mirrors the front
        wo_b = fvec(math.sin(theta), 0.0, -math.cos(theta))
        out_b = (f32 * 4)()
        lib().orc_bsdf_eval(h, two, fvec(0, 0, -1), wo_b, wl, out_b)
        assert out_b[0] == out[0]
        lib().orc_bsdf_eval(h, diffuse, fvec(0, 0, -1), wo_b, wl, out_b)
        assert out_b[0] == 0.0                              # one-sided diffuse is black from behind


def _dirs(theta, phi):
    return np.stack([np.cos(phi) * np.sin(theta), np.sin(phi) * np.sin(theta), np.cos(theta)], -1).astype(np.float32)


def test_microfacet_beckmann_eval_reference_kat():
    """src/librender/tests/test_microfacet.py:18-96 (tables from Mitsuba 0.6)."""
    L = GOLD["microfacet/test02_eval_pdf_beckmann"]["lists"]
    steps = 20
    v = _dirs(np.linspace(0, np.pi, steps, dtype=np.float32), np.full(steps, np.pi / 2, np.float32))
    ev = lambda au, av, m: lib().orc_microfacet_eval(N.BF_MICROFACET_BECKMANN, au, av, fvec(*m))
    assert np.allclose([ev(0.1, 0.3, m) for m in v], L[0], rtol=2e-4, atol=1e-8)
    assert np.allclose([ev(0.1, 0.1, m) for m in v], L[2], rtol=2e-4, atol=1e-8)
    # pdf without visible-normal sampling = D * cos(theta_m): positive-hemisphere entries of the pdf tables
    up = v[:, 2] > 0
    assert np.allclose(np.array([ev(0.1, 0.3, m) for m in v])[up] * v[up, 2], np.array(L[1])[up], rtol=2e-4, atol=1e-8)
    v = _dirs(np.full(steps, 0.1, np.float32), np.linspace(0, 2 * np.pi, steps, dtype=np.float32))
    assert np.allclose([ev(0.1, 0.3, m) for m in v], L[4], rtol=2e-4)
    assert np.allclose([ev(0.1, 0.1, m) for m in v], 11.86709118, rtol=2e-4)


def test_microfacet_beckmann_smith_g1_reference_kat():
    """src/librender/tests/test_microfacet.py:98-141."""
    L = GOLD["microfacet/test03_smith_g1_beckmann"]["lists"]
    steps = 20
    g1 = lambda au, av, vv: lib().orc_microfacet_smith_g1(N.BF_MICROFACET_BECKMANN, au, av, fvec(*vv), fvec(0, 0, 1))
    v = _dirs(np.linspace(np.pi / 3, np.pi / 2, steps, dtype=np.float32), np.full(steps, np.pi / 2, np.float32))
    assert np.allclose([g1(0.1, 0.3, m) for m in v], L[0], rtol=1e-3, atol=1e-5)
    assert np.allclose([g1(0.1, 0.1, m) for m in v], L[1], rtol=1e-3, atol=1e-5)
    v = _dirs(np.full(steps, np.pi / 2 * 0.98, np.float32), np.linspace(0, 2 * np.pi, steps, dtype=np.float32))
    assert np.allclose([g1(0.1, 0.3, m) for m in v], L[2], rtol=1e-3)
    assert np.allclose([g1(0.1, 0.1, m) for m in v], 0.67333597, rtol=1e-3)


def test_microfacet_beckmann_sample_reference_kat():
    """src/librender/tests/test_microfacet.py:143-207: mdf(Beckmann, 0.1, 0.3, sample_visible=False).sample on a 6 x 6 grid."""
    L = GOLD["microfacet/test04_sample_beckmann"]["lists"]
    ref_m, ref_pdf = np.array(L[0]).reshape(36, 3), np.array(L[1])
    u = np.linspace(0, 1, 6, dtype=np.float32)
    k = 0
    for u2 in u:
        for u1 in u:
            m, pdf = (f32 * 3)(), f32()
            lib().orc_microfacet_sample(N.BF_MICROFACET_BECKMANN, 0.1, 0.3, 0, fvec(0, 0, 1), u1, u2, m, C.byref(pdf))
            if u1 < 1.0:                                    # u1 = 1 maps to the horizon (log(0)): direction only
                assert np.allclose(m[:], ref_m[k], atol=2e-6), (k, m[:], ref_m[k])
                assert np.isclose(pdf.value, ref_pdf[k], rtol=2e-4, atol=1e-6)
            k += 1


def test_fresnel_conductor_matches_dielectric_for_real_ior():
    """src/librender/tests/test_fresnel.py:63-76: fresnel_conductor(cos, eta + 0i) equals the dielectric
    Fresnel reflectance (closed form evaluated here)."""
    for eta in (1.5, 1 / 1.5):
        for ct in np.cos(np.linspace(0, np.pi / 2, 20)):
            st2 = (1 - ct * ct) / (eta * eta)
            if st2 >= 1:
                expect = 1.0
            else:
                ctt = math.sqrt(1 - st2)
                rs, rp = (ct - eta * ctt) / (ct + eta * ctt), (eta * ct - ctt) / (eta * ct + ctt)
                expect = 0.5 * (rs * rs + rp * rp)
            assert lib().orc_fresnel_conductor(ct, eta, 0.0) == pytest.approx(expect, abs=2e-6)


def test_roughconductor_sample_eval_pdf_consistent():
    """weight * pdf == eval for samples of twosided(roughconductor) (the identity test_rough_conductor's chi2 relies on)."""
    s, _, _, rc = _bsdf_scene()
    orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
    h = C.c_void_p(orc._h)
    wl = fvec(8.5e6, 0, 0, 0)
    rng = np.random.default_rng(3)
    n_ok = 0
    for _ in range(300):
        th = rng.uniform(0, 1.3)
        wi = fvec(math.sin(th), 0, math.cos(th) * rng.choice([-1, 1]))
        wo, pdf, w = (f32 * 3)(), f32(), (f32 * 4)()
        lib().orc_bsdf_sample(h, rc, wi, rng.random(), rng.random(), rng.random(), wl, wo, C.byref(pdf), w)
        if pdf.value <= 0 or w[0] == 0:
            continue
        ev = (f32 * 4)()
        lib().orc_bsdf_eval(h, rc, wi, wo, wl, ev)
        assert lib().orc_bsdf_pdf(h, rc, wi, wo, wl) == pytest.approx(pdf.value, rel=1e-4)
        assert ev[0] == pytest.approx(w[0] * pdf.value, rel=2e-4)
        assert 0 < w[0] <= 1.0 + 1e-5                       # perfect conductor, visible-normal sampling: weight = G1 <= 1
        n_ok += 1
    assert n_ok > 200


# ---- radar-specific arithmetic: no test in the reference, analytic anchors -------------------
def test_signalblock_put_bin_edges():
    """SignalBlock::put, box filter radius 0.5 (signalblock.cpp:115,162-169): pos = pos_ - 0.5, lo = ceil(pos - 0.5)
    = ceil(pos_ - 1), so bin k collects pos_ in (k, k + 1] -- the upper edge is inclusive, pos_ = 0 falls outside."""
    L = lib()
    assert L.orc_block_bin(8, 1, 0, 0, 0.0, 0.5) == -1
    assert L.orc_block_bin(8, 1, 0, 0, 0.001, 0.5) == 0
    assert L.orc_block_bin(8, 1, 0, 0, 1.0, 0.5) == 0
    assert L.orc_block_bin(8, 1, 0, 0, 1.001, 0.5) == 1
    assert L.orc_block_bin(8, 1, 0, 0, 8.0, 0.5) == 7
    assert L.orc_block_bin(8, 1, 0, 0, 8.01, 0.5) == -1 and L.orc_block_bin(8, 1, 0, 0, -0.01, 0.5) == -1
    assert L.orc_block_bin(4, 4, 0, 0, 2.5, 3.5) == 3 * 4 + 2
    assert L.orc_block_bin(4, 4, 2, 1, 2.5, 3.5) == 2 * 4 + 0        # window offset shifts the origin


def test_rect_wigner_on_axis_is_four():
    """Rectangle::sample_wigner (rectangle.cpp:132-220) at the aperture centre, boresight: 4 tri(0)^2 sinc(0)^2 = 4."""
    s = configs.small_radar()
    orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
    h = C.c_void_p(orc._h)
    # shape 0 = receiver aperture at (0, 0, 0.3) facing +x
    w = lib().orc_rect_sample_wigner(h, 0, fvec(0.0, 0.0, 0.3), fvec(1.0, 0.0, 0.0), 3.19e6)
    assert w == pytest.approx(4.0, rel=1e-5)
    # at the aperture edge the triangular taper is zero
    w = lib().orc_rect_sample_wigner(h, 0, fvec(0.0, 0.0025, 0.3), fvec(1.0, 0.0, 0.0), 3.19e6)
    assert abs(w) < 1e-3


def test_phased_single_element_reduces_to_rect_wigner():
    """phasedtransmitter.cpp:261-290 with n_elems = 1, no steering, the element covering the parent rectangle:
    sample_wigner = wid.x wid.y x Rectangle::sample_wigner (the only difference is W_rect_2D's 4 wid.x wid.y prefactor)."""
    s = configs.small_radar(phased=dict(n_elems=1))
    orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
    h = C.c_void_p(orc._h)
    rng = np.random.default_rng(5)
    for _ in range(20):
        y, z = rng.uniform(-0.002, 0.002, 2)
        d = rng.normal(size=3); d[0] = abs(d[0]) + 0.5; d /= np.linalg.norm(d)
        p = fvec(0.0, y, 0.3 + z)
        wp = lib().orc_phased_sample_wigner(h, -1, p, fvec(*d), 3.19e6)
        wr = lib().orc_rect_sample_wigner(h, 0, p, fvec(*d), 3.19e6)
        assert wp == pytest.approx(0.005 * 0.005 * wr, rel=2e-4, abs=1e-9)


def test_phased_two_elements_boresight_sum():
    """n_elems = 2 at the aperture centre, boresight, no steering: of the four virtual elements the two cross terms
    (r_v = 0) are centred on the point (tri = 1) and the two self terms sit at their edge (tri(0.5) = 0); nu_hat is
    normal to r', so every phasor is 1 and W = 2 x 4 wid.x wid.y."""
    s = configs.small_radar(phased=dict(n_elems=2))
    orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
    w = lib().orc_phased_sample_wigner(C.c_void_p(orc._h), -1, fvec(0.0, 0.0, 0.3), fvec(1.0, 0.0, 0.0), 3.19e6)
    assert w == pytest.approx(2 * 4 * 0.0025 * 0.005, rel=1e-4)


def test_phased_steering_moves_the_main_lobe():
    """psi' = exp(i k sin(steer) . (-r')) (phasedtransmitter.cpp:160-165): steering the array tilts the direction of the
    maximum of the cross-term sum away from boresight."""
    def lobe(steer):
        s = configs.small_radar(phased=dict(n_elems=4, steer=steer))
        orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
        h = C.c_void_p(orc._h)
        ang = np.linspace(-0.6, 0.6, 241)
        # the aperture's local y axis is world y; boresight is world +x
        w = [lib().orc_phased_sample_wigner(h, -1, fvec(0.0, 0.0, 0.3), fvec(np.cos(a), np.sin(a), 0.0), 3.19e6) for a in ang]
        return ang[int(np.argmax(w))]
    assert abs(lobe((0.0, 0.0, 0.0))) < 0.01
    assert abs(lobe((0.0, 0.05, 0.0))) > 0.1


def test_phased_radar_profile_is_finite_and_additive():
    s = configs.small_radar(phased=dict(n_elems=4))
    orc = OracleScene(s.export(), accel=OracleScene.BVH)
    a = orc.receive(20000, seed=5, first=0, threads=4)
    b = orc.receive(10000, seed=5, first=0, threads=2) + orc.receive(10000, seed=5, first=10000, threads=2)
    assert np.all(np.isfinite(a)) and a[..., 0].sum() != 0
    # fp32 accumulation order differs between the thread splits and phased weights change sign: compare in norm
    assert np.linalg.norm(a.astype(np.float64) - b) <= 1e-4 * np.linalg.norm(a.astype(np.float64))
    assert a[..., 2].sum() > 0


def test_linfmcw_signal_closed_forms():
    sg = configs.fmcw_signal(False)
    # delta chirp: instantaneous frequency f_c + (B / T)(t - ti) with ti the chirp centre (wignertransmitter.cpp:152-188)
    f0 = lib().orc_sample_delta_frequency(C.byref(sg), 100e-6)
    assert f0 == pytest.approx(94e9, rel=1e-6)
    f1 = lib().orc_sample_delta_frequency(C.byref(sg), 150e-6)
    assert f1 - f0 == pytest.approx(6e9 / 200e-6 * 50e-6, rel=1e-3)
    assert lib().orc_eval_signal(C.byref(sg), 100e-6, f0) > 0


def plate_scene(distance, spectrum_width=1, bins=64, bin_m=0.25):
    """One diffuse plate facing the co-located Tx/Rx apertures at `distance` along +x: all single-bounce energy
    must land in the range bin of 2R / (c dt)."""
    from beifong_b200.meshes import Transform as T
    from beifong_b200.scene import RADAR_94GHZ
    s = FlatScene(spectrum_width=spectrum_width, **RADAR_94GHZ)
    default_bsdf = s.add_diffuse(0.5, two_sided=False)
    plate = s.add_diffuse(1.0, two_sided=True)
    ant = T.translate([0.0, 0.0, 0.0]) * T.rotate([0, 1, 0], 90.0) * T.scale([0.0025, 0.0025, 1.0])
    rx = s.add_rect(ant.matrix32(), default_bsdf)
    tx = s.add_rect((T.translate([0.0, 0.01, 0.0]) * ant).matrix32(), default_bsdf)
    s.add_rect((T.translate([distance, 0, 0]) * T.rotate([0, 1, 0], -90.0) * T.scale([50.0, 50.0, 1.0])).matrix32(), plate)
    tex = s.add_uniform(1.0)
    s.add_transmitter(tx, FlatScene.endpoint(N.BF_TX_WIGNER, texture=tex, signal=configs.fmcw_signal(True)))
    f_bw = bins * bin_m * 2.0 / configs.C0 * (6e9 / 200e-6)
    s.set_receiver(rx, FlatScene.endpoint(N.BF_RX_WIGNER, texture=tex, signal=configs.fmcw_signal(False),
                                          receive_type=N.BF_RECV_MIX_RESAMPLE, adc_sampling_start=20e-6, adc_sampling_time=160e-6))
    s.set_adc(t_bins=1, f_bins=bins, t_bandwidth=200e-6, f_bandwidth=f_bw)
    s.set_integrator(N.BF_GEN_PHASE, max_depth=2, rr_depth=5, phase_bins=1)
    return s


@pytest.mark.parametrize("width", [1, 4])
def test_flat_plate_range_bin(width):
    """Plate at R = 5 m, boresight: the closest return is at round-trip 2R; off-axis returns are longer, so the
    profile is zero below bin floor(R / bin_m) and peaks at it."""
    R, bin_m = 5.0, 0.25
    orc = OracleScene(plate_scene(R, width, bin_m=bin_m).export(), accel=OracleScene.BRUTE)
    prof = orc.receive(200000, seed=1, threads=4)[:, 0, 0]
    k = int(R / bin_m)
    # the beat frequency is a difference of two fp32 values near 94 GHz (ulp 8192 Hz = 0.16 bin), exactly as in
    # integrator.cpp:1603-1619, so returns at R can fall one bin short; nothing may land further down
    assert prof[:k - 1].sum() == 0.0 and prof[k - 1] < 0.2 * prof[k]
    assert prof[k] > 0 and prof.argmax() in (k, k + 1)
    assert prof[k:k + 8].sum() > 0.5 * prof.sum()


def test_oracle_accels_agree_on_profile():
    """Brute force, the oracle's BVH and the reference's Embree give the same range profile for the same samples."""
    s = configs.small_radar()
    blob = s.export()
    n = 50000
    base = OracleScene(blob, accel=OracleScene.BRUTE).receive(n, seed=5, threads=4)
    assert base[..., 0].sum() > 0
    assert np.array_equal(OracleScene(blob, accel=OracleScene.BVH).receive(n, seed=5, threads=4), base)
    if os.path.exists(EMBREE):
        e = OracleScene(blob, accel=OracleScene.EMBREE_ACCEL).receive(n, seed=5, threads=4)
        err = np.linalg.norm(e[..., 0].astype(np.float64) - base[..., 0]) / np.linalg.norm(base[..., 0].astype(np.float64))
        assert err <= 1e-3
        assert e[..., 2].sum() == base[..., 2].sum()


def test_oracle_sample_sharding_is_additive():
    """Samples are independent given (seed, index): any partition of the index range sums to the whole (§8e)."""
    orc = OracleScene(configs.small_radar().export(), accel=OracleScene.BVH)
    whole = orc.receive(40000, seed=9, first=0, threads=1)
    parts = orc.receive(15000, seed=9, first=0, threads=1).astype(np.float64) + orc.receive(25000, seed=9, first=15000, threads=3)
    assert np.allclose(parts[..., 1:3], whole[..., 1:3])                 # counts exact
    assert np.linalg.norm(parts[..., 0] - whole[..., 0]) <= 1e-5 * np.linalg.norm(whole[..., 0])


def test_c1_trans_rad_energy_in_expected_bins():
    """trans_rad.xml (gen 1): target at 4 m => direct return at 2 * 4 / 3e8 = 26.7 ns => outside the 50 x 0.5 ns
    window; the ground at 0.5 m below gives returns from 2 * 0.5 / 3e8 = 3.3 ns on (bin 6 upward)."""
    orc = OracleScene(configs.c1_trans_rad().export(), accel=OracleScene.BRUTE)
    prof = orc.receive(100000, seed=0, threads=4)
    assert prof.shape == (1, 1, 155) and prof[0, 0, 4] == 100000
    bins = prof[0, 0, 5:].reshape(50, 3)
    assert np.array_equal(bins[:, 0], bins[:, 1]) and np.array_equal(bins[:, 0], bins[:, 2])
    assert bins[:6].sum() == 0 and bins[6:].sum() > 0


def test_c4_animated_frame_range_bins():
    """animated_trans_rad.py (gen 2, range(pathlength), dr 0.2 x 50 bins): the ground 0.5 m below the sensor is the nearest
    surface; with the reference's path-length bookkeeping (si.t is added at pathlength.cpp:146, :161 and :209 for a lit
    surface point) nothing can land below bin 13, and every sample is counted in the 1 x 1 film."""
    s, _ = configs.c4_animated(frame=37)
    prof = OracleScene(s.export(), accel=OracleScene.BRUTE).receive(50000, seed=0, threads=4)
    assert prof.shape == (1, 1, 55) and prof[0, 0, 4] == 50000
    bins = prof[0, 0, 5:]
    assert bins[:13].sum() == 0 and bins[13:].sum() > 0
    # the pose helper turns the boresight about z
    p0, p64 = configs.c4_pose(0).matrix32(), configs.c4_pose(64).matrix32()
    assert np.allclose(p0[:3, 2], [0, -1, 0], atol=1e-6) and np.allclose(p64[:3, 2], [1, 0, 0], atol=1e-6)


def test_develop_restatement_known_answers():
    """HDRADC::bitmap(raw=False): channel 2 is the weight (hdradc.cpp:256-259), value * (1.f / weight)
    (struct.cpp:1766, 1811), float16 rounds to nearest even; an empty bin is 0 * inf = NaN."""
    from oracle.develop import develop
    raw = np.array([[[6.0, 3.0, 3.0, 1.5], [1.0, 1.0, 4096.0, 2049.0], [0.0, 0.0, 0.0, 0.0], [1e6, 1.0, 1.0, -2.0]]], np.float32)
    f32 = develop(raw, "float32")
    assert f32.shape == (1, 4, 3) and f32.dtype == np.float32
    assert f32[0, 0].tolist() == [2.0, 1.0, 0.5]
    inv = np.float32(1.0) / np.float32(4096.0)
    assert f32[0, 1, 2] == np.float32(2049.0) * inv
    assert np.isnan(f32[0, 2]).all()
    f16 = develop(raw, "float16")
    assert f16.dtype == np.float16 and f16[0, 0].tolist() == [2.0, 1.0, 0.5]
    assert f16[0, 1, 2] == np.float16(0.5)            # 2049/4096 = 0.500244..: ties to even on the 10-bit mantissa
    assert np.isinf(f16[0, 3, 0]) and f16[0, 3, 2] == np.float16(-2.0)   # 1e6 overflows float16
    with pytest.raises(ValueError):
        develop(raw, "uint32")


def _splat(size_x, size_y, kind, radius, px, py, off=(0, 0)):
    out = np.zeros((size_y, size_x), np.float32)
    lib().orc_block_splat(size_x, size_y, off[0], off[1], kind, radius, px, py, out.ctypes.data_as(C.POINTER(C.c_float)))
    return out


def test_wide_filter_splat_known_answers():
    """SignalBlock::put, wide branch (signalblock.cpp:113-161) over the discretised filter (rfilter.cpp:9-20,
    rfilter.h:62-65): analytic footprints of the tent (radius 1) and gaussian (stddev 0.5, radius 2) filters."""
    # tent, sample on a bin centre: x = -1, 0 -> table[31] = 0 and table[0] = 1: everything in that bin
    o = _splat(8, 8, 2, 1.0, 3.5, 4.5)
    assert o[4, 3] == 1.0 and o.sum() == 1.0
    # tent, a quarter bin to the right: x = -0.25, 0.75 -> table[int(.25 * 31)] = 1 - 7/31, table[23] = 1 - 23/31
    o = _splat(8, 1, 2, 1.0, 3.75, 0.5)
    a, b = np.float32(1) - np.float32(7) / np.float32(31), np.float32(1) - np.float32(23) / np.float32(31)
    assert np.isclose(o[0, 3], a, rtol=1e-6) and np.isclose(o[0, 4], b, rtol=1e-6) and np.count_nonzero(o) == 2
    # gaussian: footprint 4 x 4, separable and symmetric about a bin corner
    o = _splat(16, 16, 1, 2.0, 8.0, 8.0)          # position 8.0 = the corner between bins 7 and 8
    assert np.count_nonzero(o) == 16 and np.allclose(o, o[::-1, ::-1]) and np.allclose(o, o.T)
    g = lambda x: max(0.0, math.exp(-2 * x * x) - math.exp(-8.0))
    tab = lambda x: g(2.0 * int(abs(x) * 31 / 2.0) / 31)          # the discretised filter is a step function
    assert np.isclose(o[8, 8], tab(0.5) ** 2, rtol=1e-5) and np.isclose(o[8, 9], tab(0.5) * tab(1.5), rtol=1e-5)
    # the crop at the storage edge: a sample half a bin outside still reaches the first bins; one beyond the radius does not
    o = _splat(16, 1, 1, 2.0, -0.5, 0.5)
    assert o[0, 0] > 0 and o[0, 1] >= 0 and np.count_nonzero(o) <= 2
    assert _splat(16, 1, 1, 2.0, -2.6, 0.5).sum() == 0 and _splat(16, 1, 1, 2.0, 18.6, 0.5).sum() == 0
    # window offset: positions are relative to it
    assert np.array_equal(_splat(8, 8, 2, 1.0, 13.5, 24.5, off=(10, 20)), _splat(8, 8, 2, 1.0, 3.5, 4.5))
    # radius 0.5 keeps the single-bin branch
    assert _splat(8, 8, 0, 0.5, 3.2, 4.9)[4, 3] == 1.0


def test_cubic_and_lanczos_filter_tables():
    """mitchell (B = C = 1/3), catmullrom and lanczos through the discretised table: value at the sample position, the
    negative lobe, and the closed forms at x = 0 (mitchell 8/9, catmullrom 1, lanczos 1)."""
    for kind, radius, centre in ((3, 2.0, (8.0 / 9.0) ** 2), (4, 2.0, 1.0), (5, 3.0, 1.0)):
        o = _splat(16, 16, kind, radius, 8.5, 8.5)            # on the centre of bin (8, 8): x = 0 for that bin
        assert np.isclose(o[8, 8], centre, rtol=1e-6)
        assert (o.min() < 0) == (kind == 5)                   # at integer offsets only lanczos-3 has a negative lobe
    o = _splat(32, 1, 4, 2.0, 8.0, 0.5)                       # catmullrom at x = 1.5: (1/6)(-3 x^3 + 15 x^2 - 24 x + 12) < 0
    x = 2.0 * int(1.5 * 31 / 2.0) / 31
    want = (1 / 6) * (-3 * x ** 3 + 15 * x ** 2 - 24 * x + 12)
    assert want < 0 and np.isclose(o[0, 9], want, rtol=1e-4) and np.isclose(o[0, 6], want, rtol=1e-4)


def test_rfilter_reference_spot_checks():
    """The reference's own reconstruction-filter spot checks (src/rfilters/tests/test_rfilter.py:16-51: gaussian eval(0.2) = 0.9227,
    lanczos eval(1.4) = -0.14668, mitchell eval(0) = 0.8888, catmullrom eval(0) = 0.9765, tent eval(0.1) = 0.903, zero beyond the
    radius; same tolerances) read off the oracle's wide splat: a sample at distance x from a bin centre on a one-row film deposits
    f(x) f(0), and f(0)^2 when it sits on the centre."""
    def f_of(kind, radius, x):
        centre = _splat(32, 1, kind, radius, 16.5, 0.5)[0, 16]
        return _splat(32, 1, kind, radius, 16.5 + x, 0.5)[0, 16] / math.sqrt(centre), math.sqrt(centre)
    g, g0 = f_of(1, 2.0, 0.2)
    assert abs(g - 0.9227) <= 8e-3 and f_of(1, 2.0, 2.1)[0] == 0
    lz, _ = f_of(5, 3.0, 1.4)
    assert abs(lz - (-0.14668)) <= 1e-2 and f_of(5, 3.0, 3.1)[0] == 0
    _, m0 = f_of(3, 2.0, 0.0)
    assert abs(m0 - 0.8888) <= 1e-3 and f_of(3, 2.0, 2.1)[0] == 0
    _, c0 = f_of(4, 2.0, 0.0)
    assert abs(c0 - 0.9765) <= 5e-2 and f_of(4, 2.0, 2.1)[0] == 0
    t, _ = f_of(2, 1.0, 0.1)
    assert abs(t - 0.903) <= 5e-2 and f_of(2, 1.0, 1.1)[0] == 0


def test_develop_reference_property_unit_weights():
    """The reference's own film test (src/adcs/tests/test_hdrfilm.py:73-140, test03_develop): a 41 x 37 image put with unit
    weights and developed with float32 components comes back unchanged (atol 1e-5), the weight channel dropped.  Same seed and
    size; here the storage layout is the radar ADC's (Y, A, W, aovs) with the weight at index 2 (hdradc.cpp:256-259)."""
    from oracle.develop import develop
    np.random.seed(12345 + ord("e"))
    contents = np.random.uniform(size=(37, 41, 5)).astype(np.float32)
    contents[:, :, 2] = 1.0
    img = develop(contents, "float32")
    assert img.shape == (37, 41, 4)
    assert np.allclose(img, contents[:, :, [0, 1, 3, 4]], atol=1e-5)
    # and a constant non-unit weight rescales every channel by it
    contents[:, :, 2] = 4.0
    assert np.allclose(develop(contents, "float32") * 4.0, contents[:, :, [0, 1, 3, 4]], atol=1e-5)


def test_spectra_reference_values():
    """The reference's own spectrum tests (src/spectra/tests/test_regular.py:22-31: `regular` 500..600 nm with values 1, 2 evaluates
    to 0, 1, 1.5, 2, 0 at 450, 500, 550, 600, 650 nm; test_irregular.py:22-31: `irregular` (500, 600, 650) -> (1, 2, 0.5) evaluates
    to 0, 1, 1.5, 2, 0.5, 0 at 450 .. 700 nm), read off the oracle's spectrum evaluation of a scene that holds them."""
    s = configs.small_radar(spectrum_width=4)
    reg = s.add_spectrum(N.BF_SPEC_REGULAR, [1.0, 2.0], [500.0, 600.0])
    irr = s.add_spectrum(N.BF_SPEC_IRREGULAR, [1.0, 2.0, 0.5], [500.0, 600.0, 650.0])
    orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
    got = [lib().orc_spectrum_eval(orc._h, reg, 450.0 + 50.0 * i) for i in range(5)]
    assert np.allclose(got, [0, 1, 1.5, 2, 0])
    got = [lib().orc_spectrum_eval(orc._h, irr, 450.0 + 50.0 * i) for i in range(6)]
    assert np.allclose(got, [0, 1, 1.5, 2, 0.5, 0])
    s.close(); orc.close()


def test_wide_splat_against_independent_restatement():
    """SignalBlock::put's wide branch restated in numpy from the reference text (tests/record_binning.py: footprint lo / hi, the n
    weights per axis from the discretised filter, the border and its crop) against the oracle's splat, for all five rfilter plugins at
    random sample positions including ones outside the window."""
    from record_binning import wide_splat
    rng = np.random.default_rng(3)
    for kind, radius in ((1, 2.0), (2, 1.0), (3, 2.0), (4, 2.0), (5, 3.0)):
        worst = 0.0
        for _ in range(150):
            px, py = rng.uniform(-3.5, 15.5, size=2)
            off = (int(rng.integers(0, 3)), int(rng.integers(0, 3)))
            a = _splat(12, 9, kind, radius, float(px) + off[0], float(py) + off[1], off=off).astype(np.float64)
            b = wide_splat(12, 9, kind, radius, float(np.float32(px + off[0])), float(np.float32(py + off[1])), off=off)
            worst = max(worst, float(np.abs(a - b).max()))
            assert np.array_equal(a != 0, np.abs(b) > 1e-9) or np.abs(a - b).max() <= 2e-6, (kind, px, py)
        assert worst <= 2e-6, (kind, worst)
