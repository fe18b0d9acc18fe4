"""Independent anchor for the image-side entry point (rows a3, a6, a17: `render_sample`, range(pathlength), perspective sensor, area
emitter): tests/plate_path_render.py evaluates single samples in numpy straight from the reference's text; the CPU oracle (not gpu)
and the CUDA path (gpu, through the C ABI) must reproduce them -- film position, first-hit distance, valid flag, radiance, the
integrator's path-length scalar with the reference's double counting of the first segment, and the range bin it selects."""
import math

import numpy as np
import pytest

from beifong_b200 import _native as N
from beifong_b200 import meshes
from beifong_b200.meshes import Transform as T
from beifong_b200.scene import FlatScene
from oracle_lib import OracleScene
from plate_path_render import RenderScene, evaluate_render_sample, range_bin


def build(sc):
    s = FlatScene(spectrum_width=1)
    dflt = s.add_diffuse(0.5, two_sided=False)
    plate_m = s.add_diffuse(sc.rho, two_sided=False)
    em = s.add_rect((T.translate([sc.em_x, 0.0, 0.0]) * T.scale([sc.em_half, sc.em_half, 1.0])).matrix32(), dflt)
    s.add_rect((T.translate([0.0, 0.0, sc.R]) * T.rotate([1, 0, 0], 180.0) * T.scale([sc.plate_half, sc.plate_half, 1.0])).matrix32(), plate_m)
    s.add_rect((T.translate([0.0, 0.0, sc.floor_z]) * T.scale([sc.floor_half, sc.floor_half, 1.0])).matrix32(), plate_m)
    s.add_transmitter(em, FlatScene.endpoint(N.BF_TX_AREALIGHT, texture=s.add_uniform(sc.radiance)))
    s.set_receiver(None, FlatScene.endpoint(N.BF_RX_PERSPECTIVE, film_width=1, film_height=1, near_clip=sc.near, far_clip=sc.far,
                                            to_world=np.eye(4, dtype=np.float32),
                                            sample_to_camera=meshes.perspective_sample_to_camera(sc.fov, sc.near, sc.far, 1, 1)))
    s.set_integrator(N.BF_GEN_RANGE, max_depth=2, rr_depth=5, range_bins=sc.bins, dr=sc.dr)
    return s


def check_records(sc, rec, seed, n):
    checked = valid = lit = on_emitter = 0
    bins = np.zeros(sc.bins)
    for i in range(n):
        e = evaluate_render_sample(sc, seed, i)
        if e is None:
            continue
        r = rec[i]
        checked += 1
        assert r["tf_t"] == pytest.approx(e["pos"][0], abs=1e-7) and r["tf_f"] == pytest.approx(e["pos"][1], abs=1e-7)
        assert int(r["valid"]) == e["valid"], (i, r, e)
        if not e["valid"]:
            assert r["value"] == 0.0
            continue
        valid += 1
        assert r["first_t"] == pytest.approx(e["first_t"], rel=3e-6)
        assert int(r["depth"]) == e["depth"]
        assert r["phase"] == pytest.approx(e["scalar"], rel=1e-5), (i, r["phase"], e["scalar"])      # the integrator's scalar; t2 of a grazing continuation ray carries the fp32 error of its direction
        assert r["value"] == pytest.approx(e["value"], rel=2e-4, abs=1e-9), (i, r["value"], e["value"])
        lit += e["value"] > 0.0
        on_emitter += bool(e.get("on_emitter"))
        b = range_bin(sc, e["scalar"])
        if b >= 0 and abs(e["scalar"] / sc.dr - round(e["scalar"] / sc.dr)) > 1e-4:      # away from a bin edge
            bins[b] += e["value"]
    return checked, valid, lit, bins, on_emitter


def test_oracle_reproduces_the_independent_render_sample():
    sc = RenderScene()
    s = build(sc)
    orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
    seed, n = 5, 3000
    rec = orc.sample_records(n, seed=seed)
    checked, valid, lit, bins, on_emitter = check_records(sc, rec, seed, n)
    assert checked == n and valid == checked and lit > 0.9 * valid          # every camera ray meets the plate
    assert on_emitter >= 3                                                  # the emission + MIS branch (BSDF ray lands on the emitter) is among them
    # the range axis: 2 t1 + t2 with t1 >= R puts everything at or beyond bin floor(2 R / dr); the binned film agrees with the records
    prof = orc.receive(n, seed=seed, threads=1)
    assert prof.shape[:2] == (1, 1) and prof[0, 0, 4] == n
    aov = prof[0, 0, 5:5 + sc.bins].astype(np.float64)
    assert aov[: int(2 * sc.R / sc.dr)].sum() == 0.0 and aov.sum() > 0.0
    assert np.abs(aov - bins).sum() <= 0.005 * aov.sum()         # samples within 1e-4 of a bin edge are all that is left out
    s.close(); orc.close()


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", ["wavefront", "mega"])
def test_device_reproduces_the_independent_render_sample(native, pipeline):
    sc = RenderScene()
    s = build(sc).commit()
    s.set_tunable("mega_max", (1 << 28) if pipeline == "mega" else -1)
    seed, n = 5, 3000
    rec = s.sample_records(n, seed=seed)
    checked, valid, lit, _, on_emitter = check_records(sc, rec, seed, n)
    assert checked == n and valid == checked and lit > 0.9 * valid and on_emitter >= 3
    s.close()


# ---- gen 1: fluxmeter + spot emitter + time(pathtime) ------------------------------------------------------------------------------
from plate_path_render import FluxScene, evaluate_flux_sample      # noqa: E402


def build_flux(sc):
    s = FlatScene(spectrum_width=1)
    dflt = s.add_diffuse(0.5, two_sided=False)
    plate_m = s.add_diffuse(sc.rho, two_sided=False)
    rx = s.add_rect(T.scale([sc.a, sc.a, 1.0]).matrix32(), dflt)
    s.add_rect((T.translate([0.0, 0.0, sc.R]) * T.rotate([1, 0, 0], 180.0) * T.scale([sc.plate_half, sc.plate_half, 1.0])).matrix32(), plate_m)
    s.add_rect((T.translate([0.0, 0.0, sc.floor_z]) * T.scale([sc.floor_half, sc.floor_half, 1.0])).matrix32(), plate_m)
    s.add_transmitter(None, FlatScene.endpoint(N.BF_TX_SPOT, texture=s.add_uniform(sc.intensity), cutoff_angle_deg=sc.cutoff_deg,
                                               beam_width_deg=sc.beam_deg, to_world=T.translate(list(sc.spot)).matrix32()))
    s.set_receiver(rx, FlatScene.endpoint(N.BF_RX_FLUXMETER))
    s.set_integrator(N.BF_GEN_TIME, max_depth=2, rr_depth=5)
    return s


def check_flux_records(sc, rec, seed, n):
    """Records carry the integrator's unweighted radiance (what time.cpp bins); the film's X channel carries ray_weight * L."""
    checked = valid = lit = 0
    film_x = binned = 0.0
    for i in range(n):
        e = evaluate_flux_sample(sc, seed, i)
        if e is None:
            continue
        r = rec[i]
        checked += 1
        assert r["tf_t"] == pytest.approx(e["pos"][0], abs=1e-7) and r["tf_f"] == pytest.approx(e["pos"][1], abs=1e-7)
        assert int(r["valid"]) == e["valid"], (i, r, e)
        if not e["valid"]:
            assert r["value"] == 0.0
            continue
        valid += 1
        assert r["first_t"] == pytest.approx(e["first_t"], rel=3e-6)
        assert int(r["depth"]) == e["depth"]
        assert r["phase"] == pytest.approx(e["scalar"], rel=1e-5), (i, r["phase"], e["scalar"])       # travel time in seconds
        # inside the transition zone the falloff is (cutoff - acos(cos_theta)) / width: acosf near its steep end
        assert r["value"] == pytest.approx(e["radiance"], rel=5e-4, abs=1e-9), (i, r["value"], e["radiance"])
        lit += e["radiance"] > 0.0
        film_x += e["value"]
        if np.float32(e["scalar"]) < np.float32(49) * np.float32(0.5e-9) + np.float32(0.5e-9):      # time.cpp:118-124: 50 bins of 0.5 ns
            binned += e["radiance"]
    return checked, valid, lit, film_x, binned


def test_oracle_reproduces_the_independent_flux_sample():
    sc = FluxScene()
    s = build_flux(sc)
    orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
    seed, n = 8, 3000
    rec = orc.sample_records(n, seed=seed)
    checked, valid, lit, film_x, binned = check_flux_records(sc, rec, seed, n)
    assert checked >= 0.99 * n and valid > 0.5 * n and lit > 0.5 * valid
    # all three falloff branches occur: outside the cone (0), the transition zone, the full beam
    vals = [evaluate_flux_sample(sc, seed, i) for i in range(n)]
    lit_vals = [v for v in vals if v and v["valid"]]
    assert any(v["radiance"] == 0.0 for v in lit_vals) and any(v["radiance"] > 0.0 for v in lit_vals)
    # the film: X = sum of pi * L (FluxMeter's ray weight, fluxmeter.cpp:83), W = n, the 50 time bins hold the unweighted L
    prof = orc.receive(n, seed=seed, threads=1)
    assert prof.shape[:2] == (1, 1) and prof[0, 0, 4] == n
    assert prof[0, 0, 0] == pytest.approx(film_x, rel=0.02)          # the unmodelled < 1 % of the samples is all that differs
    assert 0.5 * film_x / math.pi < binned < film_x / math.pi          # paths longer than 25 ns fall off the time axis
    assert prof[0, 0, 5::3].sum() == pytest.approx(binned, rel=0.02)
    s.close(); orc.close()


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", ["wavefront", "mega"])
def test_device_reproduces_the_independent_flux_sample(native, pipeline):
    sc = FluxScene()
    s = build_flux(sc).commit()
    s.set_tunable("mega_max", (1 << 28) if pipeline == "mega" else -1)
    seed, n = 8, 3000
    rec = s.sample_records(n, seed=seed)
    checked, valid, lit, _, _ = check_flux_records(sc, rec, seed, n)
    assert checked >= 0.99 * n and valid > 0.5 * n and lit > 0.5 * valid
    s.close()
