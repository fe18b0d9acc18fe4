"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol include/beifong_b200/abi.h
declares, and the host side mirrors the reference's error behaviour (Throw -> RuntimeError,
NotImplementedError for plugins outside the path).  No compute call is made (no GPU here)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from beifong_b200 import _native as N
from beifong_b200 import configs, meshes
from beifong_b200.scene import FlatScene, BeifongError

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "beifong_b200", "abi.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bf_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(native):
    syms = declared_symbols()
    assert len(syms) >= 30
    for name in syms:
        assert hasattr(native, name), f"{name} declared in abi.h but not exported"
    assert set(syms) == set(N.PROTOTYPES), set(syms) ^ set(N.PROTOTYPES)
    out = subprocess.run(["nm", "-D", "--defined-only", N.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (bf_[a-z0-9_]+)", out))
    assert set(syms) <= exported
    assert native.bf_abi_version() == 1


def test_struct_layouts_match_header(native):
    # sizes the C side and the ctypes mirror must agree on (checked through the flat export round trip below too)
    assert C.sizeof(N.bf_constants) == 32 and C.sizeof(N.bf_bsdf_desc) == 48 and C.sizeof(N.bf_signal_desc) == 40
    assert C.sizeof(N.bf_adc_desc) == 40 and C.sizeof(N.bf_integrator_desc) == 32 and C.sizeof(N.bf_sample_record) == 48


def test_product_is_sm100a_only_and_has_no_oracle_dependency():
    out = subprocess.run(["cuobjdump", "--list-elf", N.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs
    ldd = subprocess.run(["ldd", N.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in ldd and "embree" not in ldd
    for dirpath, _, files in os.walk(os.path.join(ROOT, "beifong_b200")):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, fn), errors="replace").read()
                assert "liboracle" not in src and "oracle_lib" not in src, fn


def test_commit_fails_loudly_without_gpu(native):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    s = configs.small_radar()
    with pytest.raises(BeifongError, match="no CUDA device|CUDA"):
        s.commit(0)
    with pytest.raises(BeifongError, match="not committed"):
        s.receive(16)
    with pytest.raises(BeifongError, match="not committed"):
        s.set_stream(None)                                   # streams exist from bf_scene_commit on


def test_error_behaviour_mirrors_reference():
    s = FlatScene(spectrum_width=1)
    with pytest.raises(NotImplementedError):
        FlatScene(spectrum_width=3)                          # only *_mono / *_spectral exist (mitsuba.conf)
    with pytest.raises(BeifongError):
        s.add_bsdf(N.BF_BSDF_DIFFUSE, reflectance=7)         # dangling reference (xml.cpp: unresolved <ref>)
    with pytest.raises(NotImplementedError):
        s.add_bsdf(17)                                       # plugin outside the radar path
    b = s.add_diffuse(0.5)
    with pytest.raises(BeifongError):
        s.add_mesh(np.zeros((3, 3), np.float32), np.array([[0, 1, 3]], np.uint32), b)   # face index out of range
    with pytest.raises(BeifongError):
        s.set_integrator(N.BF_GEN_PHASE, max_depth=0)        # integrator.cpp:1713-1720
    with pytest.raises(NotImplementedError):
        s.set_adc(filter_radius=2.0)                         # wide reconstruction filters: not on the accelerated path
    with pytest.raises(BeifongError):
        s.commit(0)                                          # no receiver / integrator yet


def test_flat_export_round_trip_and_profile_size():
    s = configs.small_radar(spectrum_width=4)
    blob = s.export()
    magic, version = np.frombuffer(blob[:8], np.uint32)
    assert magic == 0x31464642 and version == 1
    assert int(np.frombuffer(blob[8:16], np.uint64)[0]) == blob.size
    assert s.profile_shape() == (1024, 1, 4)                 # [f][t][Y, A, W, phase bin 0]
    assert configs.c1_trans_rad().profile_shape() == (1, 1, 155)
    from oracle_lib import OracleScene                       # the oracle parses the same blob: shapes agree
    assert OracleScene(blob).profile_shape() == (1024, 1, 4)


def test_ply_obj_loaders_on_generated_files(tmp_path):
    p, f = meshes.stairs(3)
    # ascii PLY with an extra vertex property and a quad face; binary little-endian PLY; OBJ with v/vt indices
    ascii_ply = tmp_path / "a.ply"
    with open(ascii_ply, "w") as fh:
        fh.write("ply\nformat ascii 1.0\nelement vertex 4\nproperty float x\nproperty float y\nproperty float z\n"
                 "property float confidence\nelement face 1\nproperty list uchar int vertex_indices\nend_header\n")
        fh.write("0 0 0 1\n1 0 0 1\n1 1 0 1\n0 1 0 1\n4 0 1 2 3\n")
    pp, ff, nn = meshes.load_ply(str(ascii_ply))
    assert pp.shape == (4, 3) and ff.tolist() == [[0, 1, 2], [0, 2, 3]] and nn is None
    bin_ply = tmp_path / "b.ply"
    with open(bin_ply, "wb") as fh:
        fh.write(b"ply\nformat binary_little_endian 1.0\nelement vertex %d\nproperty float x\nproperty float y\nproperty float z\n"
                 b"element face %d\nproperty list uchar int vertex_indices\nend_header\n" % (len(p), len(f)))
        fh.write(p.astype("<f4").tobytes())
        for tri in f:
            fh.write(b"\x03" + tri.astype("<i4").tobytes())
    pp, ff, _ = meshes.load_ply(str(bin_ply))
    assert np.array_equal(pp, p) and np.array_equal(ff, f)
    obj = tmp_path / "c.obj"
    obj.write_text("v -1 -1 0\nv 1 -1 0\nv -1 1 0\nv 1 1 0\nvt 1 1\nvt 0 0\nf 2/1 3/2 1/2\nf 2/1 4/2 3/2\n")
    pp, ff, _ = meshes.load_obj(str(obj))
    assert ff.tolist() == [[1, 2, 0], [1, 3, 2]]


@pytest.mark.parametrize("version,double_precision", [(4, False), (3, False), (4, True)])
def test_serialized_loader_round_trip(tmp_path, version, double_precision):
    """Mitsuba .serialized meshes (serialized.cpp:196-300): v3 / v4 headers, single / double precision, several
    sub-shapes addressed by shape_index through the offset dictionary at the end of the file."""
    p0, f0 = meshes.car_body(12, 7)
    n0 = meshes.vertex_normals(p0, f0)
    p1, f1 = meshes.stairs(5)
    path = str(tmp_path / "two.serialized")
    meshes.save_serialized(path, [(p0, f0, n0), (p1, f1, None)], version=version, double_precision=double_precision)
    a = meshes.load_serialized(path)
    assert np.array_equal(a[0], p0.astype(np.float32)) and np.array_equal(a[1], f0) and np.allclose(a[2], n0, atol=1e-7)
    b = meshes.load_serialized(path, shape_index=1)
    assert np.array_equal(b[0], p1.astype(np.float32)) and np.array_equal(b[1], f1) and b[2] is None
    with pytest.raises(RuntimeError, match="out of range"):
        meshes.load_serialized(path, shape_index=2)
    m = (meshes.Transform.translate([1.0, 2.0, 3.0]) * meshes.Transform.scale(2.0)).matrix32()
    c = meshes.load_serialized(path, to_world=m)
    assert np.allclose(c[0], p0 * 2.0 + np.array([1.0, 2.0, 3.0], np.float32), atol=1e-5)
    bad = tmp_path / "bad.serialized"
    bad.write_bytes(b"\x00\x01\x04\x00garbage")
    with pytest.raises(RuntimeError, match="invalid file format"):
        meshes.load_serialized(str(bad))


REF_SERIALIZED = "/root/reference/resources/data/tests/serialized"


@pytest.mark.skipif(not os.path.isdir(REF_SERIALIZED), reason="reference tree not present")
@pytest.mark.parametrize("features", ["normals", "uv", "normals_uv"])
def test_serialized_reference_fixtures(features):
    """The reference's own fixtures and expectations (src/librender/tests/test_mesh.py:130-172, test06)."""
    pos, faces, nrm = meshes.load_serialized(os.path.join(REF_SERIALIZED, f"rectangle_{features}.serialized"))
    assert np.allclose(pos[0], [-2.85, 0.0, -7.6], atol=1e-3)
    assert np.allclose(pos[2], [2.85, 0.0, 0.599999], atol=1e-3)
    assert np.allclose(pos[3], [2.85, 0.0, -7.6], atol=1e-3)
    assert faces.shape[1] == 3 and faces.max() < len(pos)
    if "normals" in features:
        assert np.allclose(nrm[[0, 2, 3]], [0.0, 1.0, 0.0])
    else:
        assert nrm is None
        assert np.allclose(meshes.vertex_normals(pos, faces)[[0, 2, 3]], [0.0, 1.0, 0.0], atol=1e-6) or \
            np.allclose(meshes.vertex_normals(pos, faces)[[0, 2, 3]], [0.0, -1.0, 0.0], atol=1e-6)


def test_tunables_by_name():
    """bf_set_tunable: every documented knob is accepted before the commit (none changes a result), ranges and names are checked."""
    s = configs.small_radar()
    for name, value in (("wave_lanes", 1 << 22), ("wave_streams", 4), ("fused_primary", 1), ("primary_defer", -1), ("trace_defer", 0),
                        ("trace_rounds", 16), ("use_graph", 1), ("mega_batch", 0), ("path_fused", 1), ("mega_max", -1),
                        ("l2_persist_mb", 0), ("path_stash_max", 32768)):
        s.set_tunable(name, value)
    with pytest.raises(BeifongError, match="out of range"):
        s.set_tunable("wave_streams", 99)
    with pytest.raises(BeifongError, match="unknown tunable"):
        s.set_tunable("no_such_knob", 1)
    s.close()
