"""ctypes wrapper of oracle/liboracle.so — TEST INFRASTRUCTURE ONLY (never imported by beifong_b200/)."""
import ctypes as C
import os
import subprocess

import numpy as np

from beifong_b200 import _native as N
from beifong_b200.scene import RECORD_DTYPE

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "liboracle.so")
EMBREE = os.path.join(ORACLE_DIR, "_ref", "libembree3.so")

_lib = None
u32, u64, f32 = C.c_uint32, C.c_uint64, C.c_float
P = C.POINTER


def build():
    if not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(os.path.join(ORACLE_DIR, "oracle.cpp")):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "liboracle.so"], stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB)
        L.orc_scene_import.restype = C.c_void_p
        L.orc_scene_import.argtypes = [C.c_void_p, C.c_size_t]
        L.orc_scene_destroy.argtypes = [C.c_void_p]
        L.orc_last_error.restype = C.c_char_p
        L.orc_last_error.argtypes = [C.c_void_p]
        L.orc_scene_set_accel.argtypes = [C.c_void_p, C.c_int, C.c_char_p]
        L.orc_profile_size.argtypes = [C.c_void_p, P(u32), P(u32), P(u32)]
        L.orc_block_splat.argtypes = [u32, u32, C.c_int, C.c_int, u32, f32, f32, f32, P(f32)]
        L.orc_receive.argtypes = [C.c_void_p, u64, u64, u64, P(f32), C.c_int, C.c_int, C.c_int, P(u64)]
        L.orc_sample_records.argtypes = [C.c_void_p, u64, u64, u64, C.c_void_p, C.c_int]
        L.orc_trace_first_hits.argtypes = [C.c_void_p, P(f32), P(f32), P(f32), P(f32), u64, P(u32), P(u32), P(f32), P(f32), P(f32), C.c_int]
        L.orc_trace_occluded.argtypes = [C.c_void_p, P(f32), P(f32), P(f32), P(f32), u64, P(C.c_uint8), C.c_int]
        L.orc_surface_interaction.argtypes = [C.c_void_p, P(f32), P(f32), P(f32), P(f32), P(u32)]
        for name in ("orc_tea_float32",):
            getattr(L, name).restype = f32
        L.orc_tea_float64.restype = C.c_double
        L.orc_tea64.restype = u64
        L.orc_tea64_wide.restype = u64
        L.orc_tea64_wide.argtypes = [u64, u64, C.c_int]
        L.orc_pcg32_next.restype = u32
        L.orc_pcg32_next.argtypes = [P(u64), P(u64)]
        L.orc_pcg32_seed.argtypes = [u64, u64, P(u64), P(u64)]
        L.orc_pcg32_next_float.restype = f32
        L.orc_pcg32_next_float.argtypes = [P(u64), P(u64)]
        for name, n in (("orc_sinc", 1), ("orc_tri", 1), ("orc_rect", 1), ("orc_fmodulo", 2), ("orc_wchirp", 4), ("orc_erfinv", 1),
                        ("orc_fresnel_conductor", 3)):
            fn = getattr(L, name)
            fn.restype = f32
            fn.argtypes = [f32] * n
        L.orc_square_to_uniform_disk_concentric.argtypes = [f32, f32, P(f32)]
        L.orc_square_to_cosine_hemisphere.argtypes = [f32, f32, P(f32)]
        L.orc_coordinate_system.argtypes = [P(f32), P(f32), P(f32)]
        L.orc_microfacet_eval.restype = f32
        L.orc_microfacet_eval.argtypes = [C.c_int, f32, f32, P(f32)]
        L.orc_microfacet_smith_g1.restype = f32
        L.orc_microfacet_smith_g1.argtypes = [C.c_int, f32, f32, P(f32), P(f32)]
        L.orc_microfacet_sample.argtypes = [C.c_int, f32, f32, C.c_int, P(f32), f32, f32, P(f32), P(f32)]
        L.orc_bsdf_eval.argtypes = [C.c_void_p, u32, P(f32), P(f32), P(f32), P(f32)]
        L.orc_bsdf_pdf.restype = f32
        L.orc_bsdf_pdf.argtypes = [C.c_void_p, u32, P(f32), P(f32), P(f32)]
        L.orc_bsdf_sample.argtypes = [C.c_void_p, u32, P(f32), f32, f32, f32, P(f32), P(f32), P(f32), P(f32)]
        L.orc_phased_sample_wigner.restype = f32
        L.orc_phased_sample_wigner.argtypes = [C.c_void_p, C.c_int, P(f32), P(f32), f32]
        L.orc_rect_sample_wigner.restype = f32
        L.orc_rect_sample_wigner.argtypes = [C.c_void_p, u32, P(f32), P(f32), f32]
        L.orc_rect_area.restype = f32
        L.orc_rect_area.argtypes = [C.c_void_p, u32]
        L.orc_eval_signal.restype = f32
        L.orc_eval_signal.argtypes = [P(N.bf_signal_desc), f32, f32]
        L.orc_sample_delta_frequency.restype = f32
        L.orc_sample_delta_frequency.argtypes = [P(N.bf_signal_desc), f32]
        L.orc_spectrum_eval.restype = f32
        L.orc_spectrum_eval.argtypes = [C.c_void_p, u32, f32]
        L.orc_block_bin.argtypes = [u32, u32, C.c_int, C.c_int, f32, f32]
        _lib = L
    return _lib


def _fp(a):
    return a.ctypes.data_as(P(f32))


def _up(a):
    return a.ctypes.data_as(P(u32))


def fvec(*v):
    return (f32 * len(v))(*v)


class OracleScene:
    """CPU oracle instance built from the blob bf_scene_export writes."""
    BRUTE, BVH, EMBREE_ACCEL = 0, 1, 2

    def __init__(self, blob, accel=1):
        self._blob = np.ascontiguousarray(blob, dtype=np.uint8)
        self._h = lib().orc_scene_import(self._blob.ctypes.data_as(C.c_void_p), self._blob.size)
        if not self._h:
            raise RuntimeError("oracle: bad scene blob")
        self.set_accel(accel)

    def set_accel(self, accel):
        path = EMBREE.encode() if accel == 2 else None
        if accel == 2 and not os.path.exists(EMBREE):
            raise FileNotFoundError(EMBREE)
        if lib().orc_scene_set_accel(self._h, accel, path):
            raise RuntimeError("oracle: " + lib().orc_last_error(self._h).decode())

    def close(self):
        if self._h:
            lib().orc_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def profile_shape(self):
        w, h, c = u32(), u32(), u32()
        lib().orc_profile_size(self._h, C.byref(w), C.byref(h), C.byref(c))
        return (h.value, w.value, c.value)

    def receive(self, n, seed=0, first=0, rng_mode=0, threads=1):
        out = np.zeros(self.profile_shape(), np.float32)
        q = u64()
        lib().orc_receive(self._h, seed, first, n, _fp(out), 0, rng_mode, threads, C.byref(q))
        self.last_queries = q.value
        return out

    def sample_records(self, n, seed=0, first=0, rng_mode=0):
        recs = np.zeros(n, dtype=RECORD_DTYPE)
        lib().orc_sample_records(self._h, seed, first, n, recs.ctypes.data_as(C.c_void_p), rng_mode)
        return recs

    def trace_first_hits(self, o, d, mint=None, maxt=None, accel=-1):
        o = np.ascontiguousarray(o, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(d, np.float32).reshape(-1, 3)
        n = len(o)
        mi = np.ascontiguousarray(mint, np.float32) if mint is not None else None
        ma = np.ascontiguousarray(maxt, np.float32) if maxt is not None else None
        shape, prim = np.empty(n, np.uint32), np.empty(n, np.uint32)
        t, u, v = np.empty(n, np.float32), np.empty(n, np.float32), np.empty(n, np.float32)
        lib().orc_trace_first_hits(self._h, _fp(o), _fp(d), _fp(mi) if mi is not None else None, _fp(ma) if ma is not None else None,
                                   n, _up(shape), _up(prim), _fp(t), _fp(u), _fp(v), accel)
        return shape, prim, t, u, v

    def trace_occluded(self, o, d, mint=None, maxt=None, accel=-1):
        o = np.ascontiguousarray(o, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(d, np.float32).reshape(-1, 3)
        n = len(o)
        mi = np.ascontiguousarray(mint, np.float32) if mint is not None else None
        ma = np.ascontiguousarray(maxt, np.float32) if maxt is not None else None
        hit = np.empty(n, np.uint8)
        lib().orc_trace_occluded(self._h, _fp(o), _fp(d), _fp(mi) if mi is not None else None, _fp(ma) if ma is not None else None,
                                 n, hit.ctypes.data_as(P(C.c_uint8)), accel)
        return hit

    def surface_interaction(self, o, d):
        out = (f32 * 17)()
        t, prim = f32(), u32()
        rc = lib().orc_surface_interaction(self._h, fvec(*o), fvec(*d), out, C.byref(t), C.byref(prim))
        return rc, t.value, prim.value, np.array(out[:], np.float32)
