"""bf_trace_first_hits_static (abi.h): the closest-hit query with every device-side temporary in the library's static device
array -- no cudaMalloc inside the call.  No counterpart in the reference.
Checked against the allocating entry point bf_trace_first_hits on the same rays, bit for bit (the tolerance the other traversal
tests use: ids identical, t / u / v identical as uint32), on two calls in a row that share the array.  Collected last on purpose
(file name)."""
import ctypes as C

import numpy as np
import pytest

from beifong_b200 import configs
from beifong_b200 import _native as N
from beifong_b200.scene import BeifongError

pytestmark = pytest.mark.gpu


def rays(n, seed, center=(10.0, 3.0, 1.0), radius=6.0):
    rng = np.random.default_rng(seed)
    center = np.asarray(center)
    o = rng.normal(size=(n, 3))
    o = center + radius * o / np.linalg.norm(o, axis=1, keepdims=True)
    tgt = center + rng.uniform(-0.4, 0.4, size=(n, 3)) * radius
    d = tgt - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return o.astype(np.float32), d.astype(np.float32)


def same(a, b):
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])            # shape, prim
    hit = b[0] != N.BF_NONE
    for k in (2, 3, 4):                                                          # t, u, v
        assert np.array_equal(a[k][hit].view(np.uint32), b[k][hit].view(np.uint32))
    assert np.all(np.isinf(a[2][~hit])) and np.all(np.isinf(b[2][~hit]))
    return hit


def test_static_trace_equals_allocating_trace_twice(native):
    s = configs.small_radar(spectrum_width=1).commit()
    # call 1: a full-size request with explicit ray intervals
    n1 = N.BF_STATIC_TRACE_MAX_RAYS
    o, d = rays(n1, 11)
    rng = np.random.default_rng(12)
    mint = rng.uniform(0.0, 0.5, n1).astype(np.float32)
    maxt = rng.uniform(4.0, 30.0, n1).astype(np.float32)
    hit = same(s.trace_first_hits_static(o, d, mint, maxt), s.trace_first_hits(o, d, mint, maxt))
    assert 0.1 * n1 < hit.sum() < n1
    # call 2, straight after: fewer rays and default intervals, so the array still holds call 1's rays, intervals and hits
    n2 = 3001
    o2, d2 = rays(n2, 13)
    hit2 = same(s.trace_first_hits_static(o2, d2), s.trace_first_hits(o2, d2))
    assert hit2.sum() > 300
    # and call 1 again gives call 1's answer again
    same(s.trace_first_hits_static(o, d, mint, maxt), s.trace_first_hits(o, d, mint, maxt))
    s.close()


def test_static_trace_second_scene_shares_the_array(native):
    a = configs.small_radar(spectrum_width=1).commit()
    b = configs.small_radar(spectrum_width=4, variant=dict(second_tx=True)).commit()
    o, d = rays(20000, 21)
    ra = a.trace_first_hits_static(o, d)
    rb = b.trace_first_hits_static(o, d)         # same GPU, same array, another scene's tables
    same(ra, a.trace_first_hits(o, d))
    same(rb, b.trace_first_hits(o, d))
    a.close()
    b.close()


def test_static_trace_limits(native):
    s = configs.small_radar(spectrum_width=1).commit()
    n = N.BF_STATIC_TRACE_MAX_RAYS + 1
    o, d = rays(n, 31)
    with pytest.raises(BeifongError, match="at most 65536 rays"):
        s.trace_first_hits_static(o, d)
    # null outputs are skipped, zero rays is a no-op
    f32p, u32p = C.POINTER(C.c_float), C.POINTER(C.c_uint32)
    o, d = rays(1000, 32)
    t = np.empty(1000, np.float32)
    rc = native.bf_trace_first_hits_static(s._h, o.ctypes.data_as(f32p), d.ctypes.data_as(f32p), None, None, 1000,
                                           C.cast(None, u32p), C.cast(None, u32p), t.ctypes.data_as(f32p), C.cast(None, f32p), C.cast(None, f32p))
    assert rc == 0
    assert np.array_equal(t.view(np.uint32), s.trace_first_hits(o, d)[2].view(np.uint32))
    assert native.bf_trace_first_hits_static(s._h, o.ctypes.data_as(f32p), d.ctypes.data_as(f32p), None, None, 0,
                                             C.cast(None, u32p), C.cast(None, u32p), C.cast(None, f32p), C.cast(None, f32p), C.cast(None, f32p)) == 0
    s.close()


def test_static_trace_device_pointers(native):
    """Inputs and outputs in device memory the CALLER owns (here: torch tensors), as a caller under graph capture would pass."""
    torch = pytest.importorskip("torch")
    s = configs.small_radar(spectrum_width=1).commit()
    n = 5000
    o, d = rays(n, 41)
    dev = torch.device("cuda", 0)
    to, td = torch.from_numpy(o).to(dev), torch.from_numpy(d).to(dev)
    shape = torch.empty(n, dtype=torch.int32, device=dev)
    prim = torch.empty(n, dtype=torch.int32, device=dev)
    t, u, v = (torch.empty(n, dtype=torch.float32, device=dev) for _ in range(3))
    torch.cuda.synchronize()
    f32p, u32p = C.POINTER(C.c_float), C.POINTER(C.c_uint32)
    rc = native.bf_trace_first_hits_static(s._h, C.cast(to.data_ptr(), f32p), C.cast(td.data_ptr(), f32p), None, None, n,
                                           C.cast(shape.data_ptr(), u32p), C.cast(prim.data_ptr(), u32p),
                                           C.cast(t.data_ptr(), f32p), C.cast(u.data_ptr(), f32p), C.cast(v.data_ptr(), f32p))
    assert rc == 0
    got = (shape.cpu().numpy().view(np.uint32), prim.cpu().numpy().view(np.uint32), t.cpu().numpy(), u.cpu().numpy(), v.cpu().numpy())
    same(got, s.trace_first_hits(o, d))
    s.close()
