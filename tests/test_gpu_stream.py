"""Caller-supplied stream: bf_scene_set_stream makes every entry point of a scene enqueue on the application's
stream; the waves still fan out over internal streams but fork from / join into it.  Also the L2 access-policy window over the top
of the BVH (tunable `l2_persist_mb`): a cache hint, results must not move."""
import numpy as np
import pytest

from beifong_b200 import configs

pytestmark = pytest.mark.gpu


def same_profile(a, b):
    # counts are integers and exact; the value channel is a sum of fp32 atomics whose order differs from run to run
    assert np.array_equal(a[..., 1:3], b[..., 1:3])
    assert np.linalg.norm(a[..., 0].astype(np.float64) - b[..., 0]) <= 1e-5 * np.linalg.norm(b[..., 0].astype(np.float64))


@pytest.mark.parametrize("pipeline", ["wavefront", "mega"])
def test_receive_on_a_caller_stream(native, pipeline):
    import torch
    s = configs.small_radar(spectrum_width=1).commit()
    s.set_tunable("mega_max", (1 << 28) if pipeline == "mega" else -1)
    n = 1 << 18
    s.receive(n, seed=4)
    base = s.profile()
    own = s.profile_device()[2]
    st = torch.cuda.Stream(device=0)
    s.set_stream(st.cuda_stream)
    assert s.profile_device()[2] == st.cuda_stream
    # ordered behind the caller's earlier work on that stream and in front of its later work: a device-side copy of the profile
    # enqueued by the CALLER on the same stream right after bf_receive (no host synchronisation in between) sees the finished sums
    ptr, nf, _ = s.profile_device()
    import ctypes
    rt = ctypes.CDLL("libcudart.so.12")           # the runtime torch has loaded; pointers and streams are per context, not per runtime copy
    rt.cudaMemcpyAsync.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
    with torch.cuda.stream(st):
        busy = torch.empty(1 << 26, device="cuda:0").normal_()        # earlier work
        s.receive(n, seed=4)
        snap = torch.zeros(nf, dtype=torch.float32, device="cuda:0")
        assert rt.cudaMemcpyAsync(snap.data_ptr(), ptr, nf * 4, 3, st.cuda_stream) == 0      # 3 = cudaMemcpyDeviceToDevice
    st.synchronize()
    same_profile(snap.cpu().numpy().reshape(base.shape), base)
    del busy
    # records / first hits go through the same stream
    o = np.array([[0.0, 0.0, 0.3]], np.float32); d = np.array([[1.0, 0.0, -0.05]], np.float32); d /= np.linalg.norm(d)
    hit_a = s.trace_first_hits(o, d)
    s.set_stream(None)
    assert s.profile_device()[2] == own
    hit_b = s.trace_first_hits(o, d)
    assert all(np.array_equal(x, y) for x, y in zip(hit_a, hit_b))
    s.receive(n, seed=4)
    same_profile(s.profile(), base)
    s.close()


def test_l2_persistence_window_changes_no_result(native):
    s = configs.small_radar(spectrum_width=1).commit()
    n = 1 << 18
    s.receive(n, seed=6)
    base = s.profile()
    rec0 = s.sample_records(20000, seed=6)
    s.set_tunable("l2_persist_mb", 16)
    s.receive(n, seed=6)
    same_profile(s.profile(), base)
    assert s.sample_records(20000, seed=6).tobytes() == rec0.tobytes()
    s.set_tunable("l2_persist_mb", 0)
    s.receive(n, seed=6)
    same_profile(s.profile(), base)
    s.close()
