"""GPU experiment: one wave of config C3 (975 744 triangles), per-kernel times; used under ncu for the traversal kernels."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
lanes = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 24
os.environ["BF_WAVE_LANES"] = str(lanes); os.environ["BF_WAVE_STREAMS"] = "1"
s, _, _ = bench.build_config("c3")
s.commit()
s.receive(lanes, seed=1); s.synchronize()
s.set_trace_timing(True)
s.receive(lanes, seed=2); s.synchronize()
for k, ms in s.last_launch_list():
    print(f"{k:14s} {ms * 1e3:9.1f} us")
print(s.last_kernel_ms())
