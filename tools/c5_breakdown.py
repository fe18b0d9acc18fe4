"""GPU experiment: per-kernel device time of one wave of the C5 urban stand-in (1e7 triangles)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from beifong_b200 import configs
lanes = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 22
n_side = int(sys.argv[2]) if len(sys.argv) > 2 else 100
os.environ["BF_WAVE_LANES"] = str(lanes); os.environ["BF_WAVE_STREAMS"] = "1"
s, rx = configs.c5_urban(n_side=n_side)
s.commit()
s.receive(lanes, seed=1); s.synchronize()
s.set_trace_timing(True)
a = s.stats()
s.receive(lanes, seed=2); s.synchronize()
b = s.stats()
print("lanes", lanes, "total ms", s.last_receive_ms())
print("closest", b["closest_queries"] - a["closest_queries"], "shadow", b["shadow_queries"] - a["shadow_queries"])
for k, ms in s.last_launch_list():
    print(f"{k:14s} {ms * 1e3:9.1f} us")
print(s.last_kernel_ms())
