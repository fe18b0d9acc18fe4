"""GPU experiment: bf_scene_commit's estimate of the share of receiver rays entering the mesh bounds, per config, and the step time with the
deferred / k_path pipeline forced on and off."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from beifong_b200 import configs
for name, mk, n in (("c2", lambda: bench.build_config("c2")[0], 1 << 26), ("c3", lambda: bench.build_config("c3")[0], 1 << 26),
                    ("small", lambda: configs.small_radar(1), 1 << 24), ("c5", lambda: bench.build_config("c5")[0], 1 << 22)):
    s = mk(); s.commit()
    st = s.stats()
    row = [f"{name}: fraction {st['mesh_bounds_fraction']:.3f}"]
    for mode in (0, 1):
        s.set_tunable("primary_defer", mode); s.set_tunable("trace_defer", mode)
        s.receive(n, seed=1); s.synchronize()
        best = 1e9
        for r in range(3):
            s.receive(n, seed=2 + r); s.synchronize(); best = min(best, s.last_receive_ms()[0])
        row.append(f"defer={mode}: {n / best / 1e6:.3f} Gsamples/s")
    print("  ".join(row), flush=True)
    s.close()
