"""GPU experiment: microseconds per receive() call against the request size, k_mega (one launch) vs the wavefront sequence."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from beifong_b200 import configs
s = configs.c2_car_body(1).commit()
for n in (1024, 4096, 16384, 65536, 262144, 1048576):
    row = []
    for mode in (1 << 28, -1):
        s.set_tunable("mega_max", mode)
        for _ in range(10):
            s.receive(n, seed=1)
        s.synchronize()
        reps = 100
        t0 = time.perf_counter()
        for i in range(reps):
            s.receive(n, seed=i)
        s.synchronize()
        row.append((time.perf_counter() - t0) / reps * 1e6)
    print(f"n {n:8d}  mega {row[0]:9.1f} us  wavefront {row[1]:9.1f} us", flush=True)
