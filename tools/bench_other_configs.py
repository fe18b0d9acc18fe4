import sys, time
sys.path.insert(0, "/root/repo")
import numpy as np
from beifong_b200 import configs
for name, mk in (("c3", lambda: configs.c3_multi_object(1)), ("c2_s4", lambda: configs.c2_car_body(4))):
    t0 = time.time(); s = mk().commit(); t1 = time.time()
    st = s.stats()
    n = 1 << 26
    s.receive(n, seed=1); s.synchronize()
    s.receive(n, seed=2); s.synchronize()
    ms = s.last_receive_ms()[0]
    p = s.profile()
    print(name, "commit s", round(t1 - t0, 2), {k: st[k] for k in ("bvh_nodes", "bvh_tris", "bvh_depth")}, "profile", p.shape, "ms", ms, "Gsamples/s", n / ms / 1e6,
          "W sum", float(p[..., 2].sum()), "Y sum", float(p[..., 0].sum()), flush=True)
    s.close()
