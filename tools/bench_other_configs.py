"""GPU experiment: throughput of the other BASELINE configs (C3 multi-object ~1 M triangles, C2 spectral, C4 frame sequence,
C5 urban 10 M triangles x receive elements).  Not the bench line: numbers for DESIGN.md / profiles."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from beifong_b200 import configs, meshes

def run(name, s, n, reps=2, pre=None):
    t0 = time.time(); s.commit(); t1 = time.time()
    st = s.stats()
    s.receive(n, seed=1); s.synchronize()
    ms = []
    for r in range(reps):
        s.receive(n, seed=2 + r); s.synchronize(); ms.append(s.last_receive_ms()[0])
    p = s.profile()
    print(name, "commit_s", round(t1 - t0, 2), {k: st[k] for k in ("bvh_nodes", "bvh_tris", "bvh_depth", "build_ms")}, "profile", p.shape,
          "ms", min(ms), "Gsamples/s", n / min(ms) / 1e6, "W", float(np.nansum(p[..., -1 if p.shape[2] > 4 else 2])), flush=True)

which = sys.argv[1:] or ["c3", "c2s4", "c4", "c5"]
if "c3" in which:
    s = configs.c3_multi_object(1); run("C3 ~1M tris, 1024 range bins", s, 1 << 27); s.close()
if "c2s4" in which:
    s = configs.c2_car_body(4); run("C2 spectral width 4", s, 1 << 27); s.close()
if "c4" in which:
    s, tx = configs.c4_animated(0); s.commit()
    n, frames = 1 << 20, 256
    s.receive(n, seed=0); s.synchronize()
    t0 = time.time()
    for f in range(frames):
        pose = configs.c4_pose(f)
        s.update_rect(tx, (pose * meshes.Transform.scale([0.5, 0.5, 1.0])).matrix32())
        s.set_receiver(None, configs.c4_receiver(pose))
        s.receive(n, seed=0)
    s.synchronize(); dt = time.time() - t0
    print("C4 256 frames x 2^20 samples (unbounded depth, rr 5): wall s", round(dt, 3), "Gsamples/s", frames * n / dt / 1e9, flush=True)
    s.close()
if "c5" in which:
    t0 = time.time(); s, rx = configs.c5_urban(); print("C5 mesh generated s", round(time.time() - t0, 1), flush=True)
    run("C5 urban 1e7 tris element 0", s, 1 << 24)
    n, elems = 1 << 20, 8
    t0 = time.time()
    for k in range(elems):
        s.update_rect(rx, configs.c5_element_pose(k)); s.receive(n, seed=k)
    s.synchronize(); dt = time.time() - t0
    print("C5 8 elements x 2^20 samples: wall s", round(dt, 3), "Gsamples/s", elems * n / dt / 1e9, flush=True)
    s.close()
