import sys, os
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
from beifong_b200 import configs
from oracle_lib import OracleScene
s = configs.c2_car_body(1).commit()
n = 1 << 24
found = []
for chunk in range(4):
    r = s.sample_records(n, seed=2, first_sample=chunk * n)
    bad = np.nonzero(~np.isfinite(r["value"]))[0]
    print("chunk", chunk, "non-finite values:", len(bad), flush=True)
    for b in bad[:3]:
        found.append(chunk * n + int(b)); print(r[b])
    if len(found) >= 3: break
s.receive(1 << 26, seed=2); p = s.profile()
print("profile nan bins:", np.argwhere(~np.isfinite(p))[:10].tolist())
orc = OracleScene(s.export(), accel=OracleScene.BVH)
for idx in found[:3]:
    c = orc.sample_records(1, seed=2, first=idx)
    print("oracle", idx, c[0])
