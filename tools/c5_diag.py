import sys; sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
from beifong_b200 import configs
from oracle_lib import OracleScene
s, rx = configs.c5_urban(n_side=10, face_grid=4)
s.commit()
n = 1 << 17
g = s.sample_records(n, seed=0)
orc = OracleScene(s.export(), accel=OracleScene.BVH)
c = orc.sample_records(n, seed=0)
same_first = (g["first_shape"] == c["first_shape"]) & (g["first_prim"] == c["first_prim"])
print("first same", same_first.mean(), "n diff", (~same_first).sum())
agree = same_first & (g["depth"] == c["depth"]) & (g["n_closest"] == c["n_closest"]) & (g["n_shadow"] == c["n_shadow"])
print("agree", agree.mean(), (~agree).sum())
both = agree & (c["value"] != 0)
rel = np.abs(g["value"][both] - c["value"][both]) / np.abs(c["value"][both])
print("value rel: p50 %.2e p99 %.2e max %.2e" % (np.percentile(rel, 50), np.percentile(rel, 99), rel.max()))
d = np.abs(g["value"].astype(np.float64) - c["value"])
order = np.argsort(-d)[:8]
print("total |c|", np.abs(c["value"]).sum(), "sum |diff|", d.sum())
for i in order:
    print(i, d[i], g[i], c[i])
