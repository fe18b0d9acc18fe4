"""GPU experiment: distribution of BVH traversal steps per closest-hit query for primary-like and bounce-like rays."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from beifong_b200 import configs, meshes
s = configs.c2_car_body(1).commit()
p, f = meshes.car_body(264, 133)
p = meshes.transform_points(configs.car_pose().matrix32(), p)
rng = np.random.default_rng(0)
n = 1 << 20
# primary-like: from the antenna, cosine hemisphere about +x
u = rng.random((n, 2)); r = np.sqrt(u[:, 0]); ph = 2 * np.pi * u[:, 1]
d = np.stack([np.sqrt(1 - u[:, 0]), r * np.cos(ph), r * np.sin(ph)], 1)
o = np.tile(np.array([[0, 0, 0.3]]), (n, 1))
st, t = s.trace_steps(o, d)
def rep(name, st):
    q = np.percentile(st, [50, 90, 99, 99.9, 99.99])
    print(name, "mean %.2f" % st.mean(), "pcts", q, "max", st.max(), "frac0", (st == 0).mean(), flush=True)
rep("primary", st)
# bounce-like: from random surface points, cosine hemisphere about the (outward or inward) face normal
tri = rng.integers(0, len(f), n)
a, b, c = p[f[tri, 0]], p[f[tri, 1]], p[f[tri, 2]]
w = rng.random((n, 2)); sw = np.sqrt(w[:, 0]); b0 = 1 - sw; b1 = sw * (1 - w[:, 1]); b2 = sw * w[:, 1]
pt = b0[:, None] * a + b1[:, None] * b + b2[:, None] * c
nr = np.cross(b - a, c - a); nr /= np.linalg.norm(nr, axis=1, keepdims=True) + 1e-30
u = rng.random((n, 2)); r = np.sqrt(u[:, 0]); ph = 2 * np.pi * u[:, 1]
tz = np.sqrt(1 - u[:, 0])
hlp = np.where(np.abs(nr[:, :1]) > 0.9, np.array([[0, 1.0, 0]]), np.array([[1.0, 0, 0]]))
tx = np.cross(hlp, nr); tx /= np.linalg.norm(tx, axis=1, keepdims=True); ty = np.cross(nr, tx)
d = tx * (r * np.cos(ph))[:, None] + ty * (r * np.sin(ph))[:, None] + nr * tz[:, None]
st2, t2 = s.trace_steps(pt + 1e-4 * nr, d)
rep("bounce", st2)
worst = np.argsort(st2)[-5:]
for i in worst:
    print("worst", st2[i], "o", pt[i], "d", d[i], "n", nr[i], "cos", tz[i], "t", t2[i])
print(s.stats())
