"""Which builder is right where the linear BVH and the SAH tree disagree on the 400 k clustered soup? Brute force in float64."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import norib200 as nb
from test_bvh_build_gpu import synthetic_scene
n, seed = 400000, 3
rng = np.random.default_rng(seed)
c = rng.normal(size=(n, 1, 3)) * 0.05 + rng.integers(0, 5, (n, 1, 1)) * 3.0
tris = (c + rng.normal(size=(n, 3, 3)) * 0.01).astype(np.float32)
s = synthetic_scene(nb, tris)
ctx = nb.Context(0); ctx.set_option("flat_scan", 0); ctx.upload_scene(s)
lo, hi = tris.reshape(-1, 3).min(axis=0), tris.reshape(-1, 3).max(axis=0)
m = 200000
a = rng.uniform(lo, hi, (m, 3)); b = tris[rng.integers(0, n, m)].mean(axis=1) + rng.normal(size=(m, 3)) * 0.01
d = b - a; d /= np.linalg.norm(d, axis=1, keepdims=True)
rays = np.concatenate([a, d, np.full((m, 1), 1e-4), np.full((m, 1), np.inf)], axis=1).astype("<f4")
out = {}
for sah in (False, True):
    ctx.build_bvh_device(sah=sah); out[sah] = ctx.trace_rays(rays, any_hit=True)
x, y = out[False], out[True]
bad = np.nonzero((x["t"].view("<u4") != y["t"].view("<u4")) | (x["prim"] != y["prim"]))[0]
print("differing rays", bad)
T = tris.astype(np.float64)
for i in bad:
    o, dd = rays[i, :3].astype(np.float64), rays[i, 3:6].astype(np.float64)
    e1, e2 = T[:, 1] - T[:, 0], T[:, 2] - T[:, 0]
    p = np.cross(dd, e2); det = (e1 * p).sum(1)
    tv = o - T[:, 0]; u = (tv * p).sum(1) / det
    q = np.cross(tv, e1); v = (q * dd).sum(1) / det; t = (e2 * q).sum(1) / det
    ok = (np.abs(det) > 1e-8) & (u >= 0) & (v >= 0) & (u + v <= 1) & (t > 1e-4)
    k = np.nonzero(ok)[0]; k = k[np.argsort(t[k])][:3]
    print("ray", i, "lbvh", (x["prim"][i], x["t"][i]), "sah", (y["prim"][i], y["t"][i]), "brute force nearest", [(int(j), float(t[j]), float(u[j]), float(v[j])) for j in k])
