#!/usr/bin/env python
"""C3 (1.31 M-triangle synthetic mesh): hit parity against oracle answers prepared beforehand
(c3cache/, made where the oracle binary and the generator can run) and render throughput of both
strategies.  python tools/c3_check.py [spp=16]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import norib200 as nb

opt = dict(a.split("=", 1) for a in sys.argv[1:])
cache = os.path.join(ROOT, "c3cache")
scene = nb.read_nbs(os.path.join(cache, "c3.nbs"))
ctx = nb.Context(0)
t0 = time.perf_counter()
ctx.load(scene)
out = {"upload_s": time.perf_counter() - t0, "triangles": int(scene["faces"].size // 3), "nodes": int(len(scene["bvh_nodes"]))}
rays_f, hits_f = os.path.join(cache, "rays.bin"), os.path.join(cache, "hits_full.bin")
if os.path.exists(hits_f):
    rays = np.fromfile(rays_f, dtype="<f4").reshape(-1, 8)
    full = np.fromfile(hits_f, dtype=np.uint8).reshape(-1, 128)
    got = ctx.trace_rays(rays, any_hit=True)
    shape = full[:, 0:4].copy().view("<i4")[:, 0]; prim = full[:, 4:8].copy().view("<i4")[:, 0]
    tuv = full[:, 8:20].copy().view("<f4"); anyh = full[:, 28:32].copy().view("<u4")[:, 0]
    hit = shape >= 0
    out["parity"] = {"rays": int(len(rays)), "shape_equal": bool(np.array_equal(got["shape"], shape)),
                     "prim_equal": bool(np.array_equal(got["prim"], prim)),
                     "tuv_bits_equal": bool(all(got[k][hit].tobytes() == tuv[hit, i].tobytes() for i, k in enumerate("tuv"))),
                     "any_equal": bool(np.array_equal(got["any_hit"], anyh)),
                     "mean_nodes": float(full[:, 20:24].copy().view("<u4").mean()), "mean_prims": float(full[:, 24:28].copy().view("<u4").mean())}
spp = int(opt.get("spp", 16))
for strat, code in (("tile", 2), ("wavefront", 1)):
    ctx.set_option("strategy", code)
    ctx.render_begin(nb.make_render_params(scene, "path_mis", seed=1))
    for rep in range(2):
        ctx.film_clear()
        ctx.render_subframe(0, spp)
        c = ctx.counters()
    img = ctx.film_download(normalize=True)
    out[strat] = {"device_ms": c["device_ms"], "samples_per_s": c["samples"] / c["device_ms"] * 1e3,
                  "rays_per_s": (c["rays_closest"] + c["rays_any"]) / c["device_ms"] * 1e3, "mean": [float(x) for x in img.mean(axis=(0, 1))]}
print(json.dumps(out))
