#!/usr/bin/env python
"""One render of a golden scene, for profiling under ncu:
   python tools/render_once.py [scene=cbox] [spp=64] [strategy=tile|wavefront] [size=800x600]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import norib200 as nb

opt = dict(a.split("=", 1) for a in sys.argv[1:])
scene = nb.read_nbs(os.path.join(os.path.dirname(nb.__file__), "..", "tests", "golden", opt.get("scene", "cbox") + ".nbs"))
if "size" in opt:
    w, h = (int(v) for v in opt["size"].split("x"))
    scene["camera"]["width"], scene["camera"]["height"] = w, h
    scene["camera"]["inv_output_size"] = [1.0 / w, 1.0 / h]
ctx = nb.Context(0)
ctx.set_option("strategy", {"auto": 0, "wavefront": 1, "tile": 2}[opt.get("strategy", "tile")])
ctx.load(scene)
for k in ("walk", "flat_scan"):
    if k in opt:
        ctx.set_option(k, int(opt[k]))
ctx.render_begin(nb.make_render_params(scene, opt.get("integrator"), seed=1))
for _ in range(int(opt.get("repeat", 1))):
    ctx.film_clear()
    t0 = time.perf_counter()
    ctx.render_subframe(0, int(opt.get("spp", 64)))
    c = ctx.counters()
    print({"wall_ms": 1e3 * (time.perf_counter() - t0), "device_ms": c["device_ms"], "samples_per_s": c["samples"] / c["device_ms"] * 1e3,
           "rays_per_s": (c["rays_closest"] + c["rays_any"]) / c["device_ms"] * 1e3, "launches": c["kernel_launches"]})
ctx.close()
