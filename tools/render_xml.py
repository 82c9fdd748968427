#!/usr/bin/env python
"""One render of a scene XML through the C++ host loader, for profiling under ncu:
   python tools/render_xml.py scene.xml [spp=16] [size=WxH] [device_bvh=1] [repeat=1] [strategy=auto|wavefront|tile] [normalize=1] [ordered=0]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import norib200 as nb

opt = dict(a.split("=", 1) for a in sys.argv[2:])
s = nb.load_xml(sys.argv[1])
if "size" in opt:
    w, h = (int(v) for v in opt["size"].split("x"))
    cam = s["camera"].copy(); cam["width"], cam["height"] = w, h; cam["inv_output_size"] = [1.0 / w, 1.0 / h]
    s["camera"] = cam
ctx = nb.Context(0)
ctx.set_option("strategy", {"auto": 0, "wavefront": 1, "tile": 2}[opt.get("strategy", "auto")])
ctx.set_option("ordered", int(opt.get("ordered", 0)))
ctx.set_option("wide", int(opt.get("wide", 1)))
ctx.load(s, device_bvh=opt.get("device_bvh", "0") == "1")
ctx.render_begin(nb.make_render_params(s, seed=1))
for _ in range(int(opt.get("repeat", 1))):
    ctx.film_clear()
    ctx.render_subframe(0, int(opt.get("spp", 16)))
    c = ctx.counters()
    print({"device_ms": c["device_ms"], "samples_per_s": c["samples"] / c["device_ms"] * 1e3,
           "rays_per_s": (c["rays_closest"] + c["rays_any"]) / c["device_ms"] * 1e3, "launches": c["kernel_launches"], "bvh": ctx.bvh_stats()})
if opt.get("normalize", "1") == "1":
    img = ctx.film_download(normalize=True)
    print("mean", float(img.mean()))
ctx.close()
