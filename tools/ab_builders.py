#!/usr/bin/env python
"""The three tree builders on the tree scenes, production traversal: host SAH (imported), device linear BVH, device SAH.
   python tools/ab_builders.py [c3=1] [c5=1] [spp=N] [reps=3]"""
import json, os, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import norib200 as nb
opt = dict(a.split("=", 1) for a in sys.argv[1:])
G = os.path.join(ROOT, "tests", "golden", "scenes")
scenes = [("table", os.path.join(G, "table", "table_path_mis.xml"), 64)]
if opt.get("c3", "1") == "1":
    d = tempfile.mkdtemp(prefix="c3_")
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_c3_scene.py"), d, "spp=16"], check=True, stdout=subprocess.DEVNULL)
    scenes.append(("c3", os.path.join(d, "c3_path_mis.xml"), 16))
if opt.get("c5", "0") == "1":
    d = tempfile.mkdtemp(prefix="c5_")
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_c5_scene.py"), d, "width=1920", "height=1080", "spp=8"], check=True, stdout=subprocess.DEVNULL)
    scenes.append(("c5", os.path.join(d, "c5_path_mis.xml"), 8))
ctx = nb.Context(0)
for name, xml, spp in scenes:
    spp = int(opt.get("spp", spp))
    t0 = time.perf_counter(); s = nb.load_xml(xml); host = time.perf_counter() - t0
    for builder in (False, True, "sah"):
        t0 = time.perf_counter(); ctx.load(s, device_bvh=builder); load_s = time.perf_counter() - t0
        ctx.render_begin(nb.make_render_params(s, seed=1))
        best = None
        for _ in range(int(opt.get("reps", 3))):
            ctx.film_clear(); ctx.render_subframe(0, spp); c = ctx.counters()
            if best is None or c["device_ms"] < best["device_ms"]: best = c
        img = ctx.film_download(normalize=True)
        print(json.dumps({"scene": name, "builder": {False: "host_sah", True: "device_lbvh", "sah": "device_sah"}[builder], "spp": spp,
                          "ms": round(best["device_ms"], 3), "Msamples_per_s": round(best["samples"] / best["device_ms"] / 1e3, 1),
                          "mean": float(img.mean()), "bvh": ctx.bvh_stats(), "upload_and_build_s": round(load_s, 3),
                          "host_xml_and_sah_s": round(host, 2)}), flush=True)
