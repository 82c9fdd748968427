#!/usr/bin/env python
"""Throughput of every measured configuration, GPU (this library) beside the reference CPU renderer
on the same machine: C1 cbox, C2 table / veach, C3 synthetic 1.31M triangles (if c3cache/ exists),
C4 homogeneous-volume scenes. One JSON object per line.   python tools/config_table.py [cpu=1]"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import norib200 as nb

G = os.path.join(ROOT, "tests", "golden", "scenes")
NORI_REF = os.path.join(ROOT, "oracle", "_ref", "nori_ref")
opt = dict(a.split("=", 1) for a in sys.argv[1:])
CONFIGS = [  # name, xml, gpu spp, cpu spp
    ("C1 cbox path_mis 800x600", os.path.join(G, "cbox", "cbox_path_mis.xml"), 256, 16),
    ("C1 cbox path_mats 800x600", os.path.join(G, "cbox", "cbox_path_mats.xml"), 256, 16),
    ("C2 table path_mis 800x600", os.path.join(G, "table", "table_path_mis.xml"), 64, 8),
    ("C2 veach direct_mis 768x512", os.path.join(G, "veach", "veach_mis.xml"), 128, 32),
    ("C2 odyssey direct_ems 768x432", os.path.join(G, "odyssey", "odyssey_ems.xml"), 128, 32),
    ("C4 cbox_val_mis path_vol_mis 600x600", os.path.join(G, "vol", "cbox_val_mis.xml"), 128, 16),
    ("C4 cbox_homog_aniso path_vol_mats 800x600", os.path.join(G, "vol", "cbox_homog_aniso.xml"), 128, 16),
    ("C4 heterog fluid (NanoVDB-class grid) path_vol_mats 800x600", os.path.join(G, "vol", "cbox_heterog_fluid.xml"), 128, 8),
    ("C4 heterog fluid path_vol_mis 800x600", os.path.join(G, "vol", "cbox_heterog_fluid_mis.xml"), 128, 8),
    ("envmap_sphere path_mis 800x800 (1500x750 PNG environment)", os.path.join(G, "envmap", "envmap_sphere.xml"), 128, 16),
]
P = os.path.join(G, "plugins")
CONFIGS += [  # the 8(f) rows at a measurable size (their fixtures are 160x120): same XML with the camera enlarged
    ("disney + spot + directional path_mis 800x600", os.path.join(P, "disney_spot_dir_path_mis.xml"), 128, 16, (800, 600)),
    ("normal maps direct_mis 800x600", os.path.join(P, "normalmap_direct_mis.xml"), 128, 32, (800, 600)),
    ("volume lights path_vol_mis 800x600", os.path.join(P, "volumelight_mis.xml"), 64, 8, (800, 600)),
]
if "only" in opt:
    CONFIGS = [c for c in CONFIGS if opt["only"] in c[0]]
c3 = os.path.join(ROOT, "c3cache", "c3_path_mis.xml")
if os.path.exists(c3):
    CONFIGS.append(("C3 synthetic 1310720 tris path_mis 800x600", c3, 32, 4))
c5 = os.path.join(ROOT, "c5cache", "c5_path_mis.xml")
if os.path.exists(c5):
    CONFIGS.append(("C5-class 64 x camelhead 1453066 tris path_mis 1920x1080 (one GPU)", c5, 16, 2))
ctx = nb.Context(0)
for cfg in CONFIGS:
    name, xml, gspp, cspp = cfg[:4]
    if len(cfg) > 4:                     # enlarged copy next to the original (relative mesh paths keep working)
        import re
        text = open(xml).read()
        text = re.sub(r'(name="width"\s+value=")\d+', r"\g<1>%d" % cfg[4][0], text)
        text = re.sub(r'(name="height"\s+value=")\d+', r"\g<1>%d" % cfg[4][1], text)
        xml = xml[:-4] + "_cfgtable.xml"
        open(xml, "w").write(text)
    t0 = time.perf_counter()
    s = nb.load_xml(xml)
    host_s = time.perf_counter() - t0
    ctx.load(s)
    ctx.render_begin(nb.make_render_params(s, seed=1))
    for _ in range(2):
        ctx.film_clear()
        ctx.render_subframe(0, gspp)
    c = ctx.counters()
    row = {"config": name, "triangles": int(s["faces"].size // 3), "bvh_nodes": int(len(s["bvh_nodes"])),
           "host_load_flatten_bvh_s": round(host_s, 3), "gpu_spp": gspp, "gpu_ms": c["device_ms"],
           "gpu_samples_per_s": c["samples"] / c["device_ms"] * 1e3,
           "gpu_rays_per_s": (c["rays_closest"] + c["rays_any"]) / c["device_ms"] * 1e3,
           "rays_per_sample": (c["rays_closest"] + c["rays_any"]) / c["samples"]}
    if opt.get("cpu", "1") == "1" and os.path.exists(NORI_REF):
        r = subprocess.run([NORI_REF, "render", xml, "/tmp/cfg_ref.npy", "spp=%d" % cspp], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        if r.returncode == 0:
            j = json.loads(r.stdout.strip().splitlines()[-1])
            row.update({"cpu_threads": j["threads"], "cpu_spp": cspp, "cpu_samples_per_s": j["samples_per_s"],
                        "cpu_rays_raw_per_s": j["rays_raw_per_s"], "cpu_rays_dedup_per_s": j["rays_dedup_per_s"],
                        "cpu_load_bvh_s": j["load_seconds"], "speedup_samples": row["gpu_samples_per_s"] / j["samples_per_s"]})
    print(json.dumps(row), flush=True)
    if len(cfg) > 4:
        os.remove(xml)
