#!/usr/bin/env python
"""Throughput of the adaptive rounds (nb_render_adaptive_round) on the 800x600 Cornell box: M samples/s with the reference's
1000-decision cap (black blocks spin to it) and with a low cap, against uniform subframes.
   python tools/adaptive_rate.py [rounds=8]"""
import json, os, re, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import norib200 as nb
opt = dict(a.split("=", 1) for a in sys.argv[1:])
rounds = int(opt.get("rounds", 8))
G = os.path.join(ROOT, "tests", "golden", "scenes", "cbox")
xml = open(os.path.join(G, "cbox_adaptive_small.xml")).read()
xml = re.sub(r'(name="width"\s+value=")\d+', r"\g<1>800", xml); xml = re.sub(r'(name="height"\s+value=")\d+', r"\g<1>600", xml)
d = tempfile.mkdtemp(prefix="adp_")
os.symlink(os.path.join(G, "meshes"), os.path.join(d, "meshes"))
open(os.path.join(d, "s.xml"), "w").write(xml)
s = nb.load_xml(os.path.join(d, "s.xml"))
ctx = nb.Context(0)
ctx.load(s)
for cap in (0, 8):
    ctx.render_begin(nb.make_render_params(s, seed=1))
    t0 = time.perf_counter()
    infos = [ctx.render_adaptive_round(4, 2, cap) for _ in range(rounds)]
    wall = time.perf_counter() - t0
    c = ctx.counters()
    print(json.dumps({"mode": "adaptive", "round_cap": cap or 1000, "rounds": rounds, "samples": c["samples"], "passes": sum(i["passes"] for i in infos),
                      "device_ms": round(c["device_ms"], 2), "wall_ms": round(wall * 1e3, 2), "Msamples_per_s_device": round(c["samples"] / c["device_ms"] / 1e3, 1),
                      "Msamples_per_s_wall": round(c["samples"] / wall / 1e6, 1)}), flush=True)
ctx.render_begin(nb.make_render_params(s, seed=1))
spp = max(1, int(round(c["samples"] / (800 * 600))))
ctx.render_subframe(0, spp); u = ctx.counters()
print(json.dumps({"mode": "uniform subframe of the same sample count", "spp": spp, "device_ms": round(u["device_ms"], 2),
                  "Msamples_per_s_device": round(u["samples"] / u["device_ms"] / 1e3, 1)}))
