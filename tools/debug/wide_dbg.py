import os, sys, subprocess
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import norib200 as nb
G = "tests/golden"
REF = "oracle/_ref/nori_ref"
name, xml, n, seed = sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4]
subprocess.run([REF, "genrays", xml, "/tmp/r.bin", n, seed], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
subprocess.run([REF, "hits", xml, "/tmp/r.bin", "/tmp/h.bin"], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
rays = np.fromfile("/tmp/r.bin", dtype="<f4").reshape(-1, 8)
full = np.fromfile("/tmp/h.bin", dtype=np.uint8).reshape(-1, 128)
want = np.zeros(len(full), dtype=nb.hit_dtype)
want["shape"] = full[:, 0:4].copy().view("<i4")[:, 0]; want["prim"] = full[:, 4:8].copy().view("<i4")[:, 0]
tuv = full[:, 8:20].copy().view("<f4"); want["t"], want["u"], want["v"] = tuv[:, 0], tuv[:, 1], tuv[:, 2]
want["any_hit"] = full[:, 28:32].copy().view("<u4")[:, 0]
scene = nb.read_nbs(os.path.join(G, name + ".nbs"))
ctx = nb.Context(0)
res = {}
for label, opts in [("walk", dict(flat_scan=0, trace_wide=0, ordered=0)), ("ordered", dict(flat_scan=0, trace_wide=0, ordered=1)),
                    ("flat", dict(flat_scan=1, trace_wide=0, ordered=0)), ("wide", dict(flat_scan=0, trace_wide=1, ordered=0))]:
    for k, v in opts.items(): ctx.set_option(k, v)
    ctx.load(scene)
    res[label] = ctx.trace_rays(rays, any_hit=True)
    g = res[label]
    bad = (g["shape"] != want["shape"]) | (g["prim"] != want["prim"])
    print(label, "closest mismatches", int(bad.sum()), "shadow mismatches", int((g["any_hit"] != want["any_hit"]).sum()))
g = res["wide"]
bad = np.nonzero((g["shape"] != want["shape"]) | (g["prim"] != want["prim"]) | (g["any_hit"] != want["any_hit"]))[0]
for i in bad[:12]:
    print(i, "ray o", rays[i, :3], "d", rays[i, 3:6], "mint/maxt", rays[i, 6:8])
    for label in ("walk", "flat", "wide"):
        r = res[label][i]; print("   ", label, r["shape"], r["prim"], repr(float(r["t"])), r["any_hit"])
    print("    oracle", want[i]["shape"], want[i]["prim"], repr(float(want[i]["t"])), want[i]["any_hit"])
