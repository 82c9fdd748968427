import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import norib200 as nb
scene = nb.read_nbs("tests/golden/cbox.nbs")
cam = scene["camera"].copy(); cam["width"], cam["height"] = 200, 150; cam["inv_output_size"] = [1/200, 1/150]; scene["camera"] = cam
def single(first, count, stride=1, off=0, dev=0):
    with nb.Context(dev) as c:
        c.load(scene); c.render_begin(nb.make_render_params(scene, "path_mis", seed=7))
        c.render_subframe(first, count, stride, off)
        return c.film_download(normalize=False)
full = single(0, 64)
parts = [single(0, 64, 2, g, dev=g) for g in range(2)]
print("full sum", full.sum(), "parts", [p.sum() for p in parts], "sum of parts", (parts[0] + parts[1]).sum())
ctxs = nb.create_multi([0, 1])
for c in ctxs:
    c.load(scene); c.render_begin(nb.make_render_params(scene, "path_mis", seed=7))
for g, c in enumerate(ctxs):
    c.render_subframe(0, 64, 2, g)
pre = [c.film_download(normalize=False) for c in ctxs]
print("multi pre-reduce sums", [p.sum() for p in pre])
ctxs[0].film_reduce()
post = [c.film_download(normalize=False) for c in ctxs]
print("multi post-reduce sums", [p.sum() for p in post])
for c in ctxs: c.close()
