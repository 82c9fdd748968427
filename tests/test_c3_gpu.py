"""C3: the synthetic 1 310 720-triangle mesh in the Cornell box (tools/make_c3_scene.py) — a deep tree
(~900 k nodes), where a skip-link or tie-break slip would show. Hits bit-exact against the reference's own
BVH::rayIntersect on rays the oracle generates and answers on this machine (primary, bounce and NEE shadow rays), in
the left-first walk and in the ordered walk; then a small render against the reference CPU render of the same XML."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, have_oracle_binary, rel_rmse, run_oracle
from test_trace_gpu import check

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not have_oracle_binary(), reason="oracle/_ref/nori_ref not present")]


@pytest.fixture(scope="module")
def c3(nb, tmp_path_factory):
    d = str(tmp_path_factory.mktemp("c3"))
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_c3_scene.py"), d, "width=200", "height=150", "spp=32"],
                   check=True, stdout=subprocess.DEVNULL)
    xml = os.path.join(d, "c3_path_mis.xml")
    scene = nb.load_xml(xml)                        # host loader + the reference's SAH build restated (host/bvh.cpp)
    assert scene["faces"].size // 3 == 1310720 + 12 and len(scene["bvh_nodes"]) > 500000
    rays_f, hits_f = os.path.join(d, "rays.bin"), os.path.join(d, "hits.bin")
    run_oracle("genrays", xml, rays_f, "40000", "4242")
    run_oracle("hits", xml, rays_f, hits_f)
    rays = np.fromfile(rays_f, dtype="<f4").reshape(-1, 8)
    full = np.fromfile(hits_f, dtype=np.uint8).reshape(-1, 128)
    want = np.zeros(len(full), dtype=nb.hit_dtype)
    want["shape"] = full[:, 0:4].copy().view("<i4")[:, 0]
    want["prim"] = full[:, 4:8].copy().view("<i4")[:, 0]
    tuv = full[:, 8:20].copy().view("<f4")
    want["t"], want["u"], want["v"] = tuv[:, 0], tuv[:, 1], tuv[:, 2]
    want["any_hit"] = full[:, 28:32].copy().view("<u4")[:, 0]
    return {"dir": d, "xml": xml, "scene": scene, "rays": rays, "want": want}


@pytest.mark.parametrize("mode", [0, 2])           # 0: the reference's left-first order; 2: near child first + tie rule
def test_c3_hits_bit_exact(nb, ctx, c3, mode):
    assert len(c3["rays"]) >= 100000
    on_mesh = (c3["want"]["shape"] == c3["want"]["shape"].max()).sum() + (c3["want"]["prim"] > 100).sum()
    assert on_mesh > 20000                          # the ray set really lands on the big mesh, not only on the walls
    check(nb, ctx, c3["scene"], c3["rays"], c3["want"], mode)


def test_c3_render_matches_reference(nb, ctx, c3):
    ref = os.path.join(c3["dir"], "ref.npy")
    run_oracle("render", c3["xml"], ref, "spp=32")
    want = np.load(ref)
    imgs = []
    for seed in (1, 2):
        ctx.load(c3["scene"])
        ctx.render_begin(nb.make_render_params(c3["scene"], seed=seed))
        ctx.render_subframe(0, 32)
        imgs.append(ctx.film_download(normalize=True))
    a, b = imgs
    bias = (a.mean(axis=(0, 1)) - want.mean(axis=(0, 1))) / want.mean(axis=(0, 1))
    floor, err = rel_rmse(a, b), rel_rmse(a, want)
    print("C3 200x150x32spp: bias", bias, "relRMSE", err, "noise floor", floor)
    assert np.all(np.abs(bias) <= 0.01), bias
    assert err <= 1.25 * floor + 2e-3, (err, floor)


def test_c3_wide_hit_parity(nb, ctx, c3):
    """The production traversal on the deep tree: same primitive on all but at most one ray in a million."""
    from test_wide_gpu import parity
    total, bad, bad_any = parity(nb, ctx, c3["scene"], c3["rays"], c3["want"])
    st = ctx.bvh_stats()
    print("C3 wide: rays", total, "closest-hit mismatches", bad, "shadow mismatches", bad_any, st)
    assert bad <= 1 and bad_any <= 1
