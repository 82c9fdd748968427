"""C5: 64 x camelhead.obj (1 453 066 triangles), four BSDF kinds, 4 area lights + a
point light, path_mis — rendered small against the reference CPU render of the same XML, on the
host-built SAH tree (the reference's, bit for bit) and on the device-built tree."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import golden, rel_rmse

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def c5_scene(nb, tmp_path_factory):
    d = tmp_path_factory.mktemp("c5")
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_c5_scene.py"), str(d), "width=384", "height=216", "spp=16"],
                   check=True, stdout=subprocess.DEVNULL)
    return nb.load_xml(os.path.join(str(d), "c5_path_mis.xml"))


@pytest.mark.parametrize("device_bvh", [False, True])
def test_c5_matches_reference_render(nb, ctx, c5_scene, device_bvh):
    s = c5_scene
    assert s["faces"].size // 3 == 64 * 22704 + 2 + 8 and len(s["emitters"]) == 5
    want = np.load(golden("c5_small_ref_32spp.npy"))
    imgs = []
    for seed in (1, 2):
        ctx.load(s, device_bvh=device_bvh)
        ctx.render_begin(nb.make_render_params(s, seed=seed))
        ctx.render_subframe(0, 32)
        imgs.append(ctx.film_download(normalize=True))
    a, b = imgs
    st = ctx.bvh_stats()
    bias = (a.mean(axis=(0, 1)) - want.mean(axis=(0, 1))) / want.mean(axis=(0, 1))
    floor, err = rel_rmse(a, b), rel_rmse(a, want)
    c = ctx.counters()
    print("device_bvh", device_bvh, st, "bias", bias, "relRMSE", err, "noise floor", floor,
          "Msamples/s", c["samples"] / c["device_ms"] / 1e3)
    assert a.shape == want.shape and np.isfinite(a).all()
    assert np.all(np.abs(bias) <= 0.02), bias
    assert err <= 1.25 * floor + 2e-3, (err, floor)
