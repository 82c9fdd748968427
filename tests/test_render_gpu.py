"""Image parity: the wavefront renderer against the reference CPU integrators on the same scene
at equal spp (committed reference renders of the *_small scenes, tests/golden/make_fixtures.py).
Streams differ (per-pixel pcg32 here, per-block in the reference), so agreement is statistical:
  |mean bias| <= 1 %  and  relRMSE(gpu, ref) <= 1.25 x the noise floor measured between two
  independent GPU renders (both images carry the same Monte-Carlo variance)."""
import numpy as np
import pytest

from conftest import golden, rel_rmse

pytestmark = pytest.mark.gpu


def render(nb, ctx, scene, integrator, spp, size=None, seed=1, pool=0):
    s = dict(scene)
    if size:
        cam = scene["camera"].copy()
        cam["width"], cam["height"] = size
        cam["inv_output_size"] = [1.0 / size[0], 1.0 / size[1]]
        s["camera"] = cam
    ctx.load(s)
    ctx.render_begin(nb.make_render_params(s, integrator, seed=seed, pool_paths=pool))
    ctx.render_subframe(0, spp)
    return ctx.film_download(normalize=True), ctx.counters()


@pytest.mark.parametrize("name,integrator,spp,ref", [
    ("cbox", "path_mis", 256, "cbox_ref_256spp.npy"),
    ("cbox_mats", "path_mats", 256, "cbox_mats_ref_256spp.npy"),
    ("table", "path_mis", 64, "table_ref_64spp.npy"),
])
def test_image_matches_reference(nb, ctx, name, integrator, spp, ref):
    scene = nb.read_nbs(golden(name + ".nbs"))
    want = np.load(golden(ref))
    h, w, _ = want.shape
    a, ca = render(nb, ctx, scene, integrator, spp, size=(w, h), seed=1)
    b, _ = render(nb, ctx, scene, integrator, spp, size=(w, h), seed=2)
    assert ca["samples"] == w * h * spp
    assert np.isfinite(a).all()
    bias = (a.mean(axis=(0, 1)) - want.mean(axis=(0, 1))) / want.mean(axis=(0, 1))
    floor = rel_rmse(a, b)
    err = rel_rmse(a, want)
    print(name, "bias", bias, "relRMSE", err, "noise floor", floor, "invalid", ca["invalid_samples"])
    assert np.all(np.abs(bias) <= 0.01), bias
    assert err <= 1.25 * floor + 1e-3, (err, floor)


def test_sample_slices_add_up(nb, ctx):
    """Sample s of pixel p uses the same stream whatever the launch split: 8 spp in one call
    equals 2 slices of 4 (stride 2) accumulated into the same film, up to fp32 add order."""
    scene = nb.read_nbs(golden("cbox.nbs"))
    size = (96, 72)
    a, _ = render(nb, ctx, scene, "path_mis", 8, size=size, seed=5)
    ctx.film_clear()
    ctx.render_subframe(0, 8, 2, 0)
    ctx.render_subframe(0, 8, 2, 1)
    b = ctx.film_download(normalize=True)
    assert ctx.counters()["samples"] == size[0] * size[1] * 8
    np.testing.assert_allclose(a, b, rtol=2e-4, atol=2e-5)


def test_small_pool_same_image(nb, ctx):
    """The pool size only changes scheduling (regeneration order), never which samples exist."""
    scene = nb.read_nbs(golden("cbox.nbs"))
    a, _ = render(nb, ctx, scene, "path_mis", 4, size=(64, 48), seed=3, pool=0)
    b, cb = render(nb, ctx, scene, "path_mis", 4, size=(64, 48), seed=3, pool=1024)
    assert cb["samples"] == 64 * 48 * 4
    np.testing.assert_allclose(a, b, rtol=2e-4, atol=2e-5)


def test_cli_renders_reference_xml(nb, tmp_path):
    """The C++ host path end to end: nori_b200 loads the reference's own scene XML (parser, OBJ
    loader, SAH BVH build, flatten hooks), renders on the GPU through the C ABI, writes .npy."""
    import json
    import os
    import subprocess
    cli = os.path.join(os.path.dirname(nb.LIB_PATH), "nori_b200")
    out = tmp_path / "cbox.npy"
    r = subprocess.run([cli, golden("scenes/cbox/cbox_path_mis_small.xml"), str(out), "spp=128", "seed=4"],
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 0, r.stderr
    stats = json.loads(r.stdout.strip().splitlines()[-1])
    img = np.load(out)
    want = np.load(golden("cbox_ref_256spp.npy"))
    assert img.shape == want.shape and stats["samples"] == 200 * 150 * 128
    bias = (img.mean(axis=(0, 1)) - want.mean(axis=(0, 1))) / want.mean(axis=(0, 1))
    assert np.all(np.abs(bias) <= 0.01), bias
    assert rel_rmse(img, want) < 0.2
