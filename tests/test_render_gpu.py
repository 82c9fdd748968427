"""Image parity: the wavefront renderer against the reference CPU integrators on the same scene
at equal spp (committed reference renders of the *_small scenes, tests/golden/make_fixtures.py).
Streams differ (per-pixel pcg32 here, per-block in the reference), so agreement is statistical:
  |mean bias| <= 1 %  and  relRMSE(gpu, ref) <= 1.25 x the noise floor measured between two
  independent GPU renders (both images carry the same Monte-Carlo variance)."""
import numpy as np
import pytest

from conftest import golden, rel_rmse

pytestmark = pytest.mark.gpu


def render(nb, ctx, scene, integrator, spp, size=None, seed=1, pool=0, strategy=0):
    ctx.set_option("strategy", strategy)
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


@pytest.mark.parametrize("xml,ref,spp", [
    ("veach/veach_mis_small.xml", "veach_ref_64spp.npy", 64),          # direct_mis, microfacet, 4 mesh lights
    ("veach/veach_ems_small.xml", "veach_ems_ref_64spp.npy", 64),      # direct_ems
    ("veach/veach_mats_small.xml", "veach_mats_ref_64spp.npy", 64),    # direct_mats
    ("odyssey/odyssey_ems_small.xml", "odyssey_ref_64spp.npy", 64),
    ("odyssey/odyssey_mis_small.xml", "odyssey_mis_ref_64spp.npy", 64),
    # participating media (scenes/project/volume): path_vol_mats / path_vol_mis, homogeneous medium
    # in a BSDF-less sphere, isotropic and Henyey-Greenstein phase, ambient medium, mirror sphere
    ("vol/cbox_val_mats_small.xml", "vol_val_mats_ref_128spp.npy", 128),
    ("vol/cbox_val_mis_small.xml", "vol_val_mis_ref_128spp.npy", 128),
    ("vol/cbox_homog_aniso_small.xml", "vol_homog_aniso_ref_128spp.npy", 128),
    ("vol/cbox_homog_ambient_small.xml", "vol_homog_ambient_ref_128spp.npy", 128),
    ("vol/cbox_homog_small.xml", "vol_homog_ref_128spp.npy", 128),
])
def test_xml_scenes_match_reference(nb, ctx, xml, ref, spp):
    """Depth-1 and volume integrators on the reference's own scenes, loaded from the reference's
    XML by the C++ host loader, against reference CPU renders at equal spp."""
    s = nb.load_xml(golden("scenes/" + xml))
    want = np.load(golden(ref))
    imgs = []
    for seed in (1, 2):
        ctx.load(s)
        ctx.render_begin(nb.make_render_params(s, seed=seed))
        ctx.render_subframe(0, spp)
        imgs.append(ctx.film_download(normalize=True))
    a, b = imgs
    assert a.shape == want.shape and np.isfinite(a).all()
    assert ctx.counters()["samples"] == a.shape[0] * a.shape[1] * spp
    bias = (a.mean(axis=(0, 1)) - want.mean(axis=(0, 1))) / want.mean(axis=(0, 1))
    floor, err = rel_rmse(a, b), rel_rmse(a, want)
    print(xml, "bias", bias, "relRMSE", err, "noise floor", floor)
    assert np.all(np.abs(bias) <= 0.02), bias
    assert err <= 1.25 * floor + 1e-3, (err, floor)


@pytest.mark.parametrize("name,integrator,spp", [("cbox", "path_mis", 64), ("cbox_mats", "path_mats", 64), ("table", "path_mis", 16)])
def test_tile_strategy_equals_wavefront(nb, ctx, name, integrator, spp):
    """K_tile (registers, one launch) and the wavefront pipeline draw the same pcg32 stream per
    (pixel, sample) in the same order, so with one seed they trace the same paths: the images
    agree far below the Monte-Carlo noise (not bit-for-bit: the film sums in a different order
    and the two translation units may contract a*b+c differently in shading)."""
    scene = nb.read_nbs(golden(name + ".nbs"))
    size = (200, 150)
    a, ca = render(nb, ctx, scene, integrator, spp, size=size, seed=3, strategy=1)
    b, cb = render(nb, ctx, scene, integrator, spp, size=size, seed=3, strategy=2)
    c, _ = render(nb, ctx, scene, integrator, spp, size=size, seed=4, strategy=2)
    assert ca["samples"] == cb["samples"] == size[0] * size[1] * spp
    assert cb["kernel_launches"] == 1
    floor = rel_rmse(b, c)
    diff = rel_rmse(a, b)
    print(name, "tile vs wavefront relRMSE", diff, "noise floor", floor, "rays", ca["rays_closest"], cb["rays_closest"])
    assert diff <= 0.15 * floor        # a few paths in 10^4 branch differently (table: microfacet + dielectric)
    assert abs(ca["rays_closest"] - cb["rays_closest"]) <= 1e-3 * ca["rays_closest"]
    assert abs(ca["rays_any"] - cb["rays_any"]) <= 1e-3 * ca["rays_any"]


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


def test_cli_progressive_snapshots(nb, tmp_path):
    """Progressive output without a window (the reference shows the running image through GL, render.cpp:725-747): with
    snapshot=K nori_b200 writes the normalised image so far every K launches; each is the render of its sample prefix
    (per-(pixel, sample) streams), so a snapshot equals a fresh render of that many samples and the noise falls."""
    import json
    import os
    import subprocess
    cli = os.path.join(os.path.dirname(nb.LIB_PATH), "nori_b200")
    xml = golden("scenes/cbox/cbox_path_mis_small.xml")
    out = tmp_path / "prog.npy"
    r = subprocess.run([cli, xml, str(out), "spp=64", "launch=16", "snapshot=1", "seed=9"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 0, r.stderr
    stats = json.loads(r.stdout.strip().splitlines()[-1])
    assert stats["snapshots"] == 3 and stats["samples"] == 200 * 150 * 64
    snaps = [np.load(str(tmp_path / ("prog.%dspp.npy" % k))) for k in (16, 32, 48)]
    final = np.load(out)
    want = np.load(golden("cbox_ref_256spp.npy"))
    errs = [rel_rmse(s, want) for s in snaps] + [rel_rmse(final, want)]
    assert errs[0] > errs[1] > errs[3]
    fresh = tmp_path / "fresh.npy"
    r = subprocess.run([cli, xml, str(fresh), "spp=32", "seed=9"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 0, r.stderr
    a, b = np.load(fresh), snaps[1]
    assert np.abs(a - b).max() <= 1e-4 * max(1.0, np.abs(a).max())      # same samples, different order of fp32 additions


def test_cli_writes_png_like_save_to_ldr(nb, tmp_path):
    """out.png: Bitmap::saveToLDR (src/utils/bitmap.cpp:109-140) — sRGB transfer, x 255 + 0.5, clamp — decoded here by PIL and
    compared with the same render written as .npy."""
    import os
    import subprocess
    from PIL import Image
    cli = os.path.join(os.path.dirname(nb.LIB_PATH), "nori_b200")
    xml = golden("scenes/cbox/cbox_path_mis_small.xml")
    for name in ("a.png", "a.npy"):
        r = subprocess.run([cli, xml, str(tmp_path / name), "spp=16", "seed=5"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        assert r.returncode == 0, r.stderr
    png = np.asarray(Image.open(tmp_path / "a.png").convert("RGB")).astype(np.int32)
    lin = np.load(tmp_path / "a.npy").astype(np.float32)
    g = np.where(lin <= 0.0031308, 12.92 * lin, 1.055 * np.power(np.maximum(lin, 0), 1 / 2.4) - 0.055)
    want = np.clip(255.0 * g + 0.5, 0, 255).astype(np.uint8).astype(np.int32)
    assert png.shape == want.shape
    assert np.abs(png - want).max() <= 1              # the two runs sum their film in different orders: a byte may round the other way
    assert (png != want).mean() < 0.01


def test_cli_checkpoint_and_resume(nb, tmp_path):
    """The film's accumulators are the whole state of a render (they are linear, and sample s of pixel p has its own random
    stream): 24 samples saved with checkpoint=, the other 40 drawn later with resume= first=24, give the image of one
    64-sample render."""
    import json
    import os
    import subprocess
    cli = os.path.join(os.path.dirname(nb.LIB_PATH), "nori_b200")
    xml = golden("scenes/cbox/cbox_path_mis_small.xml")

    def run(out, *opts):
        r = subprocess.run([cli, xml, str(tmp_path / out), "seed=21"] + list(opts), stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        assert r.returncode == 0, r.stderr
        return json.loads(r.stdout.strip().splitlines()[-1])

    whole = run("whole.npy", "spp=64")
    part = run("part.npy", "spp=24", "checkpoint=" + str(tmp_path / "film.npy"))
    rest = run("resumed.npy", "spp=64", "first=24", "resume=" + str(tmp_path / "film.npy"), "checkpoint=" + str(tmp_path / "film2.npy"))
    assert whole["samples"] == 200 * 150 * 64 and part["samples"] == 200 * 150 * 24 and rest["samples"] == 200 * 150 * 40
    a, b = np.load(tmp_path / "whole.npy"), np.load(tmp_path / "resumed.npy")
    assert np.abs(a - b).max() <= 1e-4 * max(1.0, np.abs(a).max())          # same samples, summed in two parts
    film, film2 = np.load(tmp_path / "film.npy"), np.load(tmp_path / "film2.npy")
    assert film.shape == film2.shape == (150 + 4, 200 + 4, 4)                # ImageBlock layout: gaussian filter, border 2
    np.testing.assert_allclose(film2[2:-2, 2:-2, 3].sum() / film[2:-2, 2:-2, 3].sum(), 64 / 24, rtol=2e-3)
    assert rel_rmse(np.load(tmp_path / "part.npy"), np.load(golden("cbox_ref_256spp.npy"))) > rel_rmse(b, np.load(golden("cbox_ref_256spp.npy")))
