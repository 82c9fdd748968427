"""Emitter::sample / pdf / eval of every emitter plugin with a kernel, record by record against the
reference's own plugins (tests/golden/make_emitter_probes.py -> `nori_ref emitter`): area lights on
a mesh (cbox light, sphere mesh), on analytic spheres, the point light, the PNG environment map
(luminance CDF over 1500x750 texels, spherical texture rotation) and the constant one. Then the
environment-map and texture scenes as images against reference CPU renders."""
import numpy as np
import pytest

from conftest import golden, rel_rmse

pytestmark = pytest.mark.gpu

CASES = [("cbox_area", "scenes/cbox/cbox_path_mis.xml", 0), ("sphere_area", "scenes/sphere/sphere_ems.xml", 0),
         ("sphere2_area", "scenes/sphere/sphere2_ems.xml", 0), ("sphere_mesh_area", "scenes/sphere/sphere_mesh_ems.xml", 0),
         ("point", "scenes/sphere/point_ems.xml", 0), ("envmap_png", "scenes/envmap/envmap_sphere_small.xml", 0),
         ("envmap_const", "scenes/envmap/textures_small.xml", 0)]


@pytest.mark.parametrize("name,xml,index", CASES)
def test_emitter_sample_matches_reference(nb, ctx, name, xml, index):
    s = nb.load_xml(golden(xml))
    rec = np.fromfile(golden("emit_%s_in.bin" % name), "<f4").reshape(-1, 5)
    want = np.fromfile(golden("emit_%s_out.bin" % name), "<f4").reshape(-1, 12)
    ctx.load(s)
    got = ctx.emitter_probe(index, rec)
    wi_err = np.abs(got[:, :3] - want[:, :3]).max(axis=1)
    scale = np.maximum(np.abs(want[:, 3:7]).max(axis=1), 1e-6)
    val_err = np.abs(got[:, 3:7] - want[:, 3:7]).max(axis=1) / scale
    maxt_err = np.abs(got[:, 7] - want[:, 7]) / np.maximum(np.abs(want[:, 7]), 1.0)
    eval_err = np.abs(got[:, 8:11] - want[:, 8:11]).max(axis=1) / np.maximum(np.abs(want[:, 8:11]).max(axis=1), 1e-6)
    bad = (wi_err > 2e-5) | (val_err > 5e-4) | (maxt_err > 1e-5) | (eval_err > 5e-4)
    print(name, "records", len(rec), "max wi err", wi_err.max(), "median value err", np.median(val_err), "outside tolerance", int(bad.sum()))
    # image-backed emitters: a direction within an ulp of a texel edge may read the neighbour
    # (glibc vs CUDA acos/atan2), which changes value and pdf together; nothing else may differ
    allowed = 0.01 if name == "envmap_png" else 0.0
    # area lights seen at grazing incidence: pdf = pdfArea * dist^2 / |n . wi| amplifies the last
    # bits of the dot product (sphere_ems: one record in 4096, pdf 9e3); those are left out
    grazing = (want[:, 6] > 1e3) & (name != "envmap_png")
    bad &= ~grazing
    assert grazing.mean() <= 0.002 and bad.mean() <= allowed, (int(bad.sum()), got[bad][:3], want[bad][:3])
    assert np.all(wi_err <= 2e-5)


@pytest.mark.parametrize("xml,ref,spp", [
    ("envmap/envmap_sphere_small.xml", "env_envmap_sphere_ref_128spp.npy", 128),                    # path_mis, PNG environment
    ("envmap/envmap_sphere_path_mats_small.xml", "env_envmap_sphere_path_mats_ref_128spp.npy", 128),
    ("envmap/envmap_sphere_direct_mis_small.xml", "env_envmap_sphere_direct_mis_ref_128spp.npy", 128),
    ("envmap/envmap_sphere_direct_ems_small.xml", "env_envmap_sphere_direct_ems_ref_128spp.npy", 128),
    ("envmap/envmap_sphere_direct_mats_small.xml", "env_envmap_sphere_direct_mats_ref_128spp.npy", 128),
    ("envmap/textures_small.xml", "env_textures_ref_128spp.npy", 128),     # constant environment, PNG + checkerboard albedo, mirror
])
@pytest.mark.parametrize("strategy", [2, 1])
def test_envmap_and_texture_scenes_match_reference(nb, ctx, xml, ref, spp, strategy):
    s = nb.load_xml(golden("scenes/" + xml))
    want = np.load(golden(ref))
    integ = int(s["render"]["integrator"][0])
    if strategy == 1 and integ not in (nb.INTEGRATORS["path_mis"], nb.INTEGRATORS["path_mats"]):
        pytest.skip("the wavefront pipeline renders path_mis / path_mats only")
    imgs = []
    for seed in (1, 2):
        ctx.set_option("strategy", strategy)
        ctx.load(s)
        ctx.render_begin(nb.make_render_params(s, seed=seed))
        ctx.render_subframe(0, spp)
        imgs.append(ctx.film_download(normalize=True))
    a, b = imgs
    assert a.shape == want.shape and np.isfinite(a).all()
    bias = (a.mean(axis=(0, 1)) - want.mean(axis=(0, 1))) / want.mean(axis=(0, 1))
    floor, err = rel_rmse(a, b), rel_rmse(a, want)
    print(xml, "strategy", strategy, "bias", bias, "relRMSE", err, "noise floor", floor)
    assert np.all(np.abs(bias) <= 0.02), bias
    assert err <= 1.25 * floor + 2e-3, (err, floor)
