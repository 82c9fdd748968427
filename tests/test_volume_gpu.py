"""Heterogeneous media on the GPU: the 27-tap density sampler, delta-tracking
free paths and ratio-tracking transmittance of nb_volume.cuh against
  (1) the REFERENCE's own homogeneous-medium render, through a constant-density grid (the one pin
      the reference offers: heterogmedium.cpp as written cannot run);
  (2) oracle/ref_heterog.cpp — the corrected restatement on the reference's vendored NanoVDB
      (GridBuilder grid, nanovdb::Ray, SampleFromVoxels<.., 2>) — point by point and as renders.
Fixtures: tests/golden/make_volumes.py."""
import numpy as np
import pytest

from conftest import golden, rel_rmse

pytestmark = pytest.mark.gpu


def test_density_free_path_transmittance_vs_nanovdb_oracle(nb, ctx):
    s = nb.load_xml(golden("scenes/vol/cbox_heterog_fluid_small.xml"))
    rays = np.fromfile(golden("vol_fluid_probe_rays.bin"), "<f4").reshape(-1, 8)
    want = np.fromfile(golden("vol_fluid_probe_out.bin"), "<f4").reshape(-1, 4)     # 16384 draws per ray
    ctx.load(s)
    n = 16384
    got = ctx.medium_probe(0, rays, samples=n, seed=11)
    # NvdbVolume::getDensity: same 27 taps and the same polynomial; NanoVDB carries the weights in
    # double, the kernel in fp32 (values up to 1)
    print("density max abs diff", np.abs(got[:, 0] - want[:, 0]).max(), "nonzero points", int((want[:, 0] != 0).sum()))
    np.testing.assert_allclose(got[:, 0], want[:, 0], atol=2e-6, rtol=1e-5)
    assert (want[:, 0] != 0).sum() > 100
    # collision probability and mean transmittance: two independent Monte-Carlo means of n draws each
    p = want[:, 1]
    tol = 5.0 * np.sqrt(2.0 * np.maximum(p * (1 - p), 1e-4) / n) + 1e-4
    assert np.all(np.abs(got[:, 1] - p) <= tol), np.abs(got[:, 1] - p).max()
    assert np.all(np.abs(got[:, 3] - want[:, 3]) <= 5.0 * np.sqrt(2.0 * 0.3 / n) + 1e-4), np.abs(got[:, 3] - want[:, 3]).max()
    both = (p > 0.05)
    assert np.all(np.abs(got[both, 2] - want[both, 2]) <= 0.05 * want[both, 2] + 2e-3)
    # rays that miss the grid's box (slab test here): free path infinity, transmittance exactly 1
    lo = np.array([0.102, 0.03, -0.028]); hi = lo + np.array([21, 32, 21]) / 32.0
    o, d, maxt = rays[:, :3].astype(np.float64), rays[:, 3:6].astype(np.float64), rays[:, 7]
    with np.errstate(divide="ignore", invalid="ignore"):
        ta, tb = (lo - o) / d, (hi - o) / d
    t0 = np.maximum(np.minimum(ta, tb).max(axis=1), 1e-4); t1 = np.minimum(np.maximum(ta, tb).min(axis=1), maxt)
    miss = t0 > t1 + 1e-3
    assert miss.sum() > 10 and np.all(got[miss, 1] == 0) and np.all(got[miss, 3] == 1.0) and np.all(got[miss, 0] == 0)
    # delta tracking samples the transmittance: P(no collision before maxt) = E[ratio-tracking Tr]
    np.testing.assert_allclose(1.0 - got[:, 1], np.clip(got[:, 3], 0, 1), atol=0.03)


def test_nvdb_scene_vs_nanovdb_reading_the_same_file(nb, ctx):
    """cbox_heterog_nvdb.xml names a .nvdb written by the reference's NanoVDB: host/nvdb.cpp + the kernels here against
    nanovdb::io::readGrid + SampleFromVoxels<.., 2> + the restated tracking there (tests/golden/make_nvdb.py)."""
    s = nb.load_xml(golden("scenes/vol/cbox_heterog_nvdb.xml"))
    rays = np.fromfile(golden("nvdb_probe_rays.bin"), "<f4").reshape(-1, 8)
    want = np.fromfile(golden("nvdb_probe_out.bin"), "<f4").reshape(-1, 4)
    ctx.load(s)
    n = 16384
    got = ctx.medium_probe(0, rays, samples=n, seed=5)
    print("density max abs diff", np.abs(got[:, 0] - want[:, 0]).max(), "nonzero points", int((want[:, 0] != 0).sum()))
    np.testing.assert_allclose(got[:, 0], want[:, 0], atol=2e-6, rtol=1e-5)
    assert (want[:, 0] != 0).sum() > 100
    p = want[:, 1]
    tol = 5.0 * np.sqrt(2.0 * np.maximum(p * (1 - p), 1e-4) / n) + 1e-4
    assert p.max() > 0.9 and np.all(np.abs(got[:, 1] - p) <= tol), np.abs(got[:, 1] - p).max()
    assert np.all(np.abs(got[:, 3] - want[:, 3]) <= 5.0 * np.sqrt(2.0 * 0.3 / n) + 1e-4), np.abs(got[:, 3] - want[:, 3]).max()
    both = (p > 0.05)
    assert np.all(np.abs(got[both, 2] - want[both, 2]) <= 0.05 * want[both, 2] + 2e-3)


def test_brick_layout_is_the_dense_layout_bit_for_bit(nb, ctx, monkeypatch):
    """NB_MEDIUM_HETEROGENEOUS_BRICKS holds the same voxels behind a table: the same 27 taps in the same order, so densities and
    tracking results are the dense layout's bits, and a render is the dense layout's render."""
    xml = golden("scenes/vol/cbox_heterog_fluid_small.xml")
    rays = np.fromfile(golden("vol_fluid_probe_rays.bin"), "<f4").reshape(-1, 8)
    out = {}
    for layout in ("dense", "bricks"):
        monkeypatch.setenv("NB_VOLUME_LAYOUT", layout)
        s = nb.load_xml(xml)
        assert s["media"][0]["type"] == (nb.MEDIUM_HETEROGENEOUS if layout == "dense" else nb.MEDIUM_HETEROGENEOUS_BRICKS)
        ctx.load(s)
        probe = ctx.medium_probe(0, rays, samples=512, seed=3)
        ctx.render_begin(nb.make_render_params(s, seed=4))
        ctx.render_subframe(0, 8)
        out[layout] = (probe, ctx.film_download(normalize=True))
    assert (out["dense"][0][:, 0] != 0).sum() > 100
    assert np.array_equal(out["dense"][0], out["bricks"][0])
    # the film sums splats with atomics in whatever order the tiles finish: equal up to that reassociation
    np.testing.assert_allclose(out["dense"][1], out["bricks"][1], rtol=2e-5, atol=1e-6)


def test_sparse_scene_vs_nanovdb_reading_the_same_file(nb, ctx):
    """cbox_sparse_small.xml: a sparse .nvdb (written here, leaves only where there is smoke) -> leaves on the host -> bricks
    on the device, against nanovdb::io::readGrid + SampleFromVoxels<.., 2> on the same file (tests/golden/make_nvdb.py)."""
    s = nb.load_xml(golden("scenes/vol/cbox_sparse_small.xml"))
    assert s["media"][0]["type"] == nb.MEDIUM_HETEROGENEOUS_BRICKS
    rays = np.fromfile(golden("sparse_probe_rays.bin"), "<f4").reshape(-1, 8)
    want = np.fromfile(golden("sparse_probe_out.bin"), "<f4").reshape(-1, 4)
    ctx.load(s)
    n = 16384
    got = ctx.medium_probe(0, rays, samples=n, seed=9)
    print("density max abs diff", np.abs(got[:, 0] - want[:, 0]).max(), "nonzero points", int((want[:, 0] != 0).sum()))
    np.testing.assert_allclose(got[:, 0], want[:, 0], atol=2e-6, rtol=1e-5)
    assert (want[:, 0] != 0).sum() > 100
    p = want[:, 1]
    tol = 5.0 * np.sqrt(2.0 * np.maximum(p * (1 - p), 1e-4) / n) + 1e-4
    assert p.max() > 0.5 and np.all(np.abs(got[:, 1] - p) <= tol), np.abs(got[:, 1] - p).max()
    assert np.all(np.abs(got[:, 3] - want[:, 3]) <= 5.0 * np.sqrt(2.0 * 0.3 / n) + 1e-4), np.abs(got[:, 3] - want[:, 3]).max()
    both = (p > 0.05)
    assert np.all(np.abs(got[both, 2] - want[both, 2]) <= 0.05 * want[both, 2] + 2e-3)
    ctx.render_begin(nb.make_render_params(s, seed=1))                              # and it renders: finite, smoke visible
    ctx.render_subframe(0, 16)
    img = ctx.film_download(normalize=True)
    assert np.isfinite(img).all() and img.mean() > 0.01


def render_pair(nb, ctx, xml, spp):
    s = nb.load_xml(golden("scenes/" + xml))
    imgs = []
    for seed in (1, 2):
        ctx.load(s)
        ctx.render_begin(nb.make_render_params(s, seed=seed))
        ctx.render_subframe(0, spp)
        imgs.append(ctx.film_download(normalize=True))
    assert ctx.counters()["samples"] == imgs[0].shape[0] * imgs[0].shape[1] * spp
    return imgs


def test_constant_grid_equals_reference_homogeneous_render(nb, ctx):
    """A grid of density 1 around the medium sphere is the homogeneous medium of cbox_homog.xml:
    the image must match the reference's own CPU render of that scene (path_vol_mis)."""
    a, b = render_pair(nb, ctx, "vol/cbox_heterog_const_small.xml", 128)
    want = np.load(golden("vol_homog_ref_128spp.npy"))
    assert a.shape == want.shape and np.isfinite(a).all()
    bias = (a.mean(axis=(0, 1)) - want.mean(axis=(0, 1))) / want.mean(axis=(0, 1))
    floor, err = rel_rmse(a, b), rel_rmse(a, want)
    print("const grid vs reference homog: bias", bias, "relRMSE", err, "noise floor", floor)
    assert np.all(np.abs(bias) <= 0.02), bias
    assert err <= 1.25 * floor + 1e-3, (err, floor)
    # and the medium region itself (the sphere's pixels), where a wrong extinction would show
    h = nb.load_xml(golden("scenes/vol/cbox_homog_small.xml"))
    ctx.load(h); ctx.render_begin(nb.make_render_params(h, seed=3)); ctx.render_subframe(0, 128)
    g = ctx.film_download(normalize=True)
    ys, xs = slice(75, 125), slice(30, 75)
    r = a[ys, xs].mean() / g[ys, xs].mean()
    print("sphere region heterog/homog", r)
    assert abs(r - 1) < 0.03


@pytest.mark.parametrize("xml,ref", [("vol/cbox_heterog_fluid_small.xml", "vol_heterog_fluid_ref_128spp.npy"),
                                     ("vol/cbox_heterog_fluid_mis_small.xml", "vol_heterog_fluid_mis_ref_128spp.npy")])
def test_fluid_matches_restated_oracle(nb, ctx, xml, ref):
    a, b = render_pair(nb, ctx, xml, 128)
    want = np.load(golden(ref))
    bias = (a.mean(axis=(0, 1)) - want.mean(axis=(0, 1))) / want.mean(axis=(0, 1))
    floor, err = rel_rmse(a, b), rel_rmse(a, want)
    print(xml, "bias", bias, "relRMSE", err, "noise floor", floor)
    assert np.all(np.abs(bias) <= 0.02), bias
    assert err <= 1.25 * floor + 1e-3, (err, floor)


def test_reference_heterog_scene_loads_and_renders(nb, ctx):
    """scenes/project/volume/cbox_heterog.xml unchanged (reads the .vdb): its cube only overlaps empty
    voxels of the grid, so it renders as the same box without a medium."""
    s = nb.load_xml(golden("scenes/vol/cbox_heterog.xml"))
    cam = s["camera"].copy(); cam["width"], cam["height"] = 160, 120; cam["inv_output_size"] = [1 / 160, 1 / 120]
    s["camera"] = cam
    ctx.load(s); ctx.render_begin(nb.make_render_params(s, seed=1)); ctx.render_subframe(0, 64)
    a = ctx.film_download(normalize=True)
    e = dict(s); e["shapes"] = s["shapes"].copy(); e["shapes"]["medium"] = -1
    e["media"] = s["media"][:0]; e.pop("density", None)
    ctx.load(e); ctx.render_begin(nb.make_render_params(e, seed=1)); ctx.render_subframe(0, 64)
    b = ctx.film_download(normalize=True)
    assert np.isfinite(a).all() and abs(a.mean() / b.mean() - 1) < 0.02


def test_linear_density_ramp_closed_form(nb, ctx):
    """The one closed-form pin beyond the constant grid (the reference holds no golden for its heterogeneous medium): a
    density that is linear in x. The 27-tap triquadratic sampler reproduces a linear field exactly, the optical depth of
    a ray segment is then (density at its midpoint) x length, so transmittance = exp(-sigma_t * scale * depth) and
    P(free path < maxt) = 1 - transmittance — checked against the device's ratio- and delta-tracking means."""
    s = nb.load_xml(golden("scenes/vol/cbox_heterog_const_small.xml"))
    m = s["media"][0]
    assert m["type"] == nb.MEDIUM_HETEROGENEOUS
    nx, ny, nz = (int(v) for v in m["grid_dims"])
    first = int(m["first_voxel"])
    a, b = 0.2, 0.06                                     # density = a + b * (array x coordinate): 0.32 .. 1.28 inside the box
    ramp = (a + b * np.arange(nx, dtype=np.float32))[:, None, None] * np.ones((1, ny, nz), np.float32)
    t = dict(s)
    dens = s["density"].copy()
    dens[first:first + nx * ny * nz] = ramp.ravel()
    t["density"] = dens
    med = s["media"].copy()
    B = 2                                                # NB_GRID_BORDER
    med["max_density"][0] = a + b * (nx - B)             # the largest value the box can show
    t["media"] = med
    ctx.load(t)
    W = m["world_to_grid"].reshape(3, 4).astype(np.float64)
    rng = np.random.default_rng(5)
    n = 192
    Winv = np.linalg.inv(W[:, :3])
    og = rng.uniform([B + 0.5] * 3, [nx - B - 0.5, ny - B - 0.5, nz - B - 0.5], (n, 3))      # start inside the box
    o = (og - W[:, 3]) @ Winv.T
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    maxt = rng.uniform(0.05, 0.6, n)
    rays = np.concatenate([o, d, np.full((n, 1), 1e-4), maxt[:, None]], axis=1).astype("<f4")
    draws = 1 << 16
    got = ctx.medium_probe(0, rays, samples=draws, seed=3)
    # closed form in float64
    o64, d64 = rays[:, :3].astype(np.float64), rays[:, 3:6].astype(np.float64)
    g0, gd = o64 @ W[:, :3].T + W[:, 3], d64 @ W[:, :3].T
    lo, hi = np.array([B, B, B], float), np.array([nx - B, ny - B, nz - B], float)
    with np.errstate(divide="ignore", invalid="ignore"):
        ta, tb = (lo - g0) / gd, (hi - g0) / gd
    t0 = np.maximum(np.minimum(ta, tb).max(axis=1), 1e-4)
    t1 = np.minimum(np.maximum(ta, tb).min(axis=1), rays[:, 7].astype(np.float64))
    assert np.all(t1 > t0)
    xm = g0[:, 0] + gd[:, 0] * 0.5 * (t0 + t1)
    depth = (a + b * xm) * (t1 - t0)
    sigma = float(max(m["mu_t"])) * float(m["density_scale"])
    T = np.exp(-sigma * depth)
    np.testing.assert_allclose(got[:, 0], a + b * g0[:, 0], rtol=2e-5, atol=2e-6)          # the sampler is exact on a linear field
    se = 5.0 * np.sqrt(np.maximum(T * (1 - T), 1e-4) / draws) + 2e-4
    print("ramp: transmittance range %.3f..%.3f, max |ratio-tracking mean - exp(-tau)| %.2e, max |P(collision) - (1 - exp(-tau))| %.2e" % (
        T.min(), T.max(), np.abs(got[:, 3] - T).max(), np.abs(got[:, 1] - (1 - T)).max()))
    assert T.min() < 0.9 and T.max() > 0.95
    assert np.all(np.abs(got[:, 3] - T) <= se)           # ratio tracking: unbiased estimate of exp(-tau), variance below Bernoulli's
    assert np.all(np.abs(got[:, 1] - (1 - T)) <= se)     # delta tracking: collision before maxt with probability 1 - exp(-tau)


def test_a_1024_cubed_sparse_grid_on_the_device(nb, ctx, tmp_path):
    """Three 8^3 blocks in a 1024^3 box (4.3 GB dense): 8.6 MB of table + a few bricks on the device. At a voxel's own position
    the triquadratic stencil returns the voxel (weights 0, 1, 0 per axis), far from the blocks it returns the background."""
    import shutil
    rng = np.random.default_rng(3)
    org = np.array([[0, 0, 0], [504, 512, 520], [1016, 1016, 1016]], "<i4")
    vals = (rng.random((3, 8, 8, 8), dtype=np.float32) + 0.5).astype("<f4")
    M = np.array([[0.7 / 1024, 0, 0, 0.08], [0, 1.04 / 1024, 0, 0.02], [0, 0, 0.7 / 1024, -0.05]])
    shutil.copytree(golden("scenes/vol/meshes"), tmp_path / "meshes")
    nb.nvdb_write_leaves(str(tmp_path / "big.nvdb"), org, vals, M)
    (tmp_path / "scene.xml").write_text(open(golden("scenes/vol/cbox_sparse_small.xml")).read().replace("sparse_smoke.nvdb", "big.nvdb"))
    s = nb.load_xml(str(tmp_path / "scene.xml"))
    assert s["media"][0]["type"] == nb.MEDIUM_HETEROGENEOUS_BRICKS and list(s["media"][0]["grid_dims"]) == [1028] * 3 and s["density"].size < 3 * 10 ** 6
    ctx.load(s)
    pts, want = [], []
    for o, v in zip(org, vals):
        for ijk in ((1, 2, 3), (4, 4, 4), (6, 1, 5)):
            pts.append(M[:, :3] @ (o + np.array(ijk, np.float64)) + M[:, 3]); want.append(v[ijk])
    pts.append(M[:, :3] @ np.array([300.0, 300.0, 300.0]) + M[:, 3]); want.append(0.0)
    rays = np.concatenate([np.array(pts), np.tile([0.0, 0.0, 1.0], (len(pts), 1)), np.full((len(pts), 1), 1e-4), np.full((len(pts), 1), 0.01)], axis=1).astype("<f4")
    got = ctx.medium_probe(0, rays, samples=16, seed=1)
    # the voxel positions go through fp32 world coordinates and the fp32 world -> grid map: at 1024 voxels across 0.7 units a
    # position is good to ~1e-4 voxels, the stencil's neighbours (values 0.5 .. 1.5) weigh in with that much
    np.testing.assert_allclose(got[:, 0], np.array(want, np.float32), atol=2e-3)
    ctx.render_begin(nb.make_render_params(s, seed=1))
    ctx.render_subframe(0, 8)
    assert np.isfinite(ctx.film_download(normalize=True)).all()
