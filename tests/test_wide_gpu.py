"""The production traversal (nb_internal.h NB_WIDE_*: 4-wide tree, near-first, exact-t tie rule) against the reference's
answers. It is not a bit-parity mode by construction — its box test rounds differently from bbox.h:336-363 — so the
gate is SURVEY 7.1 step 6's: the same (shape, primitive) on >= 99.9999 % of the rays of the oracle's ray sets, and on
those rays bit-identical t, u, v (the primitive tests ARE the reference's arithmetic); images within the usual gates."""
import numpy as np
import pytest

from conftest import golden, have_oracle_binary, rel_rmse, run_oracle

pytestmark = pytest.mark.gpu


def parity(nb, ctx, scene, rays, want, flat=-1):
    ctx.set_option("flat_scan", flat)
    ctx.set_option("trace_wide", 1)
    ctx.load(scene)
    assert ctx.bvh_stats()["nodes"] > 0
    got = ctx.trace_rays(rays, any_hit=True)
    same = (got["shape"] == want["shape"]) & (got["prim"] == want["prim"])
    hit = same & (want["shape"] >= 0)
    for k in ("t", "u", "v"):
        assert got[k][hit].tobytes() == want[k][hit].tobytes(), k
    shadow_same = got["any_hit"] == want["any_hit"]
    return len(rays), int((~same).sum()), int((~shadow_same).sum())


@pytest.mark.parametrize("name,flat", [("table", -1), ("cbox", 0)])       # cbox: flat scan off, so its 3-node tree is walked
def test_wide_hits_on_golden_rays(nb, ctx, name, flat):
    scene = nb.read_nbs(golden(name + ".nbs"))
    rays = np.fromfile(golden(name + "_rays.bin"), dtype="<f4").reshape(-1, 8)
    want = np.fromfile(golden(name + "_hits.bin"), dtype=nb.hit_dtype)
    n, bad, bad_any = parity(nb, ctx, scene, rays, want, flat)
    print(name, "rays", n, "closest-hit mismatches", bad, "shadow mismatches", bad_any)
    assert bad == 0 and bad_any == 0                 # 30 k rays: one mismatch would already be 3e-5


@pytest.mark.skipif(not have_oracle_binary(), reason="oracle/_ref/nori_ref not present")
@pytest.mark.parametrize("name,xml,n,flat", [("cbox", "scenes/cbox/cbox_path_mis.xml", 200000, 0), ("table", "scenes/table/table_path_mis.xml", 150000, -1)])
def test_wide_hit_parity_on_live_oracle_rays(nb, ctx, name, xml, n, flat, tmp_path):
    rays_f, hits_f = str(tmp_path / "rays.bin"), str(tmp_path / "hits.bin")
    run_oracle("genrays", golden(xml), rays_f, str(n), "31337")
    run_oracle("hits", golden(xml), rays_f, hits_f)
    rays = np.fromfile(rays_f, dtype="<f4").reshape(-1, 8)
    full = np.fromfile(hits_f, dtype=np.uint8).reshape(-1, 128)
    want = np.zeros(len(full), dtype=nb.hit_dtype)
    want["shape"] = full[:, 0:4].copy().view("<i4")[:, 0]
    want["prim"] = full[:, 4:8].copy().view("<i4")[:, 0]
    tuv = full[:, 8:20].copy().view("<f4")
    want["t"], want["u"], want["v"] = tuv[:, 0], tuv[:, 1], tuv[:, 2]
    want["any_hit"] = full[:, 28:32].copy().view("<u4")[:, 0]
    total, bad, bad_any = parity(nb, ctx, nb.read_nbs(golden(name + ".nbs")), rays, want, flat)
    print(name, "rays", total, "closest-hit mismatches", bad, "(%.2e)" % (bad / total), "shadow mismatches", bad_any)
    assert total > 500000
    assert bad <= max(1, int(1e-6 * total)) and bad_any <= max(1, int(1e-6 * total))


def test_wide_and_reference_order_render_the_same_image(nb, ctx):
    """Same seed, same per-(pixel, sample) streams: the two traversal modes trace the same paths unless a hit differs, so
    the images agree far below the Monte-Carlo noise; and each matches the reference render."""
    scene = nb.read_nbs(golden("table.nbs"))
    want = np.load(golden("table_ref_64spp.npy"))
    h, w, _ = want.shape
    cam = scene["camera"].copy(); cam["width"], cam["height"] = w, h; cam["inv_output_size"] = [1.0 / w, 1.0 / h]
    scene["camera"] = cam
    imgs = {}
    for wide, seed in ((1, 1), (0, 1), (1, 2)):
        ctx.set_option("wide", wide)
        ctx.load(scene)
        ctx.render_begin(nb.make_render_params(scene, "path_mis", seed=seed))
        ctx.render_subframe(0, 64)
        imgs[(wide, seed)] = ctx.film_download(normalize=True)
        c = ctx.counters()
        assert c["samples"] == w * h * 64
    floor = rel_rmse(imgs[(1, 1)], imgs[(1, 2)])
    diff = rel_rmse(imgs[(1, 1)], imgs[(0, 1)])
    print("wide vs reference order, same seed: relRMSE", diff, "noise floor", floor)
    assert diff <= 0.15 * floor
    for key in ((1, 1), (0, 1)):
        a = imgs[key]
        bias = (a.mean(axis=(0, 1)) - want.mean(axis=(0, 1))) / want.mean(axis=(0, 1))
        assert np.all(np.abs(bias) <= 0.01) and rel_rmse(a, want) <= 1.25 * floor + 1e-3

