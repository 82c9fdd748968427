"""The adaptive sampler's feedback loop on the device (nb_render_adaptive_round) against the reference's
own AdaptiveSampler + block loop (src/samplers/adaptive.cpp, render.cpp:319-335) on cbox_adaptive_small.xml, 8 rounds.
The loop is a random stopping process, so what is compared is its law: how many passes blocks render, and the image.
Fixtures: tests/golden/make_adaptive.py."""
import numpy as np
import pytest

from conftest import golden, rel_rmse

pytestmark = pytest.mark.gpu
ROUNDS = 8


def run(nb, ctx, s, seed):
    assert list(s["adaptive"]) == [4, 2]
    ctx.load(s)
    ctx.render_begin(nb.make_render_params(s, seed=seed))
    infos = [ctx.render_adaptive_round(int(s["adaptive"][0]), int(s["adaptive"][1])) for _ in range(ROUNDS)]
    h, w = int(s["camera"]["height"][0]), int(s["camera"]["width"][0])
    n_blocks = ((w + 3) // 4) * ((h + 3) // 4)
    return infos, ctx.adaptive_block_passes(n_blocks), ctx.film_download(normalize=True), ctx.counters()


def test_adaptive_rounds_follow_the_reference_sampler(nb, ctx):
    s = nb.load_xml(golden("scenes/cbox/cbox_adaptive_small.xml"))
    infos, passes, img, counters = run(nb, ctx, s, seed=1)
    _, passes2, img2, _ = run(nb, ctx, s, seed=2)
    ref_passes = np.load(golden("adaptive_cbox_passes.npy")).astype(np.int64)
    ref = np.load(golden("adaptive_cbox_ref.npy"))
    # bookkeeping: every pass of a block draws one sample per pixel of the block, the counters agree with the rounds' reports
    assert all(i["blocks"] == 1900 for i in infos) and passes.sum() == sum(i["block_passes"] for i in infos)
    area = np.minimum(4, 150 - 4 * (np.arange(1900) // 50)) * 4           # 50 blocks per row; the last row of blocks is 2 pixels high
    assert sum(i["samples"] for i in infos) == counters["samples"] == int((passes * area).sum())
    # a round: initialUniform + 1 uniform passes for every block (the reference's post-increment), then the first placed
    # pass always goes on (its change is measured against 10000): no block stops before 4
    assert passes.min() >= 4 * ROUNDS
    print([(i["passes"], i["block_passes"], i["samples"]) for i in infos])
    # black blocks: their variance map is zero at every decision and they go round to the cap of 1000 decisions, 1001 passes
    # per round, in the reference and here (the same blocks: the image's black margin)
    cap_ref, cap_dev = int((ref_passes >= 1000).sum()), int((passes >= 1000).sum())
    print("blocks that reached the cap at least once: reference", cap_ref, "device", cap_dev)
    assert abs(cap_dev - cap_ref) <= 0.1 * cap_ref and int((passes == 1001 * ROUNDS).sum()) >= 0.8 * cap_ref
    assert all(i["passes"] == 1001 for i in infos)
    # everything else: the law of the passes per block over 8 rounds — blocks whose first placed decision sees a constant
    # map (4 passes that round) and blocks that go on while the change of their variance map shrinks (5 or more)
    a, b = ref_passes[ref_passes < 200], passes[passes < 200]
    qa, qb = np.percentile(a, [5, 10, 25, 50, 75, 90, 95]), np.percentile(b, [5, 10, 25, 50, 75, 90, 95])
    b2 = passes2[passes2 < 200]
    print("passes per block over", ROUNDS, "rounds: reference mean", a.mean(), "quantiles", qa, "| device mean", b.mean(), "quantiles", qb,
          "| second seed mean", b2.mean())
    assert abs(b.mean() - a.mean()) <= 0.03 * a.mean() and abs(b2.mean() - a.mean()) <= 0.03 * a.mean()
    assert np.all(np.abs(qa - qb) <= 2.0)
    lo_ref, lo_dev = float((a < 5 * ROUNDS).mean()), float((b < 5 * ROUNDS).mean())      # blocks with at least one 4-pass round
    print("share of blocks with a 4-pass round: reference", lo_ref, "device", lo_dev)
    assert abs(lo_dev - lo_ref) <= 0.05
    # the image: same expectation as the reference's, at the noise of ~45 samples per pixel
    floor, err = rel_rmse(img, img2), rel_rmse(img, ref)
    bias = (img.mean(axis=(0, 1)) - ref.mean(axis=(0, 1))) / ref.mean(axis=(0, 1))
    print("adaptive image vs reference: bias", bias, "relRMSE", err, "two-seed floor", floor)
    assert np.isfinite(img).all() and np.all(np.abs(bias) <= 0.03)
    assert err <= 1.3 * floor + 1e-3


def test_adaptive_rounds_mix_with_plain_subframes(nb, ctx):
    """The film is one film: uniform subframes before and after a round accumulate into it, and a plain subframe after a
    round does not touch the blocks' own images."""
    s = nb.load_xml(golden("scenes/cbox/cbox_adaptive_small.xml"))
    ctx.load(s)
    ctx.render_begin(nb.make_render_params(s, seed=3))
    ctx.render_subframe(0, 4)
    info = ctx.render_adaptive_round(4, 2, round_cap=6)      # black blocks stop at the 6th round instead of the 1000th
    assert info["passes"] <= 7 + 40
    ctx.render_subframe(4, 4)
    c = ctx.counters()
    assert c["samples"] == 200 * 150 * 8 + info["samples"] and info["samples"] >= 200 * 150 * 3
    img = ctx.film_download(normalize=True)
    assert np.isfinite(img).all() and img.mean() > 0.1
    with pytest.raises(nb.NbError):
        ctx.render_adaptive_round(16, 2)          # block sizes 2..8


@pytest.mark.parametrize("bs", [2, 8])
def test_adaptive_rounds_other_block_sizes(nb, ctx, bs):
    """block_size is a parameter of the entry (the reference compiles 4 in): 2 and 8 keep the books and the image."""
    s = nb.load_xml(golden("scenes/cbox/cbox_adaptive_small.xml"))
    ctx.load(s)
    ctx.render_begin(nb.make_render_params(s, seed=6))
    infos = [ctx.render_adaptive_round(bs, 2, round_cap=12) for _ in range(3)]
    n_blocks = ((200 + bs - 1) // bs) * ((150 + bs - 1) // bs)
    passes = ctx.adaptive_block_passes(n_blocks)
    by, bx = np.divmod(np.arange(n_blocks), (200 + bs - 1) // bs)
    area = np.minimum(bs, 200 - bs * bx) * np.minimum(bs, 150 - bs * by)
    assert all(i["blocks"] == n_blocks for i in infos)
    assert sum(i["samples"] for i in infos) == ctx.counters()["samples"] == int((passes * area).sum())
    assert passes.min() >= 4 * 3
    img = ctx.film_download(normalize=True)
    want = np.load(golden("cbox_ref_256spp.npy"))
    bias = (img.mean(axis=(0, 1)) - want.mean(axis=(0, 1))) / want.mean(axis=(0, 1))
    assert np.isfinite(img).all() and np.all(np.abs(bias) <= 0.03), bias


def test_adaptive_rounds_on_a_tree_scene(nb, ctx):
    """The sample list is consumed by the per-lane walk kernels too (table: 22 766 triangles, production tree)."""
    s = dict(nb.load_xml(golden("scenes/table/table_path_mis.xml")))
    cam = s["camera"].copy(); cam["width"], cam["height"] = 200, 150; cam["inv_output_size"] = [1 / 200, 1 / 150]
    s["camera"] = cam
    ctx.load(s)
    ctx.render_begin(nb.make_render_params(s, seed=2))
    infos = [ctx.render_adaptive_round(4, 2, round_cap=10) for _ in range(4)]
    c = ctx.counters()
    assert sum(i["samples"] for i in infos) == c["samples"] and c["invalid_samples"] == 0
    a = ctx.film_download(normalize=True)
    ctx.render_begin(nb.make_render_params(s, seed=3))
    ctx.render_subframe(0, max(1, int(round(c["samples"] / (200 * 150)))))
    b = ctx.film_download(normalize=True)
    bias = (a.mean(axis=(0, 1)) - b.mean(axis=(0, 1))) / b.mean(axis=(0, 1))
    print("table adaptive vs uniform at the same sample count: bias", bias, "relRMSE", rel_rmse(a, b))
    assert np.isfinite(a).all() and np.all(np.abs(bias) <= 0.03)
