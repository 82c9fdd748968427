"""The CPU + GPU sample split (render.cpp:256-263, :462-574): in the reference's `Both` device mode its CPU
threads and its GPU thread each take samples of the same image until sampleCount is reached, and the two ImageBlocks are
summed. That loop lives in the caller; what the boundary owes it is (1) subframes over ANY range of sample indices and (2) the
film as accumulators in ImageBlock layout, so that the caller's `m_block.put(..)` of its own samples and of these is one
sum. Shown with the reference's own CPU renderer as the caller's CPU side: samples [0, 32) on the CPU, [32, 128) here."""
import numpy as np
import pytest

from conftest import golden, have_oracle_binary, rel_rmse, run_oracle

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(not have_oracle_binary(), reason="needs oracle/_ref/nori_ref next to the repo")
def test_cpu_and_gpu_samples_sum_to_one_image(nb, ctx, tmp_path):
    xml = golden("scenes/cbox/cbox_path_mis_small.xml")
    want = np.load(golden("cbox_ref_256spp.npy"))
    run_oracle("render", xml, str(tmp_path / "cpu.npy"), "spp=32", "film=" + str(tmp_path / "cpu_film.npy"))
    cpu = np.load(tmp_path / "cpu_film.npy").astype(np.float64)           # rows x cols x { sum w rgb, sum w }, border included
    s = nb.load_xml(xml)
    ctx.load(s)
    ctx.render_begin(nb.make_render_params(s, seed=11))
    ctx.render_subframe(32, 96)                                            # the GPU's share: sample indices 32 .. 127
    gpu = ctx.film_download(normalize=False).astype(np.float64)
    assert gpu.shape == cpu.shape
    both = cpu + gpu                                                       # ImageBlock::put(const ImageBlock &): a sum of accumulators
    b = (gpu.shape[0] - want.shape[0]) // 2
    inner = both[b:-b, b:-b]
    img = np.where(inner[..., 3:] != 0, inner[..., :3] / np.where(inner[..., 3:] != 0, inner[..., 3:], 1), 0)
    # the weights say who drew what: a quarter of the samples came from the CPU
    share = cpu[b:-b, b:-b, 3].sum() / inner[..., 3].sum()
    assert abs(share - 0.25) < 0.01
    ctx.render_begin(nb.make_render_params(s, seed=12)); ctx.render_subframe(0, 128)
    alone = ctx.film_download(normalize=True)
    bias = (img.mean(axis=(0, 1)) - want.mean(axis=(0, 1))) / want.mean(axis=(0, 1))
    print("hybrid 32 CPU + 96 GPU spp: bias", bias, "relRMSE", rel_rmse(img, want), "GPU alone at 128 spp", rel_rmse(alone, want))
    assert np.all(np.abs(bias) <= 0.01)
    assert rel_rmse(img, want) <= 1.15 * rel_rmse(alone, want) + 1e-3
