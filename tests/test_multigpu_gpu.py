"""SURVEY 8(e) on real GPUs: G devices render disjoint sample slices, ONE NCCL sum inside the C ABI (nb_film_reduce)
gives the film of all samples. Per-(pixel, sample) pcg32 streams make the N-GPU sample set identical to the 1-GPU
one, so the reduced film must equal the single-GPU film up to the order of fp32 additions. Needs >= 2 GPUs."""
import json
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, golden

pytestmark = pytest.mark.gpu


def n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


needs2 = pytest.mark.skipif(n_gpus() < 2, reason="needs 2 GPUs (gpurun --gpus 2)")


def small_cbox(nb):
    scene = nb.read_nbs(golden("cbox.nbs"))
    cam = scene["camera"].copy()
    cam["width"], cam["height"] = 200, 150
    cam["inv_output_size"] = [1.0 / 200, 1.0 / 150]
    scene["camera"] = cam
    return scene


def single_gpu_film(nb, scene, spp, seed=7):
    with nb.Context(0) as c:
        c.load(scene)
        c.render_begin(nb.make_render_params(scene, "path_mis", seed=seed))
        c.render_subframe(0, spp)
        return c.film_download(normalize=False)


def close(a, b):
    # same samples, different fp32 summation order (atomics within a GPU, NCCL across): relative 1e-5 of the pixel sums
    scale = np.abs(b).max()
    return np.abs(a - b).max() <= 1e-5 * scale


@needs2
def test_one_process_two_devices_matches_one_device(nb, built_lib):
    scene, spp = small_cbox(nb), 64
    want = single_gpu_film(nb, scene, spp)
    ctxs = nb.create_multi([0, 1])
    try:
        assert [c.comm_info()["rank"] for c in ctxs] == [0, 1] and ctxs[0].comm_info()["n_ranks"] == 2
        for c in ctxs:
            c.load(scene)
            c.render_begin(nb.make_render_params(scene, "path_mis", seed=7))
        for first in (0, spp // 2):
            for g, c in enumerate(ctxs):           # asynchronous: both devices render at the same time
                c.render_subframe(first, spp // 2, 2, g)
            ctxs[0].film_reduce()                  # one call reduces the group
        got = ctxs[0].film_download(normalize=False)
        assert not ctxs[1].film_download(normalize=False).any()
        assert sum(c.counters()["samples"] for c in ctxs) == 200 * 150 * spp
    finally:
        for c in ctxs:
            c.close()
    assert close(got, want)


@needs2
def test_torchrun_two_ranks_match_one_device(nb, built_lib, tmp_path):
    scene, spp = small_cbox(nb), 64
    want = single_gpu_film(nb, scene, spp)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    out = str(tmp_path / "film.npy")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", str(port), os.path.join(ROOT, "tests", "helpers", "multigpu_rank.py"), out, str(spp)],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:]
    assert close(np.load(out), want)


@needs2
def test_cli_devices_list_matches_one_device(nb, built_lib, tmp_path):
    """nori_b200 devices=0,1: the C++ caller (B200State with a device list) — the reference's single render thread."""
    cli = os.path.join(ROOT, "optix-renderer_b200", "nori_b200")
    xml = golden("scenes/cbox/cbox_path_mis.xml")
    outs = []
    for devs in ("device=0", "devices=0,1"):
        out = str(tmp_path / (devs.replace("=", "_").replace(",", "_") + ".npy"))
        r = subprocess.run([cli, xml, out, "spp=8", "seed=3", devs], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        j = json.loads(r.stdout.strip().splitlines()[-1])
        assert j["samples"] == 800 * 600 * 8 and j["devices"] == (2 if "," in devs else 1)
        outs.append(np.load(out))
    assert np.abs(outs[0] - outs[1]).max() <= 1e-4 * np.abs(outs[0]).max()


def test_reduce_is_a_noop_on_one_rank_and_needs_a_communicator(nb, ctx):
    scene = small_cbox(nb)
    ctx.load(scene)
    ctx.render_begin(nb.make_render_params(scene, "path_mis", seed=7))
    ctx.render_subframe(0, 4)
    before = ctx.film_download(normalize=False)
    ctx.film_reduce()
    assert np.array_equal(ctx.film_download(normalize=False), before)
    assert ctx.comm_info() == {"rank": 0, "n_ranks": 1, "nccl_version": 0}
    assert len(nb.comm_unique_id()) == 128
