#!/usr/bin/env python
"""C5: the 1 453 066-triangle multi-mesh scene at 3840 x 2160, sharded by sample
slices over the GPUs of one box, one NCCL reduce of the film accumulators to rank 0.

    python tools/c5_scale.py [--spp 64] [--width 3840 --height 2160] [--filter gaussian|box] [--device-bvh]
    python -m torch.distributed.run --nproc-per-node N ... tools/c5_scale.py ...

STRONG scaling: --spp is the total per pixel, rank g renders samples s = g (mod N). Prints one
JSON line on rank 0 (device-timed, max over ranks; render and reduce also separately) and saves a
4x4-box-downsampled image to --out so runs at different N can be compared: every sample draws
the same pcg32 stream at any N, so the images agree up to fp32 summation order."""
import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--spp", type=int, default=64)
    ap.add_argument("--width", type=int, default=3840)
    ap.add_argument("--height", type=int, default=2160)
    ap.add_argument("--filter", default="gaussian")
    ap.add_argument("--device-bvh", action="store_true")
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    import norib200 as nb
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    work = "/tmp/c5_%dx%d_%s" % (args.width, args.height, args.filter)
    nbs = os.path.join(work, "scene.nbs")
    prep = {}
    if rank == 0 and not os.path.exists(nbs):
        subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_c5_scene.py"), work, "width=%d" % args.width,
                        "height=%d" % args.height, "spp=%d" % args.spp, "filter=%s" % args.filter], check=True, stdout=subprocess.DEVNULL)
        t0 = time.perf_counter()
        xml = [f for f in os.listdir(work) if f.endswith(".xml")][0]
        nb.load_xml(os.path.join(work, xml), nbs_path=nbs + ".tmp")      # XML + 66 OBJ loads + SAH BVH, C++ host loader
        os.replace(nbs + ".tmp", nbs)
        prep["host_load_flatten_bvh_s"] = round(time.perf_counter() - t0, 2)
    if world > 1:
        dist.barrier()
    scene = nb.read_nbs(nbs)
    ctx = nb.Context(local)
    t0 = time.perf_counter()
    ctx.load(scene, device_bvh=args.device_bvh)
    prep["upload_s"] = round(time.perf_counter() - t0, 3)
    prep["bvh"] = ctx.bvh_stats()
    ctx.render_begin(nb.make_render_params(scene, "path_mis", seed=1))
    film = nb.film_tensor(ctx, local)

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ms_total = ms_render = ms_reduce = 0.0
    samples = rays = 0
    for it in range(args.steps + 1):                    # first pass is the warm-up
        ctx.film_clear()
        sync()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        ctx.render_subframe(0, args.spp, world, rank)
        ctx.sync()
        e[1].record()
        nb.reduce_film(film, dst=0)
        e[2].record(); e[2].synchronize()
        if it == 0:
            continue
        ms_render += e[0].elapsed_time(e[1]); ms_reduce += e[1].elapsed_time(e[2]); ms_total += e[0].elapsed_time(e[2])
        c = ctx.counters()
        samples += c["samples"]; rays += c["rays_closest"] + c["rays_any"]
    t = torch.tensor([ms_total, ms_render, ms_reduce], dtype=torch.float64, device="cuda")
    tot = torch.tensor([float(samples), float(rays)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX); dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    if rank == 0:
        rows, cols, _ = ctx.film_shape()
        img = ctx.film_download(normalize=True)
        out = {"config": "C5 %d copies camelhead, %d triangles, %dx%d, path_mis, %s filter" % (64, int(scene["faces"].size // 3), args.width, args.height, args.filter),
               "n_gpus": world, "spp_total": args.spp, "steps": args.steps, "ms_per_step": float(t[0]) / args.steps,
               "render_ms": float(t[1]) / args.steps, "reduce_ms": float(t[2]) / args.steps,
               "samples_per_s": float(tot[0]) / float(t[0]) * 1e3, "rays_per_s": float(tot[1]) / float(t[0]) * 1e3,
               "film_bytes": rows * cols * 16, "scaling": "strong", "finite": bool(np.isfinite(img).all()), "mean": float(img.mean())}
        out.update(prep)
        if args.out:
            h4, w4 = img.shape[0] // 4 * 4, img.shape[1] // 4 * 4
            small = img[:h4, :w4].reshape(h4 // 4, 4, w4 // 4, 4, 3).mean(axis=(1, 3))
            np.save(args.out, small.astype(np.float32))
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
