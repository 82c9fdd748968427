#!/usr/bin/env python
"""bench.py — headline benchmark: samples/s and rays/s on the Cornell box, path_mis.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--spp S]

Workload (BASELINE.json configs[0], the scene the north-star target is quoted on):
tests/golden/scenes/cbox/cbox_path_mis.xml, 800x600, 512 spp, gaussian r=2 — one "step" is one
full render of it. N>1 (torchrun, one rank per GPU): every rank renders its own sample slice
(sample s = rank + N*k, 512 spp each: weak scaling), scene replicated, film accumulators summed
on rank 0 with one NCCL reduce inside the timed region.

`value`   camera samples/s, scene already resident in HBM, device-timed (CUDA events), max over ranks.
`e2e`     the same through the C ABI with HOST buffers: scene upload + BVH import (H2D), render,
          film download (D2H) every step, wall-clocked around the calls.
`--impl reference` times the reference's own CPU renderer (oracle/_ref/nori_ref, all host
threads) on a bounded sample of the same workload (same scene and image size, fewer spp).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
SCENE_XML = os.path.join(GOLDEN, "scenes", "cbox", "cbox_path_mis.xml")
SCENE_NBS = os.path.join(GOLDEN, "cbox.nbs")
NORI_REF = os.path.join(ROOT, "oracle", "_ref", "nori_ref")
WORKLOAD = "cbox_path_mis 800x600 512spp (scenes/pa4/cbox/cbox_path_mis.xml)"


def run_reference(spp, threads=None):
    """One bounded run of the reference CPU renderer; returns its JSON line."""
    if not os.path.exists(NORI_REF):
        raise FileNotFoundError("oracle/_ref/nori_ref is missing: run `make -C oracle ref` where /root/reference is mounted")
    args = [NORI_REF, "render", SCENE_XML, "/tmp/nb_bench_ref.npy", "spp=%d" % spp]
    if threads:
        args.append("threads=%d" % threads)
    r = subprocess.run(args, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    if r.returncode != 0:
        raise RuntimeError("nori_ref failed: " + r.stderr[-1000:])
    return json.loads(r.stdout.strip().splitlines()[-1])


class ClockSampler:
    """nvidia-smi clocks line of /opt/skills/guides/B200_PROFILING.md, sampled during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            self.proc.terminate()
            self.t.join(timeout=2)

    def summary(self):
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def impl_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    spp = args.ref_spp
    cores = os.cpu_count() or 1
    for _ in range(args.warmup):
        run_reference(max(1, spp // 4))
    t0 = time.perf_counter()
    samples = rays = 0.0
    secs = 0.0
    for _ in range(args.steps):
        j = run_reference(spp)
        samples += j["samples"]; secs += j["seconds"]
        rays += j["rays_closest"] + j["rays_any"] - j["rays_dup"]
    wall = time.perf_counter() - t0
    v = samples / secs
    line = {"impl": "reference", "metric": "samples_per_s", "value": v, "unit": "samples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sample": "%d spp per step instead of 512 (throughput is spp-independent)" % spp},
            "rays_per_s": rays / secs, "wall_s": wall,
            "cpu_baseline": {"value": v, "unit": "samples/s", "cores": cores, "kind": "reference",
                             "sample": "same scene and 800x600 image, %d spp per step, sample loop only (render.cpp:281-360)" % spp},
            "e2e": {"value": v, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


_REAL_STDOUT = None


def emit(obj):
    sys.stdout.flush()
    if _REAL_STDOUT is not None:
        os.dup2(_REAL_STDOUT, 1)
    print(json.dumps(obj), flush=True)


def ncu_capture(path):
    """metric -> value (bytes for the dram counters) of a tools/ncu_summary.py file, or None."""
    if not os.path.exists(path):
        return None
    import csv
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    out = {}
    for row in csv.DictReader(open(path)):
        try:
            out[row["metric"]] = float(row["value"]) * scale.get(row["unit"], 1.0)
        except ValueError:
            pass
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--spp", type=int, default=512)
    ap.add_argument("--ref-spp", type=int, default=16)
    ap.add_argument("--pool", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--strategy", default="auto", choices=["auto", "wavefront", "tile"])
    args = ap.parse_args()
    # rank 0 prints ONE JSON line on stdout: anything a library writes there meanwhile (NCCL prints its version on
    # the first collective when NCCL_DEBUG is set on the box) goes to stderr instead
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        return impl_reference(args)

    import torch
    import torch.distributed as dist
    import norib200 as nb

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the render path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    scene = nb.read_nbs(SCENE_NBS)
    params = nb.make_render_params(scene, "path_mis", seed=1, pool_paths=args.pool)
    ctx = nb.Context(local)
    ctx.set_option("strategy", {"auto": 0, "wavefront": 1, "tile": 2}[args.strategy])
    ctx.load(scene)
    ctx.render_begin(params)
    film = nb.film_tensor(ctx, local)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")      # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        ctx.film_clear()
        ctx.render_subframe(0, args.spp * world, world, rank)
        nb.reduce_film(film, dst=0)

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    dev_ms, c0 = 0.0, None
    launches = samples = rays = 0
    kernel_ms = 0.0
    barrier()
    with ClockSampler(local) as clk:
        for _ in range(args.steps):
            flush.zero_()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            step()
            e1.record()
            e1.synchronize()
            dev_ms += e0.elapsed_time(e1)
            c = ctx.counters()
            launches += c["kernel_launches"] + (1 if world > 1 else 0)
            kernel_ms += c["device_ms"]
            single_launch = c["kernel_launches"] == 1
            samples += c["samples"]; rays += c["rays_closest"] + c["rays_any"]
    barrier()
    t = torch.tensor([dev_ms], dtype=torch.float64, device="cuda")
    tot = torch.tensor([float(samples), float(rays)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms = float(t.item()); samples_all, rays_all = float(tot[0]), float(tot[1])

    # ---- e2e: host buffers in, host film out, every step
    h2d = sum(int(scene[k].nbytes) for k in scene if k not in ("render", "spp"))
    rows, cols, _ = ctx.film_shape()
    host_film = np.empty((rows, cols, 4), dtype=np.float32)
    barrier()
    t0 = time.perf_counter()
    e2e_samples = 0
    for _ in range(args.steps):
        ctx.load(scene)                              # H2D: scene arrays + BVH, through the C ABI
        ctx.render_begin(params)
        ctx.render_subframe(0, args.spp * world, world, rank)
        nb.reduce_film(nb.film_tensor(ctx, local), dst=0)
        ctx.film_download(normalize=False, out=host_film)   # D2H
        e2e_samples += ctx.counters()["samples"]
    barrier()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    se = torch.tensor([float(e2e_samples)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX); dist.all_reduce(se, op=dist.ReduceOp.SUM)

    if rank == 0:
        stats = json.load(open(os.path.join(GOLDEN, "cbox_raystats.json")))
        # algorithmic bytes per ray: 32 B per node visited + 48 B per primitive tested
        # + 32 B ray in + 16 B hit out, with the oracle's mean counters for this ray population
        bytes_per_ray = 32.0 * stats["mean_nodes_visited"] + 48.0 * stats["mean_prims_tested"] + 48.0
        peak, how = measured_peak()
        note = ("cbox is 14 primitives (L1-resident): the kernel is SM-issue-bound, the HBM figure is the "
                "algorithmic-bytes convention of SURVEY 8(d); see DESIGN.md for the issue roofline")
        if single_launch:
            # K_tile: the step IS one launch of the dominant kernel, timed by the library's own
            # CUDA events on its launch stream during the timed steps above
            achieved = bytes_per_ray * rays / (kernel_ms * 1e-3) / 1e9
            # the committed `ncu --set full` capture of this very launch (cbox, 512 spp, one GPU): DRAM traffic per
            # launch, and the warp-instruction count that gives the kernel's real (instruction-issue) roofline
            traffic, issue = None, None
            cap = ncu_capture(os.path.join(ROOT, "profiles", "r01_ncu_tile_cbox.csv"))
            if cap and args.spp == 512 and world == 1:
                traffic = cap.get("dram__bytes_read.sum", 0.0) + cap.get("dram__bytes_write.sum", 0.0)
                warp_inst = cap.get("smsp__inst_executed.sum")
                clk_hz = (clk.summary().get("sm_mhz") or 1965.0) * 1e6
                if warp_inst:
                    sms = torch.cuda.get_device_properties(local).multi_processor_count
                    floor_ms = warp_inst / (sms * 4 * clk_hz) * 1e3          # one warp instruction per scheduler per cycle
                    issue = {"bound": "sm_issue", "warp_inst_per_launch": warp_inst, "schedulers": sms * 4, "sm_mhz": clk_hz / 1e6,
                             "floor_ms": floor_ms, "frac": floor_ms / (kernel_ms / args.steps),
                             "lanes_per_warp_inst": cap.get("smsp__thread_inst_executed_per_inst_executed.ratio"),
                             "source": "profiles/r01_ncu_tile_cbox.csv (instruction count) / live launch time"}
            roofline = {"bound": "hbm", "kernel": "nb_k_tile<path_mis>", "achieved": achieved, "peak": peak, "unit": "GB/s",
                        "frac": achieved / peak, "traffic": traffic, "peak_source": how, "bytes_per_ray": bytes_per_ray,
                        "launches_timed": args.steps, "avg_launch_ms": kernel_ms / args.steps,
                        "kernel_share": {"tile": kernel_ms / ms}, "issue": issue, "note": note}
        else:
            # wavefront: dominant kernel K_trace, one extra step with per-kernel event timers
            ctx.render_begin(params)
            ctx.set_option("profile", 1)
            ctx.render_subframe(0, max(16, args.spp // 8))
            ctx.set_option("profile", 0)
            pc = ctx.counters()
            prof_rays = pc["rays_closest"] + pc["rays_any"]
            achieved = bytes_per_ray * prof_rays / (pc["trace_ms"] * 1e-3) / 1e9 if pc["trace_ms"] > 0 else None
            roofline = {"bound": "hbm", "kernel": "nb_k_trace_wave", "achieved": achieved, "peak": peak, "unit": "GB/s",
                        "frac": achieved / peak if achieved else None, "traffic": None, "peak_source": how,
                        "bytes_per_ray": bytes_per_ray, "launches_timed": int(pc["iterations"]),
                        "avg_launch_ms": pc["trace_ms"] / max(1, pc["iterations"]),
                        "kernel_share": {"trace": pc["trace_ms"] / pc["device_ms"], "shade": pc["shade_ms"] / pc["device_ms"], "regen": pc["regen_ms"] / pc["device_ms"]},
                        "note": note}
        cpu = None
        if not args.no_cpu_baseline:
            try:
                j = run_reference(args.ref_spp)
                cpu = {"value": j["samples_per_s"], "unit": "samples/s", "cores": j["threads"], "kind": "reference",
                       "sample": "same scene and 800x600 image at %d spp (%.1f s)" % (args.ref_spp, j["seconds"]),
                       "rays_per_s": j["rays_dedup_per_s"]}
            except Exception as ex:   # the checker is missing: say so, never substitute
                cpu = {"value": None, "unit": "samples/s", "cores": 0, "kind": "reference", "sample": "unavailable: %s" % ex}
        line = {"metric": "samples_per_s", "value": samples_all / (ms * 1e-3), "unit": "samples/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": WORKLOAD, "spp_per_gpu": args.spp, "integrator": "path_mis",
                           "strategy": "tile" if single_launch else "wavefront",
                           "parallelism": "sample slices x%d, scene replicated, 1 NCCL film reduce" % world,
                           "l2": "256 MB flush write between timed steps" + ("; K_tile keeps path state in registers, film 7.8 MB" if single_launch
                                                                                  else "; path pool (1M slots x 164 B) exceeds L2")},
                "rays_per_s": rays_all / (ms * 1e-3), "rays_per_sample": rays_all / samples_all,
                "e2e": {"value": float(se.item()) / float(te.item()), "unit": "samples/s", "h2d_bytes_per_step": h2d,
                        "d2h_bytes_per_step": int(host_film.nbytes)},
                "gpu_launches": int(launches), "clocks": clk.summary(), "roofline": roofline, "cpu_baseline": cpu}
        emit(line)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
