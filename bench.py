#!/usr/bin/env python
"""bench.py — headline benchmark: samples/s and rays/s on the Cornell box, path_mis.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--spp S] [--scaling strong|weak]

Workload (BASELINE.json configs[0], the scene the north-star target is quoted on):
tests/golden/scenes/cbox/cbox_path_mis.xml, 800x600, 512 spp, gaussian r=2 — one "step" is one
full render of it. N>1 (torchrun, one rank per GPU): STRONG scaling by default — the one 512-spp render is split
into sample slices (rank g renders the samples s with s % N == g of every pixel, 512/N spp each), scene replicated,
the per-rank film accumulators summed onto rank 0 by ONE NCCL reduce issued by the library itself (nb_film_reduce)
inside the timed region. `--scaling weak` gives every rank its own 512 spp instead (round 1's curve).

`value`   camera samples/s with the scene already resident in HBM: film clear + render + film reduce, timed with CUDA
          events recorded on the library's own stream (nb_timer_mark), max over ranks.
`e2e`     the same through the C ABI with HOST buffers every step: scene + BVH upload from pinned host memory (H2D),
          render, reduce, ImageBlock::toBitmap on the device (nb_k_film_finalize) and the download of the finished
          image into pinned host memory (D2H) on rank 0, wall-clocked around the calls, max over ranks.
`roofline` the dominant kernel against the roofline that binds it: SM instruction issue (the 14-primitive scene is
          L1-resident; SURVEY 8d names issue as C1's roof). `configs` (N=1 only): the BVH kernels on C2 table and C3.
`--impl reference` times the reference's own CPU renderer (oracle/_ref/nori_ref, all host
threads) on a bounded sample of the same workload (same scene and image size, fewer spp).
"""
import argparse
import hashlib
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
NCU_CAPTURE = os.path.join(ROOT, "profiles", "r02_ncu_tile_cbox.csv")       # ncu --set full of this very launch + .meta.json


def run_reference(spp, threads=None, xml=SCENE_XML):
    """One bounded run of the reference CPU renderer; returns its JSON line."""
    if not os.path.exists(NORI_REF):
        raise FileNotFoundError("oracle/_ref/nori_ref is missing: run `make -C oracle ref` where /root/reference is mounted")
    args = [NORI_REF, "render", xml, "/tmp/nb_bench_ref.npy", "spp=%d" % spp]
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


def kernel_source_sha():
    """Fingerprint of what the kernels are built from: an ncu capture only describes the binary it was taken on."""
    h = hashlib.sha256()
    d = os.path.join(ROOT, "optix-renderer_b200", "csrc")
    for f in sorted(os.listdir(d)):
        if f.endswith((".cu", ".cuh", ".h")):
            h.update(f.encode()); h.update(open(os.path.join(d, f), "rb").read())
    h.update(open(os.path.join(ROOT, "optix-renderer_b200", "build.py"), "rb").read())
    return h.hexdigest()[:16]


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
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sample": "%d spp per step instead of 512 (throughput is spp-independent: "
                                                       "BASELINE.md has the full 512-spp run next to the 16-spp one)" % spp},
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


def thread_inst(cap):
    """thread-instructions of a captured launch: the counter itself when the report has it, else warp instructions x lanes"""
    v = cap.get("smsp__thread_inst_executed.sum", 0.0)
    if not v:
        v = cap.get("smsp__inst_executed.sum", 0.0) * cap.get("smsp__thread_inst_executed_per_inst_executed.ratio", 0.0)
    return v


def pinned(shape, dtype=np.float32):
    """numpy view of page-locked host memory (the e2e legs copy from / to pinned buffers)."""
    import torch
    t = torch.empty(int(np.prod(shape)) * np.dtype(dtype).itemsize, dtype=torch.uint8, pin_memory=True)
    a = t.numpy().view(dtype).reshape(shape)
    a.flags.writeable = True
    return a, t


def pin_scene(scene):
    keep, out = [], {}
    for k, v in scene.items():
        v = np.ascontiguousarray(v)
        if v.size == 0:
            out[k] = v
            continue
        a, t = pinned(v.shape, v.dtype)
        a[...] = v
        out[k] = a; keep.append(t)
    return out, keep


def tree_config(nb, ctx, name, xml, spp, sms, clk_mhz, cap_csv):
    """One BVH-walking configuration, driver-timed: samples/s, rays/s and (from the committed ncu capture of the same
    kernel, when its source fingerprint matches) the lane-issue fraction."""
    t0 = time.perf_counter()
    s = nb.load_xml(xml)
    host_s = time.perf_counter() - t0
    ctx.load(s)
    ctx.render_begin(nb.make_render_params(s, seed=1))
    best = None
    for _ in range(3):
        ctx.film_clear()
        ctx.render_subframe(0, spp)
        c = ctx.counters()
        if best is None or c["device_ms"] < best["device_ms"]:
            best = c
    rays = best["rays_closest"] + best["rays_any"]
    row = {"config": name, "triangles": int(s["faces"].size // 3), "spp": spp, "ms": best["device_ms"],
           "samples_per_s": best["samples"] / best["device_ms"] * 1e3, "rays_per_s": rays / best["device_ms"] * 1e3,
           "host_load_flatten_bvh_s": round(host_s, 2)}
    # the same scene on the tree the device builds itself (NB_BVH_DEVICE_SAH: no host build at all)
    t0 = time.perf_counter()
    ctx.load(s, device_bvh="sah")
    dev_load_s = time.perf_counter() - t0
    ctx.render_begin(nb.make_render_params(s, seed=1))
    dbest = None
    for _ in range(3):
        ctx.film_clear()
        ctx.render_subframe(0, spp)
        c = ctx.counters()
        if dbest is None or c["device_ms"] < dbest["device_ms"]:
            dbest = c
    row["device_sah"] = {"build_ms": round(ctx.bvh_stats()["build_ms"], 3), "upload_build_collapse_s": round(dev_load_s, 3), "ms": dbest["device_ms"],
                         "samples_per_s": dbest["samples"] / dbest["device_ms"] * 1e3,
                         "rays_per_s": (dbest["rays_closest"] + dbest["rays_any"]) / dbest["device_ms"] * 1e3}
    cap = ncu_capture(cap_csv)
    meta = cap_csv[:-4] + ".meta.json"
    if cap and os.path.exists(meta):
        m = json.load(open(meta))
        per_sample = thread_inst(cap) / max(1.0, m.get("samples", 0.0))
        lanes = sms * 128 * clk_mhz * 1e6
        row["lane_issue_frac"] = per_sample * row["samples_per_s"] / lanes if per_sample else None
        row["lanes_per_warp_inst"] = cap.get("smsp__thread_inst_executed_per_inst_executed.ratio")
        row["capture"] = os.path.relpath(cap_csv, ROOT)
        row["capture_matches_build"] = m.get("src_sha") == kernel_source_sha()
    return row


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--spp", type=int, default=512)
    ap.add_argument("--ref-spp", type=int, default=16)
    ap.add_argument("--pool", type=int, default=0)
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true")
    ap.add_argument("--no-flush", action="store_true", help="skip the 256 MB L2-flush write between steps (for the ncu capture that attributes DRAM writes)")
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
    if args.scaling == "strong" and args.spp % world:
        raise SystemExit("bench.py: --spp %d does not split over %d ranks" % (args.spp, world))
    spp_total = args.spp if args.scaling == "strong" else args.spp * world      # samples of one step, all ranks together
    scene = nb.read_nbs(SCENE_NBS)
    params = nb.make_render_params(scene, "path_mis", seed=1, pool_paths=args.pool)
    ctx = nb.Context(local)
    ctx.set_option("strategy", {"auto": 0, "wavefront": 1, "tile": 2}[args.strategy])
    if world > 1:
        nb.comm_init_torch(ctx)             # the library's own NCCL communicator; torch only carries the 128-byte id
    ctx.load(scene)
    ctx.render_begin(params)
    flush = None if args.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device="cuda")      # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ctx.sync()

    def step():
        """film clear -> this rank's sample slice -> one NCCL sum onto rank 0; all queued on the library's stream"""
        ctx.film_clear()
        ctx.render_subframe(0, spp_total, world, rank)
        ctx.film_reduce()

    for _ in range(args.warmup):
        step()
    ctx.sync()
    dev_ms = 0.0
    launches = samples = rays = 0
    kernel_ms = 0.0
    barrier()
    with ClockSampler(local) as clk:
        t_region = time.perf_counter()
        for _ in range(args.steps):
            if flush is not None:
                flush.zero_()
            barrier()                       # every rank starts the step together: the reduce then waits for the slowest
            ctx.timer_mark(0)
            step()
            ctx.timer_mark(1)
            dev_ms += ctx.timer_elapsed()
            c = ctx.counters()
            launches += c["kernel_launches"] + (1 if world > 1 else 0)
            kernel_ms += c["device_ms"]
            single_launch = c["kernel_launches"] == 1
            samples += c["samples"]; rays += c["rays_closest"] + c["rays_any"]
        # nvidia-smi answers every 100 ms at best: a timed region shorter than that (8 GPUs: 5 x 18 ms) would end with no
        # clock sample at all. Keep the same load running, untimed, until half a second has been covered; the sampler sees the
        # timed steps and these.
        region_s = time.perf_counter() - t_region
        extra = torch.tensor([0 if region_s >= 0.5 else int(np.ceil((0.5 - region_s) / max(region_s / max(1, args.steps), 1e-4)))],
                             dtype=torch.int64, device="cuda")
        if world > 1:
            dist.broadcast(extra, src=0)
        clock_extra_steps = int(extra.item())
        for _ in range(clock_extra_steps):
            step()
        if clock_extra_steps:
            ctx.sync()
    barrier()
    t = torch.tensor([dev_ms], dtype=torch.float64, device="cuda")
    tot = torch.tensor([float(samples), float(rays)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms = float(t.item()); samples_all, rays_all = float(tot[0]), float(tot[1])

    # ---- e2e: pinned host buffers in, finished image in pinned host memory out, every step
    host_scene, _keep = pin_scene(scene)
    h2d = sum(int(host_scene[k].nbytes) for k in host_scene if k not in ("render", "spp"))
    cam = scene["camera"][0] if scene["camera"].shape else scene["camera"]
    host_img, _keep2 = pinned((int(cam["height"]), int(cam["width"]), 3))
    d2h = int(host_img.nbytes)

    def e2e_step():
        ctx.load(host_scene)                         # H2D: scene arrays + BVH, through the C ABI
        ctx.render_begin(params)
        ctx.render_subframe(0, spp_total, world, rank)
        ctx.film_reduce()
        if rank == 0:
            ctx.film_download(normalize=True, out=host_img)   # ImageBlock::toBitmap on the device + D2H
        else:
            ctx.sync()
        return ctx.counters()["samples"]

    for _ in range(min(2, args.warmup)):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    e2e_samples = 0
    e2e_step_ms = []
    for _ in range(args.steps):
        t1 = time.perf_counter()
        e2e_samples += e2e_step()
        e2e_step_ms.append(round(1e3 * (time.perf_counter() - t1), 3))
    barrier()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    se = torch.tensor([float(e2e_samples)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX); dist.all_reduce(se, op=dist.ReduceOp.SUM)
    assert np.isfinite(host_img).all() or rank != 0

    if rank == 0:
        stats = json.load(open(os.path.join(GOLDEN, "cbox_raystats.json")))
        clocks = clk.summary()
        clocks["window"] = "the timed steps" if not clock_extra_steps else "the timed steps + %d untimed steps of the same load (the timed region is shorter than nvidia-smi's 100 ms period)" % clock_extra_steps
        clk_mhz = clocks.get("sm_mhz") or 1965.0
        sms = torch.cuda.get_device_properties(local).multi_processor_count
        # algorithmic bytes per ray (SURVEY 8d's convention): 32 B per node visited + 48 B per primitive tested
        # + 32 B ray in + 16 B hit out, with the oracle's mean counters for this ray population
        bytes_per_ray = 32.0 * stats["mean_nodes_visited"] + 48.0 * stats["mean_prims_tested"] + 48.0
        peak, how = measured_peak()
        avg_launch_ms = kernel_ms / args.steps
        lanes_peak = sms * 128 * clk_mhz * 1e6       # thread-instructions/s: SMs x 4 schedulers x 32 lanes x clock (8d, C1)
        roofline = {"bound": "sm_issue", "kernel": "nb_k_tile<path_mis, FLAT>" if single_launch else "wavefront (K_shade + K_trace)",
                    "achieved": None, "peak": lanes_peak / 1e12, "unit": "T lane-inst/s", "frac": None, "traffic": None,
                    "peak_source": "%d SMs x 128 lanes x %.0f MHz (median SM clock sampled during the timed steps)" % (sms, clk_mhz),
                    "launches_timed": args.steps, "avg_launch_ms": avg_launch_ms, "kernel_share": {"tile": kernel_ms / ms},
                    "note": "cbox is 14 primitives (L1-resident): the kernel is bound by SM instruction issue, not by HBM; "
                            "frac = thread-instructions/s / (SMs x 128 lanes x clock), instruction counts per sample from the "
                            "committed ncu --set full capture of this launch, launch time measured live"}
        cap = ncu_capture(NCU_CAPTURE)
        meta_p = NCU_CAPTURE[:-4] + ".meta.json"
        if cap and os.path.exists(meta_p) and single_launch:
            meta = json.load(open(meta_p))
            n_cap = float(meta["samples"])
            t_inst = thread_inst(cap) / n_cap * (samples_all / args.steps / world)
            warp_inst = cap.get("smsp__inst_executed.sum", 0.0) / n_cap * (samples_all / args.steps / world)
            roofline.update({
                "achieved": t_inst / (avg_launch_ms * 1e-3) / 1e12,
                "frac": t_inst / (avg_launch_ms * 1e-3) / lanes_peak,
                "warp_issue_frac": warp_inst / (avg_launch_ms * 1e-3) / (sms * 4 * clk_mhz * 1e6),
                "lanes_per_warp_inst": cap.get("smsp__thread_inst_executed_per_inst_executed.ratio"),
                "traffic": (cap.get("dram__bytes_read.sum", 0.0) + cap.get("dram__bytes_write.sum", 0.0)) / n_cap * (samples_all / args.steps / world),
                "capture": os.path.relpath(NCU_CAPTURE, ROOT), "capture_src_sha": meta.get("src_sha"),
                "capture_matches_build": meta.get("src_sha") == kernel_source_sha()})
        algorithmic = {"convention": "SURVEY 8(d) algorithmic bytes: (32 B x nodes + 48 B x primitives + 48 B) per ray; on cbox these "
                                     "bytes never leave L1, so this is NOT an HBM roofline — kept as the contract's per-ray figure",
                       "bytes_per_ray": bytes_per_ray, "gb_per_s": bytes_per_ray * rays / (kernel_ms * 1e-3) / 1e9 / 1.0,
                       "hbm_peak_gb_per_s": peak, "peak_source": how}
        roofline["algorithmic_bytes"] = algorithmic
        cpu = None
        if not args.no_cpu_baseline:
            try:
                j = run_reference(args.ref_spp)
                cpu = {"value": j["samples_per_s"], "unit": "samples/s", "cores": j["threads"], "kind": "reference",
                       "sample": "same scene and 800x600 image at %d spp (%.1f s)" % (args.ref_spp, j["seconds"]),
                       "rays_per_s": j["rays_dedup_per_s"]}
            except Exception as ex:   # the checker is missing: say so, never substitute
                cpu = {"value": None, "unit": "samples/s", "cores": 0, "kind": "reference", "sample": "unavailable: %s" % ex}
        configs = None
        if world == 1 and not args.no_configs:
            configs = []
            try:
                configs.append(tree_config(nb, ctx, "C2 table path_mis 800x600 (22 766 triangles)",
                                           os.path.join(GOLDEN, "scenes", "table", "table_path_mis.xml"), 64, sms, clk_mhz,
                                           os.path.join(ROOT, "profiles", "r02_ncu_walk_table.csv")))
                import tempfile
                d = tempfile.mkdtemp(prefix="c3_")
                subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_c3_scene.py"), d, "spp=16"], check=True,
                               stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
                configs.append(tree_config(nb, ctx, "C3 synthetic 1 310 720 triangles path_mis 800x600",
                                           os.path.join(d, "c3_path_mis.xml"), 16, sms, clk_mhz,
                                           os.path.join(ROOT, "profiles", "r02_ncu_walk_c3.csv")))
            except Exception as ex:
                configs.append({"error": str(ex)[:300]})
        line = {"metric": "samples_per_s", "value": samples_all / (ms * 1e-3), "unit": "samples/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": WORKLOAD, "spp_total": spp_total, "spp_per_gpu": spp_total // world, "integrator": "path_mis",
                           "strategy": "tile" if single_launch else "wavefront",
                           "parallelism": "sample slices x%d of ONE render, scene replicated, 1 NCCL film reduce issued by the library "
                                          "(nb_film_reduce) inside the timed region" % world,
                           "timed": "film clear + render + reduce, CUDA events on the library's stream; every rank starts each step at a barrier",
                           "l2": ("no flush (--no-flush)" if args.no_flush else "256 MB flush write between timed steps") + ("; K_tile keeps path state in registers, film 7.8 MB" if single_launch
                                                                                  else "; path pool (1M slots x 164 B) exceeds L2")},
                "rays_per_s": rays_all / (ms * 1e-3), "rays_per_sample": rays_all / samples_all,
                "e2e": {"value": float(se.item()) / float(te.item()), "unit": "samples/s", "h2d_bytes_per_step": h2d,
                        "d2h_bytes_per_step": d2h, "step_ms_rank0": e2e_step_ms,
                        "what": "pinned scene + BVH upload, render, reduce, toBitmap on device, finished image to pinned host memory (rank 0)"},
                "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu}
        if configs is not None:
            line["configs"] = configs
        emit(line)
    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
