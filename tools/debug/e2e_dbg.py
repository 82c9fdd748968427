import os, sys, time
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import norib200 as nb
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1: dist.init_process_group("nccl", device_id=torch.device("cuda", local))
scene = nb.read_nbs("tests/golden/cbox.nbs")
params = nb.make_render_params(scene, "path_mis", seed=1)
ctx = nb.Context(local)
if world > 1: nb.comm_init_torch(ctx)
img = np.empty((600, 800, 3), np.float32)
for it in range(4):
    if world > 1: dist.barrier()
    torch.cuda.synchronize()
    t = [time.perf_counter()]
    ctx.upload_scene(scene); t.append(time.perf_counter())
    ctx.build_bvh(scene["bvh_nodes"], scene["bvh_indices"]); t.append(time.perf_counter())
    ctx.render_begin(params); t.append(time.perf_counter())
    ctx.render_subframe(0, 512, world, rank); t.append(time.perf_counter())
    ctx.film_reduce(); t.append(time.perf_counter())
    ctx.sync(); t.append(time.perf_counter())
    if rank == 0: ctx.film_download(normalize=True, out=img)
    t.append(time.perf_counter())
    print(rank, it, "upload %.2f bvh %.2f begin %.2f subframe %.2f reduce %.2f sync %.2f download %.2f ms" % tuple(1e3 * (b - a) for a, b in zip(t, t[1:])), flush=True)
ctx.close()
if world > 1: dist.destroy_process_group()
