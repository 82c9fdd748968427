"""One rank of a torchrun job (tests/test_multigpu_gpu.py): renders its sample slice of the Cornell box, joins the
library's own NCCL communicator (nb_comm_init_rank; torch.distributed only carries the 128-byte id), reduces the film
with nb_film_reduce and, on rank 0, writes the summed accumulators."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import norib200 as nb  # noqa: E402


def main():
    out, spp = sys.argv[1], int(sys.argv[2])
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")                 # the id needs no GPU collective
    scene = nb.read_nbs(os.path.join(os.path.dirname(__file__), "..", "golden", "cbox.nbs"))
    cam = scene["camera"].copy()
    cam["width"], cam["height"] = 200, 150
    cam["inv_output_size"] = [1.0 / 200, 1.0 / 150]
    scene["camera"] = cam
    with nb.Context(local) as ctx:
        nb.comm_init_torch(ctx)
        info = ctx.comm_info()
        assert info["n_ranks"] == world and info["rank"] == rank and info["nccl_version"] > 0
        ctx.load(scene)
        ctx.render_begin(nb.make_render_params(scene, "path_mis", seed=7))
        for first in (0, spp // 2):                 # two subframes, two reduces: the second must not count the first twice
            ctx.render_subframe(first, spp // 2, world, rank)
            ctx.film_reduce()
        film = ctx.film_download(normalize=False)
        c = ctx.counters()
        if rank == 0:
            np.save(out, film)
        else:
            assert not film.any()                   # the samples moved to the root
        assert c["samples"] == 200 * 150 * (spp // world)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
