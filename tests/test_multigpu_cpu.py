"""N>1 host logic on CPU (gloo, world_size 2): sample slices partition the sample range, and the
film reduce is a plain sum to rank 0 (the path's only collective)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def test_slices_partition(nb, built_lib):
    for first, count, stride in [(0, 512, 1), (0, 512, 8), (3, 17, 4), (5, 2, 8), (7, 0, 2), (100, 33, 5)]:
        seen = []
        for off in range(stride):
            s0, n = nb.sample_slice(first, count, stride, off)
            seen += [s0 + k * stride for k in range(n)]
            assert all(s % stride == off for s in seen[-n:] if n)
        assert sorted(seen) == list(range(first, first + count))
    import pytest
    with pytest.raises(nb.NbError):
        nb.sample_slice(0, 8, 2, 2)


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import norib200 as nb
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(100 + rank)
    film = torch.from_numpy(rng.random((20, 24, 4), dtype=np.float32))
    mine = film.clone()
    nb.reduce_film(film, dst=0)
    gathered = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine)
    if rank == 0:
        want = gathered[0] + gathered[1]
        out.put(bool(torch.allclose(film, want, rtol=0, atol=0)))
        # each rank's slice of an 8-sample budget
        out.put([nb.sample_slice(0, 8, world, r) for r in range(world)])
    dist.destroy_process_group()


def test_film_reduce_gloo_world2():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs: p.start()
    for p in procs: p.join(120)
    assert all(p.exitcode == 0 for p in procs)
    assert q.get(timeout=10) is True
    assert q.get(timeout=10) == [(0, 4), (1, 4)]
