"""World-size-2 gloo tests of the multi-GPU host logic (no GPU): tile dealing partitions the film, sample-index sharding
covers every index once, and the all-reduce helper sums film accumulators and table gradients across ranks."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch
    import torch.distributed as dist
    import lensrt as mi
    from mitsuba3_camera_b200 import parallel as par
    r, w = par.init("gloo")
    assert (r, w) == (rank, world)
    W, H = 100, 37                                   # ragged: not a multiple of the 32-pixel tile
    film = torch.zeros((H, W, 4))
    film[torch.from_numpy(par.pixel_owner_mask(W, H, r, w))] = 1.0      # stand-in for this rank's accumulator
    par.all_reduce_sum(film)
    n_total = W * H * 3
    mine = torch.tensor([par.samples_of_shard(n_total, r, w)], dtype=torch.int64)
    par.all_reduce_sum(mine)
    grad = torch.full((4, 16), float(r + 1))         # table gradient of this rank's samples
    par.all_reduce_sum(grad)
    ms = torch.tensor([10.0 + r], dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)        # bench.py's max-over-ranks timing
    q.put((rank, bool((film == 1).all()), int(mine.item()) == n_total, float(grad[0, 0]), float(ms.item())))
    dist.destroy_process_group()


def test_tile_and_sample_sharding_partition():
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import lensrt  # noqa
    from mitsuba3_camera_b200 import parallel as par
    for (W, H, world) in [(100, 37, 2), (3840, 2160, 8), (33, 5, 3), (1, 1, 4)]:
        total = np.zeros((H, W), int)
        for r in range(world):
            total += par.pixel_owner_mask(W, H, r, world)
        assert (total == 1).all()
        n = W * H * 5
        assert sum(par.samples_of_shard(n, r, world) for r in range(world)) == n
        idx = np.concatenate([np.arange(r, n, world) for r in range(world)])
        assert len(np.unique(idx)) == n


@pytest.mark.timeout(300)
def test_gloo_world_size_2_allreduce():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q, port = ctx.Queue(), _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=240) for _ in ps)
    for p in ps:
        p.join(60)
        assert p.exitcode == 0
    for rank, film_ok, samples_ok, g, ms in res:
        assert film_ok and samples_ok and g == 3.0 and ms == 11.0
