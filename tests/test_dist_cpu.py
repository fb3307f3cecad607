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


def _synthetic_spot(theta, idx, n_total):
    """A differentiable stand-in for one spot render: sample `i` lands at a theta-dependent position and is splatted with a
    tent weight into a 16 x 16 x 3 image scaled by 1 / n_total (the spot kernel's 1 / spp): pure torch, float64."""
    import torch
    u = ((idx.double() * 0.6180339887) % 1.0) * 2 - 1
    v = ((idx.double() * 0.7548776662) % 1.0) * 2 - 1
    px = 8 + 3 * theta[0] * u + theta[1] * u * u
    py = 8 + 3 * theta[2] * v + theta[3] * u * v
    ii = torch.arange(16, dtype=torch.float64)
    wx = torch.clamp(1 - (ii[None, :] + 0.5 - px[:, None]).abs(), min=0)
    wy = torch.clamp(1 - (ii[None, :] + 0.5 - py[:, None]).abs(), min=0)
    img = torch.einsum("ni,nj->ij", wy, wx) / n_total
    return img[..., None].expand(16, 16, 3)


def _design_worker(rank, world, port, q):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch
    import torch.distributed as dist
    import lensrt as mi
    from mitsuba3_camera_b200 import parallel as par
    par.init("gloo")
    L = mi.lens_design
    n_total = 4000
    theta = torch.tensor([0.7, 0.3, 0.5, -0.2], dtype=torch.float64, requires_grad=True)
    img = _synthetic_spot(theta, torch.arange(rank, n_total, world), n_total)       # this rank's share of the sample indices
    loss = L.rms_loss(L.sum_over_ranks(img))                                         # the loss sees the SUMMED image
    loss.backward()
    par.all_reduce_sum(theta.grad)
    # the wrong recipe of round 1, for contrast: loss of the partial image, gradients summed
    t2 = theta.detach().clone().requires_grad_(True)
    L.rms_loss(_synthetic_spot(t2, torch.arange(rank, n_total, world), n_total)).backward()
    par.all_reduce_sum(t2.grad)
    q.put((rank, float(loss), theta.grad.tolist(), t2.grad.tolist()))
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_sharded_design_loss_matches_single_process():
    """Sample-index sharding of the lens-design loss: all-reduce the spot image BEFORE the non-linear rms_loss, back-propagate
    the common image gradient on every rank, sum the parameter gradients: equal to the single-process value and gradient
    (the per-rank-loss recipe is not)."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import torch
    import torch.multiprocessing as mp
    import lensrt as mi
    theta = torch.tensor([0.7, 0.3, 0.5, -0.2], dtype=torch.float64, requires_grad=True)
    loss1 = mi.lens_design.rms_loss(_synthetic_spot(theta, torch.arange(0, 4000), 4000)); loss1.backward()
    ctx = mp.get_context("spawn")
    q, port = ctx.Queue(), _free_port()
    ps = [ctx.Process(target=_design_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=240) for _ in ps)
    for p in ps:
        p.join(60)
        assert p.exitcode == 0
    for rank, loss, grad, wrong in res:
        assert abs(loss - float(loss1)) <= 1e-9 * abs(float(loss1))
        assert np.allclose(grad, theta.grad.numpy(), rtol=1e-8, atol=1e-12)
        assert not np.allclose(wrong, theta.grad.numpy(), rtol=0.2)          # sum of per-rank-loss gradients is a different quantity
