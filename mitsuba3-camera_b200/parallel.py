"""Multi-GPU plumbing: one process per GPU (torch.distributed; NCCL on GPUs, gloo in the CPU tests).

The reference has no multi-device code. The path shards without any data-path exchange: every path is
a function of (seed, sample index) only. Render: pixel tiles of 32x1 are dealt round-robin over ranks (the kernel does the
dealing from shard_index / shard_count, lrt_render.cu); spot path: sample idx = rank + k * world. The one collective per
step is a sum of the film accumulator (and of the lens-table gradient)."""
from __future__ import annotations

import os

TILE_W, TILE_H = 32, 1      # lrt_render.cu kTileW / kTileH


def shard_from_env():
    """(rank, world, local_rank) from the torchrun environment (single process: (0, 1, 0))."""
    return int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))


def init(backend=None):
    """Join the process group if launched under torchrun; returns (rank, world)."""
    import torch
    import torch.distributed as dist
    rank, world, local = shard_from_env()
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    return rank, world


def tiles_of_shard(width, height, rank, world):
    """Tile ids (row-major over ceil(W/32) x ceil(H/1)) rendered by `rank`: tile_id = rank + k * world — the same formula as
    the kernel (lrt_render.cu: tile_id = shard_index + (item / n_chunks) * shard_count)."""
    tx, ty = (width + TILE_W - 1) // TILE_W, (height + TILE_H - 1) // TILE_H
    return list(range(rank, tx * ty, world)), tx


def pixel_owner_mask(width, height, rank, world):
    """bool [H, W]: pixels whose samples `rank` generates."""
    import numpy as np
    ids, tx = tiles_of_shard(width, height, rank, world)
    m = np.zeros((height, width), bool)
    for t in ids:
        y0, x0 = (t // tx) * TILE_H, (t % tx) * TILE_W
        m[y0:y0 + TILE_H, x0:x0 + TILE_W] = True
    return m


def samples_of_shard(n_total, rank, world):
    """Number of sample indices idx = rank + k * world below n_total (lrt_spot.cu n_my)."""
    return (n_total - rank + world - 1) // world


def all_reduce_sum(t):
    """Sum a tensor over ranks in place (no-op for a single process). Host tensors under NCCL go through the device."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return t
    if dist.get_backend() == "nccl" and not t.is_cuda:
        g = t.cuda(); dist.all_reduce(g); t.copy_(g.cpu())
    else:
        dist.all_reduce(t)
    return t


def render_sharded(mi, scene, seed=0, spp=0, storage=None, develop=True):
    """mi.render over all ranks: every rank renders its tiles into its own accumulator, one all-reduce merges them."""
    rank, world, _ = shard_from_env()
    st = mi.render(scene, seed=seed, spp=spp, develop=False, shard=(rank, world), storage=storage)
    all_reduce_sum(st)
    return mi.develop_rgb(st) if develop else st
