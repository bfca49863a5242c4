"""Host-side sharding rules of the multi-GPU path (mirrors RendererImpl::init / render in csrc/renderer_impl.cuh).

Frames shard by sample range or by 32x32 image tile; the only exchange is one sum-reduce of the
{sum r, sum g, sum b, count} buffer per frame.  These helpers are what bench.py and multi-process drivers use to
decide who renders what and to combine partial estimator buffers through torch.distributed (NCCL on GPUs, gloo in tests).
"""
from __future__ import annotations

TILE = 32  # PATHS_SHARD_TILE in csrc/renderer_impl.cuh


def sample_slice(done, spp, rank, world):
    """Samples [lo, hi) of a render(spp) call that `rank` traces under sample sharding."""
    world = max(int(world), 1)
    return done + spp * rank // world, done + spp * (rank + 1) // world


def tile_owner(x, y, width, world):
    """Rank that owns pixel (x, y) under tile sharding: tiles dealt round-robin in row-major tile order."""
    tiles_x = (width + TILE - 1) // TILE
    return ((y // TILE) * tiles_x + (x // TILE)) % max(int(world), 1)


def owned_pixels(width, height, rank, world):
    import numpy as np

    ys, xs = np.mgrid[0:height, 0:width]
    tiles_x = (width + TILE - 1) // TILE
    owner = ((ys // TILE) * tiles_x + (xs // TILE)) % max(int(world), 1)
    return np.nonzero(owner.reshape(-1) == rank)[0].astype(np.uint32)


def combine(sums, dst=0, group=None):
    """Sum-reduces a partial estimator buffer (torch tensor, any device) onto rank `dst`; returns it on dst, None elsewhere."""
    import torch.distributed as dist

    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return sums
    dist.reduce(sums, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return sums if dist.get_rank(group) == dst else None
