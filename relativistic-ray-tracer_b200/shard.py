"""Host-side logic of the multi-GPU path: which rank renders which pixels, how many samples a
rank takes under bench.py's weak-scaling rule, and the one collective that assembles the frame.

The reference distributes 32x32 image tiles to host threads through a work queue (reference
src/pathtracer.cpp:224-282, WorkQueue in src/work_queue.h); across GPUs the same tiles are dealt round-robin, which
needs no queue and no communication while rendering: a pixel's estimate depends only on (seed, pixel, sample).
"""
import numpy as np

DEFAULT_TILE = 8


def shard_stride(world):
    """Row-to-row shift of the owner pattern: the smallest integer >= round(sqrt(world)) coprime with world (mirrors
    rrt_shard_stride in csrc/rrt_kernels.cuh): 1 -> 1, 2 -> 1, 3 -> 2, 4 -> 3, 6 -> 5, 8 -> 3."""
    s = 1
    while (s + 1) * (s + 1) <= world + s:
        s += 1
    while np.gcd(s, world) != 1:
        s += 1
    return int(s)


def tile_owner(width, height, tile=DEFAULT_TILE, world=1):
    """[height, width] int32: rank that renders each pixel.  Mirrors KeepOwned in csrc/rrt_kernels.cuh:
    owner of tile (tx, ty) = (tx + ty * shard_stride(world)) % world — every run of `world` tiles along a row or a column holds
    every rank once, whatever the frame width (the tile index of round 1, tx + ty * tiles_per_row, degenerates into vertical
    stripes when tiles_per_row is a multiple of world: 1920 / 16 = 120 tiles with 8 ranks)."""
    if width <= 0 or height <= 0 or tile <= 0 or world <= 0:
        raise ValueError("tile_owner: width, height, tile and world must be positive")
    x = np.arange(width, dtype=np.int64) // tile
    y = np.arange(height, dtype=np.int64) // tile
    return ((x[None, :] + y[:, None] * shard_stride(world)) % world).astype(np.int32)


def owned_pixels(width, height, tile, rank, world):
    """number of pixels rank `rank` renders"""
    return int((tile_owner(width, height, tile, world) == rank).sum())


def weak_scaling_params(spp_per_gpu, world, default_batch=32):
    """bench.py's rule: the frame is fixed and the job takes spp_per_gpu * world samples per pixel, so every rank traces
    about width * height * spp_per_gpu paths.  Returns (ns_aa, samples_per_batch): the adaptive batch stays the
    reference default while ns_aa fits in it, else one batch = the whole job (no adaptive stop inside a bench step)."""
    ns_aa = int(spp_per_gpu) * int(world)
    return ns_aa, (default_batch if ns_aa <= default_batch else ns_aa)


def assemble(frame, dst=0, group=None):
    """Sum the per-rank frames (unowned pixels are zero) onto rank `dst`: the path's only collective.
    `frame` is a torch tensor on the backend's device (CUDA for nccl, CPU for gloo); returned for convenience."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(frame, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return frame
