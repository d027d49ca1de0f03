"""Multi-GPU host logic on CPU (world size 2, gloo): tile ownership, the weak-scaling rule and the frame assembly
(relativistic-ray-tracer_b200/shard.py).  The device side of the same rule (KeepOwned in csrc/rrt_kernels.cuh) is checked
against tile_owner() on the GPU in test_gpu_parity.py::test_c2_full_size_properties."""
import importlib
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

shard = importlib.import_module("relativistic-ray-tracer_b200.shard")


@pytest.mark.parametrize("w,h,tile,world", [(480, 360, 16, 2), (481, 359, 16, 3), (33, 17, 32, 8), (5, 5, 16, 4), (1920, 1080, 16, 8)])
def test_tile_owner_is_a_disjoint_cover(w, h, tile, world):
    own = shard.tile_owner(w, h, tile, world)
    assert own.shape == (h, w) and own.min() >= 0 and own.max() < world
    assert sum(shard.owned_pixels(w, h, tile, r, world) for r in range(world)) == w * h
    # constant inside a tile, and neighbouring tiles in a row go to consecutive ranks
    assert np.all(own[: min(tile, h), : min(tile, w)] == own[0, 0])
    if w > tile:
        assert own[0, tile] == (own[0, 0] + 1) % world
    ntx = (w + tile - 1) // tile
    stride = shard.shard_stride(world)
    assert np.gcd(stride, world) == 1
    if h > tile:
        assert own[tile, 0] == stride % world        # the next tile row is shifted by the stride, whatever the frame width
    # balance: no rank holds more than one tile row's worth above the mean
    counts = np.bincount(own.ravel(), minlength=world)
    if w * h >= world * tile * tile * 4:
        assert counts.max() - counts.min() <= tile * tile * (ntx + 1)
    # any window of `world` tiles along a row or along a column holds every rank once (what round 1's index rule lost at 1920 / 16 = 120)
    if w >= world * tile and h >= world * tile:
        assert sorted(own[0, ::tile][:world].tolist()) == list(range(world))
        assert sorted(own[::tile, 0][:world].tolist()) == list(range(world))


def test_box_region_of_the_1080p_frame_is_balanced_over_8_ranks():
    """the pixels that carry the work of C4 (the central 480x360 cell, bench.py ADAPTIVE) split within 0.5 % over 8 ranks"""
    own = shard.tile_owner(1920, 1080, shard.DEFAULT_TILE, 8)[360:720, 720:1200]
    counts = np.bincount(own.ravel(), minlength=8)
    assert counts.max() - counts.min() <= 0.005 * counts.mean()


def test_tile_owner_rejects_bad_arguments():
    for bad in [(0, 4, 16, 1), (4, 0, 16, 1), (4, 4, 0, 1), (4, 4, 16, 0)]:
        with pytest.raises(ValueError):
            shard.tile_owner(*bad)


def test_weak_scaling_rule():
    assert shard.weak_scaling_params(16, 1) == (16, 32)
    assert shard.weak_scaling_params(16, 2) == (32, 32)
    assert shard.weak_scaling_params(16, 8) == (128, 128)   # one batch: no adaptive stop inside a bench step


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, w, h, tile, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(7)
        full = torch.rand((h, w, 4), generator=g, dtype=torch.float32)        # what a single rank would render
        own = torch.from_numpy(shard.tile_owner(w, h, tile, world))
        mine = full * (own == rank).unsqueeze(-1)                             # unowned pixels stay zero, as on the device
        got = shard.assemble(mine.clone(), dst=0)
        # max-over-ranks timing, as bench.py does it
        t = torch.tensor([float(rank + 1)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ok = bool(t.item() == world)
        if rank == 0:
            ok = ok and torch.equal(got, full)                                 # exact: every pixel has one non-zero term
        out[rank] = ok
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("w,h,tile", [(96, 72, 16), (50, 35, 16)])
def test_assemble_over_gloo_world2(w, h, tile):
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), w, h, tile, out), nprocs=world, join=True)
    assert dict(out) == {0: True, 1: True}


def test_assemble_is_identity_without_a_process_group():
    f = torch.ones((2, 2, 4))
    assert shard.assemble(f) is f
