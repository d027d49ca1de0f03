"""The algorithm of csrc/rrt_bvh_device.cuh part 1, restated in numpy and checked on the CPU: construct_bvh's depth-first primitive
order (reference src/bvh.cpp:49-96) is the fixed point of a LEVEL-SYNCHRONOUS stable partition — every segment of more than 4
primitives is split at once by `centroid[axis] < midpoint(segment box)`, positions computed from one exclusive scan — and equals
the recursive host restatement (rrt_bvh_build_host, which tests/test_oracle_kat.py pins to the reference's own traversal
order) integer for integer.  The device kernels do the same steps with atomics and cub; they are tested in test_gpu_parity.py."""
import importlib

import numpy as np
import pytest

from util import load_scene, sc

api = importlib.import_module("relativistic-ray-tracer_b200.api")
synth = importlib.import_module("relativistic-ray-tracer_b200.synth")


def prim_boxes(s):
    lo = np.empty((s.n_prims, 3)); hi = np.empty((s.n_prims, 3))
    for p in range(s.n_prims):
        i = s.prim_index[p]
        if s.prim_kind[p] == 0:
            q = s.tri_pos[i].reshape(3, 3); lo[p] = q.min(0); hi[p] = q.max(0)
        else:
            c = s.sph[i]; lo[p] = c[:3] - c[3]; hi[p] = c[:3] + c[3]
    return lo, hi


def levelwise_order(lo, hi):
    n = len(lo)
    cen = 0.5 * (lo + hi)
    order = np.arange(n); sstart = np.zeros(n, np.int64); send = np.full(n, n, np.int64)
    nodes, leaves, levels = 1, (1 if n <= 4 else 0), 0
    while True:
        act = send - sstart > 4
        if not act.any():
            break
        levels += 1
        starts = np.unique(sstart[act])
        # k_ref_accum: box of every active segment
        blo = np.minimum.reduceat(lo[order], starts.tolist() if len(starts) else [0], axis=0)
        bhi = np.maximum.reduceat(hi[order], starts.tolist(), axis=0)
        # reduceat runs to the next start: inactive segments in between would pollute -> recompute per segment where needed
        seg_of = {int(s0): k for k, s0 in enumerate(starts)}
        for k, s0 in enumerate(starts):
            e0 = int(send[s0])
            nxt = int(starts[k + 1]) if k + 1 < len(starts) else n
            if nxt != e0:
                blo[k] = lo[order[s0:e0]].min(0); bhi[k] = hi[order[s0:e0]].max(0)
        # k_ref_flag
        is_left = np.zeros(n + 1, np.int64)
        idx = np.nonzero(act)[0]
        k = np.array([seg_of[int(x)] for x in sstart[idx]])
        ext = bhi[k] - blo[k]
        axis = np.where((ext[:, 0] > ext[:, 1]) & (ext[:, 0] > ext[:, 2]), 0, np.where((ext[:, 1] > ext[:, 0]) & (ext[:, 1] > ext[:, 2]), 1, 2))
        mid = 0.5 * (blo[k, axis] + bhi[k, axis])
        is_left[idx] = cen[order[idx], axis] < mid
        pre = np.concatenate([[0], np.cumsum(is_left)[:-1]])          # exclusive scan over n + 1 entries
        # k_ref_scatter
        s0 = sstart; e0 = send; cnt = e0 - s0
        nl = pre[e0] - pre[s0]
        halve = act & ((nl == 0) | (nl == cnt))
        L = pre[:n] - pre[s0]
        pos = np.where(is_left[:n] == 1, s0 + L, s0 + nl + (np.arange(n) - s0 - L))
        pos = np.where(act & ~halve, pos, np.arange(n))
        nl = np.where(halve, cnt // 2, nl)
        left = pos < s0 + nl
        ns = np.where(act, np.where(left, s0, s0 + nl), s0); ne = np.where(act, np.where(left, s0 + nl, e0), e0)
        new_order = np.empty_like(order); new_s = np.empty_like(sstart); new_e = np.empty_like(send)
        new_order[pos] = order; new_s[pos] = ns; new_e[pos] = ne
        first = act & (pos == ns)
        nodes += int(first.sum()); leaves += int((first & (ne - ns <= 4)).sum())
        order, sstart, send = new_order, new_s, new_e
    rank = np.empty(n, np.int64); rank[order] = np.arange(n)
    return rank, nodes, leaves, levels


def soup(base, tri):
    tri = np.asarray(tri, np.float64).reshape(-1, 9)
    nt = base.tri_pos.shape[0] + len(tri)
    z = np.zeros_like(tri)
    return sc.SceneData(np.zeros(nt, np.uint8), np.arange(nt, dtype=np.uint32), np.concatenate([base.tri_pos, tri]), np.concatenate([base.tri_nrm, z]),
                        np.concatenate([base.tri_mat, np.zeros(len(tri), np.int32)]), np.zeros((0, 4)), np.zeros(0, np.int32), base.mats, base.lights, base.camera)


@pytest.mark.parametrize("case", ["cbspheres", "cbgems", "synth_mesh", "identical", "two_clusters"])
def test_levelwise_partition_equals_recursive_construct_bvh(case):
    base = load_scene("cbspheres_lambertian_96x72")
    rng = np.random.default_rng(3)
    if case == "cbspheres":
        s = base
    elif case == "cbgems":
        s = load_scene("cbgems_480x360")
    elif case == "synth_mesh":
        s = synth.c5_scene(base, 40, 41)
    elif case == "identical":
        s = soup(base, np.tile(np.array([[-0.2, 0.3, 0.0, 0.2, 0.3, 0.0, 0.0, 0.6, 0.1]]), (700, 1)))
    else:
        c = np.concatenate([rng.normal([-0.6, 0.3, -0.5], 1e-3, (400, 3)), rng.normal([0.6, 1.2, 0.5], 1e-3, (400, 3))])
        s = soup(base, np.concatenate([c, c + rng.normal(0, 2e-3, c.shape), c + rng.normal(0, 2e-3, c.shape)], 1))
    host = api.bvh_build_host(s, 0)
    rank, nodes, leaves, levels = levelwise_order(*prim_boxes(s))
    assert np.array_equal(rank, host["dfs_rank"].astype(np.int64))
    assert (nodes, leaves, levels) == (host["n_nodes"], host["n_leaves"], host["depth"])
