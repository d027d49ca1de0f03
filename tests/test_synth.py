"""Procedural C5 scene (relativistic-ray-tracer_b200/synth.py): host-side checks, CPU only."""
import importlib

import numpy as np
import pytest

from util import load_scene

synth = importlib.import_module("relativistic-ray-tracer_b200.synth")
api = importlib.import_module("relativistic-ray-tracer_b200.api")


def test_displaced_sphere_is_closed_outward_and_deterministic():
    v, f = synth.displaced_sphere(48, 37, seed=184)
    assert len(f) == 2 * 48 * 36 and f.min() == 0 and f.max() == len(v) - 1
    # closed 2-manifold: every directed edge appears once and its reverse once
    e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])
    key = e[:, 0] * len(v) + e[:, 1]; rkey = e[:, 1] * len(v) + e[:, 0]
    assert len(np.unique(key)) == len(key) and np.array_equal(np.sort(key), np.sort(rkey))
    fn = np.cross(v[f[:, 1]] - v[f[:, 0]], v[f[:, 2]] - v[f[:, 0]])
    assert ((fn * (v[f].mean(1) - [0, 0.5, 0])).sum(1) > 0).all()
    r = np.linalg.norm(v - [0, 0.5, 0], axis=1)
    assert 0.39 < r.min() and r.max() < 0.51 and r.std() > 1e-3
    v2, f2 = synth.displaced_sphere(48, 37, seed=184)
    assert np.array_equal(v, v2) and np.array_equal(f, f2)
    v3, _ = synth.displaced_sphere(48, 37, seed=185)
    assert not np.array_equal(v, v3)
    with pytest.raises(ValueError):
        synth.displaced_sphere(2, 5)


def test_default_triangle_count_is_the_c5_figure():
    # 2 * n_u * (n_v - 1) without building the mesh
    assert 2 * 708 * 708 == 1002528


def test_scene_assembly_and_host_bvh():
    base = load_scene("cbspheres_lambertian_480x360")
    s = synth.c5_scene(base, 40, 41)
    assert s.sph.shape[0] == 0 and s.tri_pos.shape[0] == 12 + 2 * 40 * 40 and s.n_prims == s.tri_pos.shape[0]
    assert np.array_equal(s.tri_pos[:12], base.tri_pos) and len(s.lights) == len(base.lights)
    n = s.tri_nrm[12:].reshape(-1, 3)
    np.testing.assert_allclose(np.linalg.norm(n, axis=1), 1.0, atol=1e-12)
    info = api.bvh_build_host(s, 1)
    assert info["leaf_records"] == s.n_prims and 0 < info["flat_depth"] < info["stack"]
    assert sorted(info["dfs_rank"].tolist()) == list(range(s.n_prims))


def test_sah_depth_budget_on_a_degenerate_distribution():
    """primitives spaced geometrically along a line make binned SAH peel one off per level; the depth budget has to
    keep the flattened tree below the device stack"""
    base = load_scene("cbspheres_lambertian_480x360")
    n = 3000
    x = 2.0 ** (-np.arange(n) / 8.0)
    tri = np.zeros((n, 9)); tri[:, 0] = x; tri[:, 3] = x * 1.0001; tri[:, 4] = 1e-6 * x; tri[:, 6] = x; tri[:, 8] = 1e-6 * x
    sc_mod = importlib.import_module("relativistic-ray-tracer_b200.scene")
    s = sc_mod.SceneData(np.zeros(n, np.uint8), np.arange(n, dtype=np.uint32), tri, np.tile([0.0, 0, 1] * 3, (n, 1)), np.zeros(n, np.int32),
                         np.zeros((0, 4)), np.zeros(0, np.int32), base.mats, base.lights, base.camera)
    info = api.bvh_build_host(s, 1)
    assert info["flat_depth"] < info["stack"], info["flat_depth"]
