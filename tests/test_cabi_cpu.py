"""CPU-side checks of the product library: it loads, exports every symbol include/rrt_b200.h declares, refuses to
compute without a device (no CPU fallback), and its host-side BVH construction reproduces the reference's tree."""
import ctypes as C
import importlib
import os
import re

import numpy as np
import pytest

from util import ROOT, load_scene, sc
from oracle import orc

api = importlib.import_module("relativistic-ray-tracer_b200.api")


def declared_functions():
    src = open(os.path.join(ROOT, "include", "rrt_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rrt_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    L = api.lib()
    names = declared_functions()
    assert len(names) >= 24 and set(names) == set(api.EXPORTS)
    for n in names:
        assert hasattr(L, n), n
    assert L.rrt_abi_version() == 5


def test_struct_layouts_match_header():
    assert C.sizeof(sc.Material) == 36 and C.sizeof(sc.Light) == 120 and C.sizeof(sc.Hit) == 96
    assert C.sizeof(sc.Stats) == 11 * 8 and C.sizeof(sc.Params) == 48 and C.sizeof(sc.Camera) == 18 * 8


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    assert api.lib().rrt_device_count() == 0
    with pytest.raises(api.RrtError, match="no CPU fallback"):
        api.Context(0)


@pytest.mark.parametrize("scene", ["cbspheres_lambertian_480x360", "cbgems_480x360", "cbbunny_microfacet_480x360", "cbempty_480x360"])
def test_host_bvh_matches_reference_tree(scene):
    s = load_scene(scene)
    got = api.bvh_build_host(s, 0)
    ref = orc.Oracle(s).bvh_info()
    assert (got["n_nodes"], got["n_leaves"], got["depth"]) == (ref["n_nodes"], ref["n_leaves"], ref["depth"])
    assert np.array_equal(got["dfs_rank"], ref["dfs_rank"])
    assert got["flat_nodes"] == ref["n_nodes"] - ref["n_leaves"] and got["leaf_records"] == s.n_prims
    sah = api.bvh_build_host(s, 1)
    assert np.array_equal(sah["dfs_rank"], ref["dfs_rank"])  # tie-break ranks always come from the reference shape
    assert sah["leaf_records"] == s.n_prims


def _triangles_only(tri):
    from util import sc
    base = load_scene("cbempty_480x360")
    n = tri.shape[0]
    return sc.SceneData(np.zeros(n, np.uint8), np.arange(n, dtype=np.uint32), np.ascontiguousarray(tri, np.float64), np.tile([0, 0, 1.0], (n, 3)).reshape(n, 9),
                        np.zeros(n, np.int32), np.zeros((0, 4)), np.zeros(0, np.int32), base.mats[:1], base.lights, base.camera)


def test_host_bvh_survives_degenerate_scenes():
    """a scene is input, not code: the reference's midpoint split has no depth bound, and triangles whose boxes all contain one point give a chain as
    deep as the scene is large.  The reference recurses (and so did the builder here: 50 000 such triangles overflowed the stack); the builder now
    walks the same tree with an explicit stack, and the traversal tree (binned SAH with a depth budget) stays under the device's stack regardless."""
    n = 50000
    chain = np.zeros((n, 9)); chain[:, 0] = np.linspace(0, 1, n)       # (t, 0, 0), (0, 0, 0), (0, 0, 0)
    sah = api.bvh_build_host(_triangles_only(chain), 1)               # builds the reference-shape tree (ranks) and the traversal tree
    assert sah["depth"] > 40000 and sah["leaf_records"] == n and sorted(sah["dfs_rank"].tolist()) == list(range(n))
    assert sah["flat_depth"] < sah["stack"]
    rng = np.random.default_rng(3)
    for tri in (np.tile([0, 0, 0, 1, 0, 0, 0, 1, 0.0], (20000, 1)),                                                   # identical triangles
                np.where(rng.uniform(0, 1, (5000, 9)) < 0.01, np.nan, rng.uniform(-1, 1, (5000, 9))),                 # NaN vertices
                np.where(rng.uniform(0, 1, (5000, 9)) < 0.01, np.inf, rng.uniform(-1, 1, (5000, 9))),                 # infinite vertices
                rng.uniform(-1, 1, (3000, 9)) * 1e300, rng.uniform(-1, 1, (1, 9)), np.zeros((0, 9))):
        for topology in (0, 1):
            b = api.bvh_build_host(_triangles_only(tri), topology)
            assert b["leaf_records"] == tri.shape[0] and sorted(b["dfs_rank"].tolist()) == list(range(tri.shape[0]))


def test_survey_bvh_shapes():
    assert api.bvh_build_host(load_scene("cbspheres_lambertian_480x360"))["n_nodes"] == 7
    b = api.bvh_build_host(load_scene("cbbunny_microfacet_480x360"))
    assert (b["n_nodes"], b["n_leaves"], b["depth"]) == (19103, 9552, 19)


@pytest.mark.parametrize("scene", ["cbspheres_lambertian_480x360", "cbgems_480x360", "cbbunny_microfacet_480x360", "banana_480x360", "cbsynth_480x360"])
def test_dfs_rank_equals_the_reference_trees_own_leaf_order(scene):
    """tests/golden/kat_bvh_order.npz: BVHAccel::get_root() of the reference's OWN tree (src/bvh.cpp:49-96) walked left-then-right as intersect_micro
    walks it (src/bvh.cpp:120-135), primitives of a leaf in vector order.  The tie-break ranks (Triangle::intersect keeps the last primitive among
    equal t) of the host builder and of the oracle restatement equal that order, node / leaf counts and depth included."""
    from util import golden
    g = golden("kat_bvh_order")
    s = load_scene(scene)
    nodes, leaves, depth = [int(v) for v in g[scene + "_info"]]
    for topology in (0, 1):
        got = api.bvh_build_host(s, topology)
        assert (got["n_nodes"], got["n_leaves"], got["depth"]) == (nodes, leaves, depth)
        assert np.array_equal(got["dfs_rank"], g[scene + "_rank"])
    ref = orc.Oracle(s).bvh_info()
    assert np.array_equal(ref["dfs_rank"], g[scene + "_rank"]) and (ref["n_nodes"], ref["n_leaves"], ref["depth"]) == (nodes, leaves, depth)
    assert sorted(g[scene + "_rank"].tolist()) == list(range(s.n_prims))
