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
