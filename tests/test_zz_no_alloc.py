"""rrt_trace_rays_static — the allocation-free variant of the BVH parity hook (include/rrt_b200.h; no counterpart in the
reference) — against the existing entry point rrt_trace_rays, through the
C-ABI.  Collected last on purpose (file name).

Tolerance: the tests of rrt_trace_rays assert bit-exact hit records; both entry points launch the same kernel, so the same
bar holds here: every field of every rrt_hit equal, bit for bit.  Two calls in a row check that the shared static array,
which keeps the previous call's contents, is cleared where it must start at zero (work cursor, counters)."""
import importlib

import numpy as np
import pytest

from util import golden, load_scene, sc

pytestmark = pytest.mark.gpu


def same_records(a, b):
    for f in a.dtype.names:
        x, y = a[f], b[f]
        if not np.all((x == y) | (np.isnan(x) & np.isnan(y)) if x.dtype.kind == "f" else (x == y)):
            return False
    return True


@pytest.fixture(scope="module")
def api():
    return importlib.import_module("relativistic-ray-tracer_b200.api")


def test_static_scratch_entry_point_matches_trace_rays_twice_in_a_row(api):
    scene = "cbbunny_microfacet_480x360"
    g = golden("kat_trace_" + scene)
    c = api.Context(0)
    c.upload_scene(load_scene(scene))
    c.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_LITERAL)
    rays = g["rays_bh"]
    o, d = np.ascontiguousarray(rays[:, 0:3]), np.ascontiguousarray(rays[:, 3:6])
    assert 0 < o.shape[0] <= api.STATIC_MAX_RAYS
    want, _ = c.trace_rays(o, d)
    assert (want["status"] == 1).any() and (want["status"] != 1).any()
    # first call: a different (smaller, reversed) batch, so the second call finds stale rays, hits, cursor and counters
    k = max(1, o.shape[0] // 3)
    first = c.trace_rays_static(o[:k][::-1], d[:k][::-1])
    assert same_records(first, want[:k][::-1])
    second = c.trace_rays_static(o, d)
    assert same_records(second, want)
    third = c.trace_rays_static(o, d)
    assert same_records(third, want)
    # shadow-ray form, and flat spacetime with caller ranges (min_t / max_t are staged in the same array)
    assert np.array_equal(c.trace_rays_static(o, d, any_hit=True)["status"], c.trace_rays(o, d, any_hit=True)[0]["status"])
    c.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_FLAT)
    rays = g["rays_flat"]
    o, d, mn, mx = rays[:, 0:3], rays[:, 3:6], rays[:, 6], rays[:, 7]
    flat = c.trace_rays(o, d, mn, mx)[0]
    assert (flat["status"] == 1).any() and same_records(c.trace_rays_static(o, d, mn, mx), flat)


def test_static_scratch_entry_point_rejects_sizes_beyond_its_maximum(api):
    c = api.Context(0)
    c.upload_scene(load_scene("cbspheres_lambertian_480x360"))
    n = api.STATIC_MAX_RAYS + 1
    o = np.zeros((n, 3)); d = np.tile([0.0, 0.0, -1.0], (n, 1))
    with pytest.raises(api.RrtError, match="static scratch limit of 65536"):
        c.trace_rays_static(o, d)
    # exactly the maximum is served, and equals the allocating entry point
    o = np.random.default_rng(5).uniform(-0.9, 0.9, (api.STATIC_MAX_RAYS, 3)) + [0, 1, 3]
    d = np.tile([0.0, 0.0, -1.0], (api.STATIC_MAX_RAYS, 1))
    assert same_records(c.trace_rays_static(o, d), c.trace_rays(o, d)[0])
    assert c.trace_rays_static(o[:0], d[:0]).shape == (0,)
