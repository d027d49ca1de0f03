"""Pins the oracle's estimator (raytrace_pixel and below) against images rendered by the reference's own
PathTracer::raytrace_pixel with its rand() fed from the same per-path Philox stream (tests/golden/render_*.npz).
Both sides consume identical uniforms, so images must agree to float rounding and sample counts exactly."""
import numpy as np
import pytest

from util import golden, load_scene, sc
from oracle import orc

RENDERS = ["c1_flat_lambertian", "c2_bh_lambertian", "bh_glass_mirror", "flat_glass_mirror", "flat_microfacet", "bh_microfacet",
           "flat_gems", "flat_bunny_microfacet", "bh_bunny_microfacet", "flat_adaptive", "bh_adaptive", "flat_depth0",
           "flat_depth1", "bh_depth1_spp1", "flat_pointlight", "flat_hemisphere", "bh_alt_hole", "flat_envmap", "bh_envmap", "bh_envmap_gems",
           "flat_thin_lens", "bh_thin_lens", "flat_synth_mesh", "bh_synth_mesh", "flat_keenan_lights", "bh_keenan_lights"]


def setup_from_golden(g, obj):
    """apply the golden's black-hole / mode / frame to an Oracle or a product PathTracer-like object"""
    h, w = g["counts"].shape
    bh = g["bh"]
    mode = sc.MODE_FLAT if str(g["mode"]) == "flat" else sc.MODE_LITERAL
    if bh.size == 5:
        obj.set_blackhole(bh[:3], bh[3], bh[4], mode)
    else:
        obj.set_blackhole((0.0, 1.0, 0.0), 0.1, 0.1, mode)
    obj.set_frame(w, h)
    if "envmap" in g.files and bool(g["envmap"]):
        obj.set_envmap(golden("kat_envmap")["env"])          # the synthetic map the reference rendered with (-e)
    spp, l, m, batch, tol, seed = g["params"]
    return sc.make_params(ns_aa=spp, max_ray_depth=m, ns_area_light=l, samples_per_batch=batch, max_tolerance=tol,
                          direct_hemisphere_sample=bool(g["hemi"]), thin_lens=("lens" in g.files and bool(g["lens"])), seed=int(seed))


@pytest.mark.parametrize("name", RENDERS)
def test_oracle_render_matches_reference(name):
    g = golden("render_" + name)
    o = orc.Oracle(load_scene(str(g["scene"])))
    p = setup_from_golden(g, o)
    hdr, cnt, st = o.render(p)
    assert np.array_equal(cnt, g["counts"]), "sample counts (adaptive stop decisions) differ"
    ref = g["hdr"]
    assert np.isfinite(ref).all()
    # identical IEEE operations: expect equality; allow 2 float ulps for libm (glibc here vs glibc at golden time is the same)
    np.testing.assert_allclose(hdr, ref, rtol=3e-7, atol=1e-9)
    assert st["paths"] == int(cnt.sum())
