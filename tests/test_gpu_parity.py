"""GPU parity tests proper: the CUDA path (through the C-ABI, csrc/librrt_b200.so) against
  (a) the golden vectors produced by the reference's own objects (tests/golden/, made by make_golden.py), and
  (b) the CPU oracle (oracle/rrt_oracle.c) on the same seeded inputs.
Tolerances (BASELINE.json north_star): BVH hit / miss / primitive id bit-exact, t within 1e-5 relative; geodesic
endpoints and directions within 1e-4 relative; images RMSE <= 1 % of mean radiance.  What this implementation
actually delivers is stricter and is what the asserts state: chords, hit points and normals are BIT-identical
(IEEE double, same operation order, no FMA), BSDF / light values agree to libm rounding.
"""
import numpy as np
import pytest

from util import RAND_MAX, golden, hits_from_probe, load_scene, sc
from test_oracle_render import FULLSIZE, RENDERS, setup_from_golden, setup_full, stat_blocks

pytestmark = pytest.mark.gpu

import importlib
shard = importlib.import_module("relativistic-ray-tracer_b200.shard")


def same(a, b):
    a = np.asarray(a); b = np.asarray(b)
    return bool(np.all((a == b) | (np.isnan(a) & np.isnan(b))))


@pytest.fixture(scope="module")
def api():
    import importlib
    return importlib.import_module("relativistic-ray-tracer_b200.api")


@pytest.fixture(scope="module")
def orc():
    from oracle import orc as o
    return o


def set_opt(api, c, opt, value):
    """set_option; the vote-scheduled kernel (variant 1) and 4-wide nodes exist only in A/B builds of the library (make EXTRA=-DRRT_AB=1)"""
    try:
        c.set_option(opt, value)
    except api.RrtError as e:
        if "A/B builds" in str(e):
            pytest.skip("variant compiled only into A/B builds")
        raise


def make_ctx(api, scene_name):
    c = api.Context(0)
    c.upload_scene(load_scene(scene_name))
    return c


def test_philox_known_answers(api):
    c = api.Context(0)
    assert [hex(x) for x in c.philox(0, [[0, 0, 0, 0]])[0]] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    assert [hex(x) for x in c.philox(0xFFFFFFFFFFFFFFFF, [[0xFFFFFFFF] * 4])[0]] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    assert [hex(x) for x in c.philox(0x299F31D0A4093822, [[0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344]])[0]] == \
        ["0xd16cfe09", "0x94fdcceb", "0x5001e420", "0x24126ea1"]


@pytest.mark.parametrize("tag", ["default", "alt", "r0"])
def test_geodesic_chords_bit_exact(api, tag):
    g = golden("kat_geodesic")
    c = make_ctx(api, "cbempty_480x360")
    bh = g["bh_" + tag]
    c.set_blackhole(bh[:3], bh[3], bh[4], sc.MODE_LITERAL)
    ref = g["seg_" + tag]
    got = c.geodesic(g["o"], g["d"], ref.shape[1])
    assert same(got, ref)  # north_star asks for 1e-4 relative; delivered: identical bits, NaN chords of radial rays included
    assert c.chords_per_ray() == (63 if bh[4] == 0.1 else 32)


@pytest.mark.parametrize("topology,leaf,variant,grid,width", [(0, 4, 1, 0, 2), (1, 2, 1, 32, 2), (1, 1, 0, 64, 4), (1, 4, 0, 0, 8), (1, 2, 0, 128, 2),
                                                              (1, 2, 0, 0, 8), (1, 7, 0, 0, 4), (0, 4, 0, 0, 2),
                                                              (0, 4, 2, 0, 2), (1, 2, 2, 32, 2), (1, 2, 2, 0, 8), (1, 2, 0, 64, 8)])
@pytest.mark.parametrize("scene", ["cbspheres_lambertian_480x360", "cbgems_480x360", "cbbunny_microfacet_480x360"])
@pytest.mark.parametrize("mode", ["flat", "bh", "alt"])
def test_bvh_intersect_matches_reference(api, scene, mode, topology, leaf, variant, grid, width):
    g = golden("kat_trace_" + scene)
    if "hit_" + mode not in g.files:
        pytest.skip("no alt-hole vectors for this scene")
    c = api.Context(0)
    c.set_option(api.OPT_BVH_TOPOLOGY, topology)
    c.set_option(api.OPT_LEAF_SIZE, leaf)
    set_opt(api, c, api.OPT_TRACE_KERNEL, variant)   # 0 = chord-lockstep (default); A/B builds: 1 = vote-scheduled, 2 = decoupled stepping
    c.set_option(api.OPT_CLEARANCE_GRID, grid)
    set_opt(api, c, api.OPT_BVH_WIDTH, width)
    c.upload_scene(load_scene(scene))
    if mode == "flat":
        c.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_FLAT)
    elif mode == "alt":
        bh = g["bh_alt"]; c.set_blackhole(bh[:3], bh[3], bh[4], sc.MODE_LITERAL)
    rays = g["rays_" + mode]; ref = hits_from_probe(g["hit_" + mode])
    got, st = c.trace_rays(rays[:, 0:3], rays[:, 3:6], rays[:, 6], rays[:, 7])
    assert np.array_equal(got["status"] == 1, ref["hit"])                    # hit / miss: exact
    h = ref["hit"]
    assert np.array_equal(got["prim"][h], ref["prim"][h])                    # primitive id: exact
    assert same(got["hit_p"][h], ref["hit_p"][h]) and same(got["w_out"][h], ref["w_out"][h]) and same(got["n"][h], ref["n"][h])
    anyh, _ = c.trace_rays(rays[:, 0:3], rays[:, 3:6], rays[:, 6], rays[:, 7], any_hit=True)
    assert np.array_equal(anyh["status"] == 1, ref["any"])
    assert st["rays"] == len(rays)


@pytest.mark.parametrize("scene,mode", [("cbspheres_lambertian_480x360", "bh"), ("cbbunny_microfacet_480x360", "flat"),
                                        ("cbbunny_microfacet_480x360", "bh")])
def test_trace_matches_oracle_records(api, orc, scene, mode):
    """full rrt_hit records (status incl. capture, chord index, step count, t) against the CPU oracle"""
    g = golden("kat_trace_" + scene)
    s = load_scene(scene)
    c = api.Context(0); c.upload_scene(s)
    o = orc.Oracle(s)
    m = sc.MODE_FLAT if mode == "flat" else sc.MODE_LITERAL
    c.set_blackhole((0, 1, 0), 0.1, 0.1, m); o.set_blackhole((0, 1, 0), 0.1, 0.1, m)
    rays = g["rays_" + mode]
    got, gst = c.trace_rays(rays[:, 0:3], rays[:, 3:6], rays[:, 6], rays[:, 7])
    ref, ost = o.trace_rays(rays[:, 0:3], rays[:, 3:6], rays[:, 6], rays[:, 7])
    for f in ("status", "prim", "chord", "steps"):
        assert np.array_equal(got[f], ref[f]), f
    assert same(got["t"], ref["t"])  # north_star: 1e-5 relative; delivered: identical
    assert gst["geodesic_steps"] == ost["geodesic_steps"] and gst["rays"] == ost["rays"]
    assert (got["status"] == 2).sum() > 0 or mode == "flat"
    # without the 8-wide record for the top of the tree (RRT_OPT_TOP_NODE = 0: start at the binary root): the same records, more node visits
    c0 = api.Context(0); c0.set_option(api.OPT_TOP_NODE, 0); c0.upload_scene(s); c0.set_blackhole((0, 1, 0), 0.1, 0.1, m)
    g0, st0 = c0.trace_rays(rays[:, 0:3], rays[:, 3:6], rays[:, 6], rays[:, 7])
    assert all(np.array_equal(g0[f], got[f]) for f in ("status", "prim", "chord", "steps")) and same(g0["t"], got["t"]) and same(g0["hit_p"], got["hit_p"])
    assert st0["tri_tests"] > 0 and (scene.startswith("cbspheres") or st0["node_visits"] != gst["node_visits"])
    try:
        c.set_option(api.OPT_ESCAPE, 1)      # A/B builds: with the escape hand-off, the same records
    except api.RrtError as e:
        assert "A/B builds" in str(e)
        return
    on, ost2 = c.trace_rays(rays[:, 0:3], rays[:, 3:6], rays[:, 6], rays[:, 7])
    assert all(np.array_equal(on[f], got[f]) for f in ("status", "prim", "chord", "steps")) and same(on["t"], got["t"]) and same(on["hit_p"], got["hit_p"])
    assert ost2["geodesic_steps"] == gst["geodesic_steps"]


@pytest.mark.parametrize("builder", [2, 1])   # host binned-SAH tree, device LBVH
def test_grazing_rays_against_the_literal_double_traversal(api, orc, builder):
    """DESIGN.md section 2 names one possible deviation from bit-exact hit / miss: the FP32 slab test is conservative, so it can visit a leaf
    the reference's double slab test (src/bbox.cpp:10-25) culled, and the exact primitive test then reports a hit the reference's traversal
    never reaches.  1.2 M rays aimed at exactly those places (util.grazing_rays) against the oracle's literal double traversal of the
    reference tree, counted per family:
      B through points of triangle edges, D from origins exactly on a vertex (generic directions): the mismatch count must be 0.
      A through triangle vertices (generic directions): the hit point is a corner of the leaf box and of every ancestor box the vertex is
        extreme in, and the reference's six rounded quotients can come out as tmin > tmax by an ulp there: it skips that leaf and reports the
        hit on another triangle of the vertex fan or a farther one.  Measured on B200: 817 of 240 000 such rays (0.34 %).
      C inside the axis-aligned plane of a vertex coordinate (origin coordinate == a box plane, that direction component == 0 exactly) and
        E exactly axis-parallel rays through a vertex: the reference's slab quotient is 0 / 0 there, and std::max(NaN, t) / std::min(t, NaN)
        (argument order, src/bbox.cpp:19-20) make it reject that node — the reference then reports a FARTHER hit (or none) than the primitive
        the ray geometrically meets first.  The library returns the closest exact hit.  These rays cannot come out of the estimator (a
        direction component is exactly 0 only for a hand-made ray); the count is measured and stated, and every such ray must differ in
        exactly that way: library hit, at a smaller t than the reference's."""
    from util import grazing_rays
    s = load_scene("cbbunny_microfacet_480x360")
    c = api.Context(0); c.set_option(api.OPT_BVH_BUILDER, builder); c.upload_scene(s)
    c.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_FLAT)
    o = orc.Oracle(s); o.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_FLAT)
    n = 1200000
    ro, rd = grazing_rays(s, n, 77)
    got, _ = c.trace_rays(ro, rd)
    ref, _ = o.trace_rays(ro, rd)
    assert (ref["status"] == 1).mean() > 0.5            # the rays do find the mesh
    diff = (got["status"] != ref["status"]) | (got["prim"] != ref["prim"])
    agree = (got["status"] == 1) & ~diff
    diff |= agree & ((got["t"] != ref["t"]) | (got["hit_p"] != ref["hit_p"]).any(1))
    m = n // 5
    fam = {k: diff[i * m:(i + 1) * m if i < 4 else n] for i, k in enumerate("ABCDE")}
    print("[grazing rays, builder %d] mismatches per family: %s of %d each" % (builder, {k: int(v.sum()) for k, v in fam.items()}, m))
    for k in "BD":
        assert int(fam[k].sum()) == 0, "family %s: %d rays differ (first %s)" % (k, fam[k].sum(), np.flatnonzero(fam[k])[:5])
    d = np.flatnonzero(diff)
    assert (got["status"][d] == 1).all()                                          # the library always has the hit ...
    far = ref["status"][d] == 1
    assert (got["t"][d][far] <= ref["t"][d][far]).all()                           # ... and never a farther one than the reference
    tie = far & (got["t"][d] == ref["t"][d])
    print("    of the %d differing rays the reference misses %d, reports a farther hit for %d and another primitive at the same t for %d"
          % (len(d), (~far).sum(), (far & ~tie).sum(), tie.sum()))
    zero_comp = (rd[d] == 0).sum(1) >= 1
    in_a = d < m
    assert (zero_comp | in_a).all()                  # outside family A only rays with an exactly-zero direction component differ
    assert fam["A"].mean() < 0.01 and diff.mean() < 0.05


@pytest.mark.parametrize("variant,width", [(0, 2), (1, 2), (0, 8), (0, 4), (2, 2), (2, 8)])
@pytest.mark.parametrize("mode", ["flat", "bh"])
def test_far_origin_chords_match_oracle(api, orc, mode, variant, width):
    """chords whose origin lies 1e2..1e7 units from the scene are box-tested from a shifted origin (rrt_trace.cuh boxctx_make):
    same records as the oracle, for rays aimed into the scene and for the escaped rays of a 1080p frame"""
    s = load_scene("cbbunny_microfacet_480x360")
    c = api.Context(0); set_opt(api, c, api.OPT_TRACE_KERNEL, variant); set_opt(api, c, api.OPT_BVH_WIDTH, width); c.upload_scene(s)
    o = orc.Oracle(s)
    m = sc.MODE_FLAT if mode == "flat" else sc.MODE_LITERAL
    c.set_blackhole((0, 1, 0), 0.1, 0.1, m); o.set_blackhole((0, 1, 0), 0.1, 0.1, m)
    rng = np.random.default_rng(11)
    n = 3000
    target = np.c_[rng.uniform(-1, 1, n), rng.uniform(0, 2, n), rng.uniform(-1, 1, n)]
    dirs = rng.normal(size=(n, 3)); dirs /= np.linalg.norm(dirs, axis=1)[:, None]
    dirs[: n // 10, rng.integers(0, 3)] = 0.0                                  # some axis-parallel ones
    dirs /= np.linalg.norm(dirs, axis=1)[:, None]
    dist = 10.0 ** rng.uniform(2, 7, n)
    org = target - dirs * dist[:, None]
    if mode == "bh":                                                           # plus the escaped rays of a wide frame
        s2 = load_scene("cbbunny_microfacet_480x360"); s2.set_screen_size(1920, 1080)
        c2 = api.Context(0); c2.upload_scene(s2)
        co, cd = c2.generate_rays(rng.uniform(0, 1, (3000, 2)))
        org = np.r_[org, co]; dirs = np.r_[dirs, cd]
    got, gst = c.trace_rays(org, dirs)
    ref, ost = o.trace_rays(org, dirs)
    for f in ("status", "prim", "chord", "steps"):
        assert np.array_equal(got[f], ref[f]), f
    assert same(got["t"], ref["t"])
    if mode == "flat":
        assert (got["status"] == 1).mean() > 0.5
        # the shifted boxes prune: a far ray costs about what a near one does, not the whole tree
        assert gst["node_visits"] < 400 * len(org)
    anyh, _ = c.trace_rays(org, dirs, any_hit=True)
    refany, _ = o.trace_rays(org, dirs, any_hit=True)
    assert np.array_equal(anyh["status"], refany["status"])


def test_camera_rays(api):
    g = golden("kat_trace_cbspheres_lambertian_480x360")
    c = make_ctx(api, "cbspheres_lambertian_480x360")
    o, d = c.generate_rays(g["xy"])
    assert same(o, g["cam"][:, 0:3])
    np.testing.assert_allclose(d, g["cam"][:, 3:6], rtol=0, atol=4e-16)  # tan() is CUDA libm, not glibc


@pytest.mark.parametrize("scene", ["cbspheres_lambertian_480x360", "cbspheres_480x360", "cbspheres_microfacet_480x360"])
def test_bsdf_f_and_sample_f(api, scene):
    g = golden("kat_bsdf_" + scene)
    s = load_scene(scene)
    c = api.Context(0); c.upload_scene(s)
    for mi in range(len(s.mats)):
        inp, ref = g["in_%d" % mi], g["out_%d" % mi]
        u = inp[:, 6:10] / RAND_MAX
        f, wi, pdf = c.bsdf_sample(mi, inp[:, 0:3], u)
        fe = c.bsdf_eval(mi, inp[:, 0:3], inp[:, 3:6])
        # float results: a few float ulps (double libm differences are rounded away except at float ties)
        np.testing.assert_allclose(fe, ref[:, 0:3], rtol=2e-6, atol=1e-30, err_msg="%s mat %d f" % (scene, mi))
        live = ref[:, 9] != 0
        assert np.array_equal(pdf != 0, live)
        np.testing.assert_allclose(f[live], ref[live, 3:6], rtol=2e-6, atol=1e-30, err_msg="%s mat %d sample_f" % (scene, mi))
        np.testing.assert_allclose(pdf[live], ref[live, 9], rtol=2e-6, err_msg="%s mat %d pdf" % (scene, mi))
        np.testing.assert_allclose(wi[live], ref[live, 6:9], rtol=0, atol=1e-13, err_msg="%s mat %d wi" % (scene, mi))


@pytest.mark.parametrize("scene", ["cbspheres_lambertian_480x360", "cbempty_480x360"])
def test_light_sample_L(api, scene):
    g = golden("kat_light_" + scene)
    c = make_ctx(api, scene)
    inp, ref = g["inp"], g["out"]
    rad, wi, dist, pdf = c.light_sample(0, inp[:, 0:3], inp[:, 3:5] / RAND_MAX)
    assert same(rad.astype(np.float64), ref[:, 0:3]) and same(wi, ref[:, 3:6])  # no transcendental on this path: exact
    assert same(dist.astype(np.float64), ref[:, 6]) and same(pdf.astype(np.float64), ref[:, 7])


def test_hemisphere_and_directional_light_sample_L(api):
    """InfiniteHemisphereLight (uniform hemisphere sample through the reference's fixed sampleToWorld) and DirectionalLight: vectors from
    the reference's own objects on dae/keenan/banana.dae"""
    g = golden("kat_light_banana_480x360")
    c = make_ctx(api, "banana_480x360")
    inp = g["inp"]
    for k, key in ((0, "out"), (1, "out_1")):
        ref = g[key]
        rad, wi, dist, pdf = c.light_sample(k, inp[:, 0:3], inp[:, 3:5] / RAND_MAX)
        assert same(rad.astype(np.float64), ref[:, 0:3]) and same(dist.astype(np.float64), ref[:, 6]) and same(pdf.astype(np.float64), ref[:, 7]), k
        # hemisphere: the reference's UniformHemisphereSampler3D goes through sinf / cosf (src/sampler.cpp:21-29): CUDA's float libm against
        # glibc's, a float ulp (1.2e-7 measured); directional: exact
        np.testing.assert_allclose(wi, ref[:, 3:6], rtol=0, atol=2.5e-7 if k == 0 else 0)
    assert same(c.light_sample(1, inp[:, 0:3], inp[:, 3:5] / RAND_MAX)[1], g["out_1"][:, 3:6])


def test_environment_light_matches_reference_vectors(api):
    """EnvironmentLight::sample_L / sample_dir (golden vectors made by the reference's own class, tests/golden/kat_envmap.npz)"""
    g = golden("kat_envmap")
    c = make_ctx(api, str(g["scene"]))
    c.set_envmap(g["env"])
    inp, ref = g["inp"], g["out"]
    rad, wi, dist, pdf = c.light_sample(int(g["light_index"]), inp[:, 0:3], inp[:, 3:5] / RAND_MAX)
    # same table cell for every draw (integer work: exact); direction / pdf go through CUDA's sin / cos instead of glibc's
    np.testing.assert_allclose(wi, ref[:, 3:6], rtol=0, atol=1e-15)
    np.testing.assert_allclose(pdf, ref[:, 7], rtol=3e-7)
    assert same(rad.astype(np.float64), ref[:, 0:3]) and np.isinf(dist).all()
    got = c.env_lookup(g["dirs"]).astype(np.float64)
    np.testing.assert_allclose(got, g["dir_out"], rtol=2e-5, atol=1e-6)       # acos / atan2 feed interpolation weights
    assert (np.abs(got - g["dir_out"]) > 1e-6 * np.maximum(1, g["dir_out"])).mean() < 0.01
    c.set_envmap(None)
    with pytest.raises(api.RrtError):
        c.env_lookup(g["dirs"][:1])


def test_tonemap_matches_oracle(api, orc):
    rng = np.random.default_rng(5)
    hdr = np.concatenate([rng.uniform(0, 2, (4000, 3)), [[0, 0, 0], [1, 1, 1], [0.5, 10, 1e-6]]]).astype(np.float32)
    got = api.Context(0).tonemap(hdr).view(np.uint8).reshape(-1, 4).astype(int)
    ref = orc.tonemap(hdr).view(np.uint8).reshape(-1, 4).astype(int)
    assert np.abs(got - ref).max() <= 1  # powf: CUDA vs glibc, at most one 8-bit level at a truncation boundary
    assert (got != ref).mean() < 0.01


def image_report(hdr, ref):
    rel = np.abs(hdr - ref) / np.maximum(np.abs(ref), 1e-3)
    pix = rel.max(-1)
    rmse = float(np.sqrt(np.mean((hdr.astype(np.float64) - ref) ** 2)))
    return pix, rmse


@pytest.mark.parametrize("name", ["flat_hemisphere", "flat_pointlight", "bh_envmap", "flat_microfacet", "bh_glass_mirror"])
def test_render_is_identical_with_direct_lighting_inside_k_shade(api, name):
    g = golden("render_" + name)
    s = load_scene(str(g["scene"]))
    out = []
    for split in (1, 0):
        c = api.Context(0); c.set_option(api.OPT_SHADE_SPLIT, split); c.upload_scene(s)
        hdr, cnt, st = c.render(setup_from_golden(g, c))
        out.append((hdr, cnt, st["rays"], st["geodesic_steps"]))
    assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1]) and out[0][2:] == out[1][2:]


@pytest.mark.parametrize("name", RENDERS)
def test_render_matches_reference_image(api, orc, name):
    """Same Philox stream as the golden image rendered by the reference's raytrace_pixel: images agree path by path.
    A path can only differ where a CUDA-libm direction differs from glibc's in the last bit AND that flips a hit;
    the asserts allow 0.5 % of pixels to do so and bound the whole-image error far below the 1 % RMSE target."""
    g = golden("render_" + name)
    s = load_scene(str(g["scene"]))
    c = api.Context(0); c.upload_scene(s)
    p = setup_from_golden(g, c)
    hdr, cnt, st = c.render(p)
    ref = g["hdr"]
    pix, rmse = image_report(hdr, ref)
    assert (cnt != g["counts"]).mean() <= 0.005, "adaptive stop decisions differ"
    assert (pix > 1e-4).mean() <= 0.005, "fraction of pixels off by more than 1e-4 relative: %g" % (pix > 1e-4).mean()
    # thin lens: sin / cos of the lens angle make EVERY camera ray differ from glibc's in the last bit (pinhole rays only share one tan),
    # so a path flips more often; at 4 spp on 96x72 one flipped path is already 1.3 % of the mean.  The agreeing pixels must still be exact.
    assert rmse <= (0.02 if "thin_lens" in name else 0.01) * float(ref.mean()) + 1e-12
    same = pix <= 1e-4
    assert float(np.sqrt(np.mean((hdr[same].astype(np.float64) - ref[same]) ** 2))) <= 1e-4 * float(ref.mean())
    assert st["paths"] == int(cnt.sum())
    # the same work as the CPU oracle does on this input: ray and chord counts agree to the rare libm-flipped path
    o = orc.Oracle(s)
    po = setup_from_golden(g, o)
    _, _, ost = o.render(po)
    assert abs(st["rays"] - ost["rays"]) <= 1e-3 * ost["rays"] + 2
    assert abs(st["geodesic_steps"] - ost["geodesic_steps"]) <= 1e-3 * ost["geodesic_steps"] + 64


def c2_setup(api, w=480, h=360):
    s = load_scene("cbspheres_lambertian_480x360")
    if (w, h) != (480, 360):
        s.set_screen_size(w, h)
    c = api.Context(0); c.upload_scene(s)
    c.set_frame(w, h)
    return c, s


@pytest.mark.parametrize("name", FULLSIZE)
def test_full_size_frames_match_the_reference(api, name):
    """BASELINE.json's own configs at their real frame sizes against frames rendered by the reference's raytrace_pixel (Philox-fed rand()):
    C1 and C2 whole (480x360, 16 spp), a 256x144 cell of C3's 1280x720 / 256 spp frame and of the 1080p hole frame (64 spp, adaptive).  Every
    path consumes the same uniforms on both sides, so pixels agree to float rounding except where a last-bit libm difference (CUDA vs glibc
    sin / cos / pow ...) flips a hit or a roulette decision; with 37 k - 173 k pixels per fixture the share of such pixels is a measurement,
    not an artefact of a small frame: it is bounded at 0.5 % and reported."""
    g = golden("render_full_" + name)
    c = make_ctx(api, str(g["scene"]))
    p, (x0, y0, cw, ch) = setup_full(g, c)
    hdr, cnt, st = c.render(p)
    got, gotc = hdr[y0:y0 + ch, x0:x0 + cw], cnt[y0:y0 + ch, x0:x0 + cw]
    ref, refc = g["hdr"], g["counts"]
    assert cnt.sum() == gotc.sum() and st["paths"] == int(cnt.sum())
    rel = np.abs(got - ref) / np.maximum(np.abs(ref), 1e-3)
    bad = rel.max(-1) > 1e-4
    badc = gotc != refc
    rmse = float(np.sqrt(np.mean((got - ref) ** 2)))
    print("[full-size %s] %d of %d pixels differ by > 1e-4 (%.3f %%), %d sample counts differ, rmse / mean %.2e"
          % (name, bad.sum(), bad.size, 100.0 * bad.mean(), badc.sum(), rmse / ref.mean()))
    assert bad.mean() <= 0.005 and badc.mean() <= 0.002
    assert rmse <= 0.01 * ref.mean()                                     # north_star: RMSE <= 1 % of mean radiance
    assert np.median(rel.max(-1)) <= 2e-6


def test_converged_image_against_the_reference_on_an_independent_rng(api):
    """north_star's third check: converged images within 1 % of mean radiance of the reference's.  The reference image comes from the
    reference's own raytrace_pixel driven by glibc's generator (random_r; tests/golden/render_independent_rng.npz: 96x72, 4096 spp, hole on,
    adaptive stop off) — nothing in it depends on Philox or on the order the product consumes its draws, so a mistake made identically on
    both sides of the path-by-path goldens would show here as bias.  The library renders 16 384 spp.  The estimator is heavy-tailed (the
    light seen through the lens), so single pixels of a 4096-spp reference are not converged; what is converged, and asserted, is the
    16x12-pixel block mean: RMSE over the 36 blocks <= 1 % of the frame's mean radiance, frame mean within 0.5 %, and per-pixel differences
    no larger than the reference's own sampling noise predicts."""
    g = golden("render_independent_rng")
    ref = g["hdr"].astype(np.float64)
    spp, l, m = int(g["params"][0]), int(g["params"][1]), int(g["params"][2])
    c = make_ctx(api, str(g["scene"]))
    c.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_LITERAL); c.set_frame(96, 72)
    n = 16384
    acc = np.zeros((72, 96, 3)); acc2 = np.zeros((72, 96, 3))
    k = 8                                      # 8 independent renders of n / 8 spp: their spread estimates the library's own noise
    for i in range(k):
        hdr, cnt, st = c.render(sc.make_params(ns_aa=n // k, max_ray_depth=m, ns_area_light=l, samples_per_batch=n, max_tolerance=0.0, seed=1000 + i))
        assert (cnt == n // k).all()
        acc += hdr; acc2 += hdr.astype(np.float64) ** 2
    got = acc / k
    var_got = np.maximum(acc2 / k - got ** 2, 0) / (k - 1)              # variance of the 16 k-spp mean, per pixel and channel
    M = ref.mean()
    d = stat_blocks(got) - stat_blocks(ref)
    rmse_blocks = float(np.sqrt((d ** 2).mean()))
    print("[independent rng] frame mean %.6f vs %.6f (%.3f %%), block rmse / mean %.4f, pixel rmse / mean %.3f"
          % (got.mean(), M, 100 * (got.mean() / M - 1), rmse_blocks / M, np.sqrt(((got - ref) ** 2).mean()) / M))
    assert abs(got.mean() / M - 1) < 0.005
    assert rmse_blocks <= 0.01 * M
    # per pixel: the difference is explained by sampling noise (reference at 4096 spp has n / spp = 4x the library's variance)
    z = (got - ref) / np.sqrt(var_got * (1 + n / spp) + 1e-12)
    assert np.median(np.abs(z)) < 1.2 and (np.abs(z) > 6).mean() < 0.01


def test_shadow_rays_sorted_by_origin_cell_give_the_same_frame(api):
    """RRT_OPT_SORT_SHADOW (meshes, hole on): the shadow rays of a launch handed to the trace kernel cell by cell of their origin instead of
    in queue order.  Which lane traces a ray changes; frame, sample counts and every counter must not."""
    s = load_scene("cbbunny_microfacet_480x360")
    p = sc.make_params(ns_aa=40, max_ray_depth=4, ns_area_light=2, samples_per_batch=16, max_tolerance=0.2, seed=11)   # adaptive stop acts: 3 rounds
    out = {}
    for sort in (0, 1, 2):
        c = api.Context(0); c.set_option(api.OPT_SORT_SHADOW, sort); c.upload_scene(s); c.set_frame(240, 180)
        c.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_LITERAL)
        out[sort] = c.render(p)
        c.close()
    h0, c0, st0 = out[0]
    assert 0 < c0.min() < 40 and c0.max() == 40 and np.isfinite(h0).all() and h0.max() > 0
    for sort in (1, 2):
        h, cn, st = out[sort]
        assert np.array_equal(h, h0) and np.array_equal(cn, c0), sort
        assert all(st[k] == st0[k] for k in ("paths", "rays", "geodesic_steps", "node_visits", "tri_tests")), sort
        assert st["kernel_launches"] > st0["kernel_launches"] and st["trace_launches"] == st0["trace_launches"]
    # not for flat spacetime (a flat shadow ray is one traversal: nothing to repay a sort) and not for the 8-wide kernel of small scenes
    c = api.Context(0); c.set_option(api.OPT_SORT_SHADOW, 2); c.upload_scene(s); c.set_frame(240, 180); c.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_FLAT)
    cf = api.Context(0); cf.set_option(api.OPT_SORT_SHADOW, 0); cf.upload_scene(s); cf.set_frame(240, 180); cf.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_FLAT)
    assert c.render(p)[2]["kernel_launches"] == cf.render(p)[2]["kernel_launches"]


def test_c2_full_size_properties(api):
    """BASELINE config 2 at full size: size-independent properties instead of a CPU run."""
    c, s = c2_setup(api)
    p = sc.make_params(ns_aa=16, max_ray_depth=5, ns_area_light=4, seed=7)
    hdr, cnt, st = c.render(p)
    assert np.isfinite(hdr).all() and (hdr >= 0).all()
    assert (cnt == 16).all() and st["paths"] == 480 * 360 * 16          # spp < batch: no adaptive stop
    assert st["rays"] > st["paths"] and st["geodesic_steps"] > st["rays"]
    hdr2, cnt2, st2 = c.render(p)
    assert np.array_equal(hdr, hdr2) and st2["rays"] == st["rays"] and st2["geodesic_steps"] == st["geodesic_steps"]  # deterministic
    # the other trace-kernel variant and the other BVH shape must not change a single bit either
    for variant in (1, 2):                                                 # A/B builds only: vote-scheduled loop, decoupled stepping
        try:
            c.set_option(api.OPT_TRACE_KERNEL, variant)
        except api.RrtError:
            continue
        hv, _, stv = c.render(p)
        c.set_option(api.OPT_TRACE_KERNEL, 0)
        assert np.array_equal(hdr, hv) and stv["rays"] == st["rays"] and stv["geodesic_steps"] == st["geodesic_steps"]
    c0 = api.Context(0); c0.set_option(api.OPT_BVH_TOPOLOGY, 0); c0.set_option(api.OPT_CLEARANCE_GRID, 96)   # reference tree shape + empty-space grid
    c0.upload_scene(s); c0.set_frame(480, 360)
    h0, _, st0 = c0.render(p)
    assert np.array_equal(hdr, h0) and st0["rays"] == st["rays"]
    c.set_option(api.OPT_SHADE_SPLIT, 1)                                   # direct lighting as its own kernel instead of inside k_shade
    hs, _, sts = c.render(p)
    c.set_option(api.OPT_SHADE_SPLIT, 2)
    assert np.array_equal(hdr, hs) and sts["rays"] == st["rays"] and sts["kernel_launches"] > st["kernel_launches"]
    for width in (2,):                                                     # default is 8 children per node
        cw = api.Context(0); cw.set_option(api.OPT_BVH_WIDTH, width); cw.upload_scene(s); cw.set_frame(480, 360)
        hw, _, stw = cw.render(p)
        assert np.array_equal(hdr, hw) and stw["rays"] == st["rays"] and stw["geodesic_steps"] == st["geodesic_steps"]
    # a different chunking of the wavefront must not change a single bit
    p2 = sc.make_params(ns_aa=16, max_ray_depth=5, ns_area_light=4, seed=7, max_paths_in_flight=300000)
    hdr3, _, st3 = c.render(p2)
    assert np.array_equal(hdr, hdr3) and st3["rays"] == st["rays"]
    # escape hand-off on (rays outside the scene box carried on by a lean kernel): same image, same ray / chord / primitive-test counts
    try:
        c.set_option(api.OPT_ESCAPE, 1)       # A/B builds only
        he, ce, ste = c.render(p)
        c.set_option(api.OPT_ESCAPE, 0)
        assert np.array_equal(hdr, he) and np.array_equal(cnt, ce)
        assert all(ste[k] == st[k] for k in ("rays", "geodesic_steps", "tri_tests", "sphere_tests", "paths"))
    except api.RrtError as e:
        assert "A/B builds" in str(e)
    c.set_option(api.OPT_OVERLAP, 0)          # shadow rays of depth i no longer beside the extension rays of depth i + 1
    ho, co, sto = c.render(p)
    c.set_option(api.OPT_OVERLAP, 1)
    assert np.array_equal(hdr, ho) and np.array_equal(cnt, co) and sto["rays"] == st["rays"]
    # the round's chunks dealt over 1, 3 or 4 streams instead of 2: not a bit may change
    for ns in (1, 3, 4):
        c.set_option(api.OPT_STREAMS, ns)
        hn, cn, stn = c.render(p2)
        assert np.array_equal(hdr, hn) and np.array_equal(cnt, cn) and stn["rays"] == st["rays"], ns
    c.set_option(api.OPT_STREAMS, 2)
    # interleaved tile shards: the union of the ranks' frames is the single-GPU frame (multi-GPU contract)
    acc = np.zeros_like(hdr); accn = np.zeros_like(cnt)
    for r in range(3):
        c.set_tile_shard(32, r, 3)
        hr, cr, _ = c.render(p)
        assert ((hr != 0).any(-1) <= (cr > 0)).all()
        assert np.array_equal(cr > 0, shard.tile_owner(480, 360, 32, 3) == r)   # device owner rule == host owner rule
        acc += hr; accn += cr
    c.set_tile_shard(32, 0, 1)
    assert np.array_equal(acc, hdr) and np.array_equal(accn, cnt)
    # cell render (-p): pixels outside the cell stay zero, inside equal the full frame
    c.set_cell(100, 50, 64, 40)
    hc, cc, _ = c.render(p)
    c.set_cell(0, 0, 0, 0)
    assert np.array_equal(hc[50:90, 100:164], hdr[50:90, 100:164]) and cc.sum() == 64 * 40 * 16
    hc[50:90, 100:164] = 0
    assert not hc.any()


def test_c1_flat_full_size_against_oracle_stats(api, orc):
    """BASELINE config 1 (flat space) at 480x360, 1 spp: per-pixel agreement with the CPU oracle."""
    c, s = c2_setup(api)
    o = orc.Oracle(s)
    c.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_FLAT); o.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_FLAT)
    o.set_frame(480, 360)
    p = sc.make_params(ns_aa=2, max_ray_depth=5, ns_area_light=4, seed=3)
    hdr, cnt, st = c.render(p)
    ref, rcnt, ost = o.render(p)
    pix, rmse = image_report(hdr, ref)
    assert np.array_equal(cnt, rcnt)
    assert (pix > 1e-4).mean() <= 0.002 and rmse <= 0.01 * float(ref.mean())
    assert abs(st["rays"] - ost["rays"]) <= 1e-3 * ost["rays"]


@pytest.mark.parametrize("mode", ["flat", "bh"])
def test_procedural_mesh_scene_matches_oracle(api, orc, mode):
    """C5's generator at a size the oracle finishes in seconds (18k triangles): hit records and a small render"""
    synth = importlib.import_module("relativistic-ray-tracer_b200.synth")
    s = synth.c5_scene(load_scene("cbspheres_lambertian_96x72"), 96, 97)
    c = api.Context(0); c.upload_scene(s); c.set_frame(96, 72)
    o = orc.Oracle(s); o.set_frame(96, 72)
    m = sc.MODE_FLAT if mode == "flat" else sc.MODE_LITERAL
    c.set_blackhole((0, 1, 0), 0.1, 0.1, m); o.set_blackhole((0, 1, 0), 0.1, 0.1, m)
    rng = np.random.default_rng(5)
    ro, rd = c.generate_rays(rng.uniform(0, 1, (4000, 2)))
    got, gst = c.trace_rays(ro, rd)
    ref, ost = o.trace_rays(ro, rd)
    for f in ("status", "prim", "chord", "steps"):
        assert np.array_equal(got[f], ref[f]), f
    assert same(got["t"], ref["t"]) and same(got["hit_p"], ref["hit_p"]) and same(got["n"], ref["n"])
    assert (got["prim"] >= 12).sum() > 100                     # the mesh is actually hit
    p = sc.make_params(ns_aa=2, max_ray_depth=3, ns_area_light=2, seed=9)
    hdr, cnt, st = c.render(p)
    rimg, rcnt, rst = o.render(p)
    pix, rmse = image_report(hdr, rimg)
    assert np.array_equal(cnt, rcnt) and (pix > 1e-4).mean() <= 0.005 and rmse <= 0.01 * float(rimg.mean())
    assert abs(st["rays"] - rst["rays"]) <= 2e-3 * rst["rays"]


def test_adaptive_rounds_match_oracle(api, orc):
    """spp > batch: several adaptive rounds, active-pixel compaction between them"""
    s = load_scene("cbspheres_lambertian_96x72")
    c = api.Context(0); c.upload_scene(s); c.set_frame(96, 72)
    o = orc.Oracle(s); o.set_frame(96, 72)
    p = sc.make_params(ns_aa=40, max_ray_depth=3, ns_area_light=1, samples_per_batch=8, max_tolerance=0.25, seed=21)
    hdr, cnt, st = c.render(p)
    ref, rcnt, _ = o.render(p)
    assert (cnt != rcnt).mean() <= 0.005
    assert set(np.unique(cnt)) <= {8, 16, 24, 32, 40} and len(np.unique(cnt)) >= 3
    pix, rmse = image_report(hdr, ref)
    assert (pix > 1e-4).mean() <= 0.01 and rmse <= 0.01 * float(ref.mean())


def test_empty_and_degenerate_inputs(api):
    c = make_ctx(api, "cbempty_480x360")
    hits, st = c.trace_rays(np.zeros((0, 3)), np.zeros((0, 3)))
    assert len(hits) == 0 and st["rays"] == 0
    # radial ray: NaN chords, never hits, walks all chords (reference src/static_scene/blackhole.cpp:24-27)
    hits, st = c.trace_rays([[0.0, 1.0, 3.0]], [[0.0, 0.0, -1.0]])
    assert hits["status"][0] == 0 and hits["steps"][0] == 63
    # a scene with no primitives at all
    e = sc.SceneData(np.zeros(0, np.uint8), np.zeros(0, np.uint32), np.zeros((0, 9)), np.zeros((0, 9)), np.zeros(0, np.int32),
                     np.zeros((0, 4)), np.zeros(0, np.int32), np.zeros(0, sc.MAT_DTYPE), np.zeros(0, sc.LIGHT_DTYPE), load_scene("cbempty_480x360").camera)
    c2 = api.Context(0); c2.upload_scene(e); c2.set_frame(16, 8)
    hdr, cnt, st = c2.render(sc.make_params(ns_aa=2, max_ray_depth=2))
    assert not hdr.any() and (cnt == 2).all()
    with pytest.raises(api.RrtError):
        api.Context(0).render(sc.make_params())  # nothing set: RRT_E_STATE, loudly
    # arguments that would hang or silently do nothing are refused: a NaN / non-positive / microscopic geodesic step makes the
    # reference's loop `j * dtheta < 2 pi` endless, undefined option ids, a cell without height
    for bad in (0.0, -0.1, float("nan"), float("inf"), 1e-9):
        with pytest.raises(api.RrtError, match="delta_theta"):
            c.set_blackhole((0, 1, 0), 0.1, bad, sc.MODE_LITERAL)
    for bad_r in (-0.1, float("nan")):
        with pytest.raises(api.RrtError, match="radius"):
            c.set_blackhole((0, 1, 0), bad_r, 0.1, sc.MODE_LITERAL)
    c.set_blackhole((0, 1, 0), 0.0, 0.1, sc.MODE_LITERAL)      # r = 0 is legal (the reference's -B 0 1 0 0 0.1)
    with pytest.raises(api.RrtError, match="unknown option"):
        c.set_option(15, 1)
    with pytest.raises(api.RrtError, match="positive height"):
        c.set_cell(0, 0, 16, 0)
    c.set_cell(0, 0, 0, 0)


def test_pathtracer_mirror_state_machine(api, tmp_path):
    s = load_scene("cbspheres_lambertian_96x72")
    pt = api.PathTracer(ns_aa=4, max_ray_depth=3, ns_area_light=2, seed=5)
    assert pt.state == pt.INIT
    pt.start_raytracing()          # wrong state: silently ignored, as the reference does
    assert pt.state == pt.INIT
    pt.set_scene(s); pt.set_camera(s)
    assert pt.state == pt.READY
    pt.set_frame_size(96, 72)
    out = tmp_path / "img.png"
    pt.render_to_file(str(out))
    assert pt.state == pt.DONE and out.exists() and (tmp_path / "img_rate.png").exists()
    assert pt.sampleBuffer.mean() > 0.05 and (pt.sampleCountBuffer == 4).all()
    assert pt.bvh.ctx.bvh_info()["n_nodes"] == 7


def test_thin_lens_rays_match_reference(api):
    """rrt_generate_rays_thin_lens vs the reference's Camera::generate_ray_for_thin_lens: origins exact (sin / cos of the lens
    angle enter the direction through CUDA's libm, so directions to 2 ulp)"""
    g = golden("kat_thin_lens")
    c = api.Context(0); c.upload_scene(load_scene(str(g["scene"])))
    o, d = c.generate_rays_thin_lens(g["inp"][:, :2], g["inp"][:, 2:])
    np.testing.assert_allclose(o, g["out"][:, :3], rtol=0, atol=3e-16)
    np.testing.assert_allclose(d, g["out"][:, 3:6], rtol=0, atol=4e-16)


def test_progress_callback_snapshots_and_stop(api):
    """rrt_set_progress / rrt_snapshot: called between adaptive batches; snapshots are the running estimate; a render
    with a callback equals one without; a truthy return abandons the remaining batches and leaves the partial estimate."""
    s = load_scene("cbspheres_lambertian_96x72")
    c = api.Context(0); c.upload_scene(s); c.set_frame(96, 72)
    p = sc.make_params(ns_aa=32, max_ray_depth=3, ns_area_light=1, samples_per_batch=8, max_tolerance=0.0, seed=3)   # tol 0: only black pixels stop early
    ref_hdr, ref_cnt, _ = c.render(p)
    seen = []
    def cb(done, total, active):
        hdr, cnt = c.snapshot()
        seen.append((done, total, active, hdr, cnt))
        return False
    c.set_progress(cb)
    hdr, cnt, st = c.render(p)
    c.set_progress(None)
    assert np.array_equal(hdr, ref_hdr) and np.array_equal(cnt, ref_cnt)
    assert [x[0] for x in seen] == [8, 16, 24] and all(x[1] == 32 for x in seen)
    for done, _, active, shdr, scnt in seen:
        assert scnt.max() == done and set(np.unique(scnt)) <= {8, done} and int((scnt == done).sum()) >= active > 0
        assert np.isfinite(shdr).all() and shdr.max() > 0
    # the 24-sample snapshot of a pixel that went all the way is the mean of its first 24 samples: close to, not equal to, the final mean
    full = seen[-1][4] == 24
    assert np.abs(seen[-1][3][full] - ref_hdr[full]).mean() < 0.05
    with pytest.raises(api.RrtError):
        c.snapshot()                                  # outside a callback
    # stop after the second batch
    calls = []
    c.set_progress(lambda done, total, active: calls.append(done) or done >= 16)
    hdr2, cnt2, st2 = c.render(p)
    c.set_progress(None)
    assert calls == [8, 16] and cnt2.max() == 16 and st2["paths"] == int(cnt2.sum())
    assert np.array_equal(cnt2, seen[1][4]) and np.array_equal(hdr2, seen[1][3])


# ---------------------------------------------------------------- FP32 stepper (RRT_OPT_STEPPER = 1, opt-in)
def _chord_errors(ref, got, c):
    end_r = ref[..., 0:3] + ref[..., 3:6] * ref[..., 6:7]; end_g = got[..., 0:3] + got[..., 3:6] * got[..., 6:7]
    with np.errstate(all="ignore"):
        perr = np.linalg.norm(end_r - end_g, axis=-1) / np.linalg.norm(end_r - c, axis=-1)
        derr = np.linalg.norm(ref[..., 3:6] - got[..., 3:6], axis=-1)
    captured_before = np.cumsum(ref[..., 7] > 0, axis=1) - (ref[..., 7] > 0) > 0
    return perr, derr, np.isfinite(perr) & np.isfinite(derr) & ~captured_before


@pytest.mark.parametrize("tag", ["default", "alt"])
def test_fp32_stepper_chords_within_tolerance(api, tag):
    """BASELINE.json: geodesic end points and directions within 1e-4 relative at the same step size.  Against the chords of
    the reference's own BlackHole::next_micro_ray (golden): directions always (< 1e-5), end points for 99 % of the chords
    (median ~1e-6); the remainder are far points of escaped rays, see tests/test_plane_recurrence.py."""
    g = golden("kat_geodesic")
    c = make_ctx(api, "cbempty_480x360")
    bh = g["bh_" + tag]
    c.set_blackhole(bh[:3], bh[3], bh[4], sc.MODE_LITERAL)
    c.set_option(api.OPT_STEPPER, 1)
    ref = g["seg_" + tag]
    got = c.geodesic(g["o"], g["d"], ref.shape[1])
    perr, derr, ok = _chord_errors(ref, got, bh[:3])
    assert ok.sum() > 0.5 * ok.size
    assert np.median(perr[ok]) < 2e-6 and np.percentile(perr[ok], 99) < 1e-4 and perr[ok].max() < 1e-3
    assert np.median(derr[ok]) < 1e-6 and derr[ok].max() < 1e-5
    assert np.array_equal((ref[..., 7] > 0)[ok], (got[..., 7] > 0)[ok])          # horizon verdicts
    c.set_option(api.OPT_STEPPER, 0)
    assert same(c.geodesic(g["o"], g["d"], ref.shape[1]), ref)                   # and the default is still bit-exact


@pytest.mark.parametrize("scene", ["cbspheres_lambertian_480x360", "cbbunny_microfacet_480x360", "cbgems_480x360"])
@pytest.mark.parametrize("width", [0, 2])
def test_fp32_stepper_hits_agree_with_reference(api, scene, width):
    """The reference's own hit records (camera, shadow-style and bounce rays, hole on) through the FP32 stepper: hit / miss
    and primitive id agree for all but rays that graze an edge (< 0.2 %); where they agree the hit point is within 1e-5
    (99 %; beyond 1e-4 for < 0.1 %)."""
    g = golden("kat_trace_" + scene)
    c = api.Context(0)
    c.set_option(api.OPT_STEPPER, 1); c.set_option(api.OPT_BVH_WIDTH, width)
    c.upload_scene(load_scene(scene))
    rays = g["rays_bh"]; ref = hits_from_probe(g["hit_bh"])
    got, st = c.trace_rays(rays[:, 0:3], rays[:, 3:6], rays[:, 6], rays[:, 7])
    hit = got["status"] == 1
    assert (hit != ref["hit"]).mean() < 2e-3
    both = hit & ref["hit"]
    same_prim = got["prim"][both] == ref["prim"][both]
    assert (~same_prim).mean() < 2e-3
    d = np.linalg.norm(got["hit_p"][both][same_prim] - ref["hit_p"][both][same_prim], axis=-1)
    assert np.percentile(d, 99) < 1e-5 and (d > 1e-4).mean() < 1e-3   # the odd ray meets the same primitive on another chord
    anyh, _ = c.trace_rays(rays[:, 0:3], rays[:, 3:6], rays[:, 6], rays[:, 7], any_hit=True)
    assert ((anyh["status"] == 1) != ref["any"]).mean() < 2e-3
    assert st["rays"] == len(rays)


@pytest.mark.parametrize("scene", ["cbspheres_lambertian_96x72", "cbspheres_96x72", "cbbunny_microfacet_96x72"])
def test_fp32_stepper_image_within_one_percent_rmse(api, scene):
    """BASELINE.json: converged images within 1 % RMSE of mean radiance.  Same seed, same Philox streams, 1024 spp without
    the adaptive stop: the FP32-stepper image against the bit-exact one (which the golden renders pin to the reference)."""
    s = load_scene(scene)
    p = sc.make_params(ns_aa=1024, max_ray_depth=5, ns_area_light=4, seed=11, samples_per_batch=1025)
    out = []
    for stepper in (0, 1):
        c = api.Context(0); c.set_option(api.OPT_STEPPER, stepper); c.upload_scene(s); c.set_frame(96, 72)
        hdr, cnt, st = c.render(p)
        out.append((hdr, st))
        c.close()
    (a, sa), (b, sb) = out
    mean = float(a.mean())
    assert np.sqrt(np.mean((a - b) ** 2)) <= 0.01 * mean
    assert abs(float(b.mean()) - mean) <= 2e-3 * mean
    assert abs(sb["rays"] - sa["rays"]) <= 1e-3 * sa["rays"] and abs(sb["geodesic_steps"] - sa["geodesic_steps"]) <= 2e-3 * sa["geodesic_steps"]
    c = api.Context(0); c.set_option(api.OPT_STEPPER, 1); c.upload_scene(s); c.set_frame(96, 72)
    hdr2, _, _ = c.render(p)
    assert np.array_equal(hdr2, b)                                               # the FP32 path is deterministic too


# ---------------------------------------------------------------- BVH built on the device (RRT_OPT_BVH_BUILDER = 1, opt-in)
def _device_built(api, s):
    c = api.Context(0); c.set_option(api.OPT_BVH_BUILDER, 1); c.upload_scene(s)
    return c


@pytest.mark.parametrize("scene", ["cbbunny_microfacet_480x360", "cbgems_480x360", "synth"])
def test_device_bvh_build_reproduces_reference_order(api, scene):
    """construct_bvh's depth-first primitive order (reference src/bvh.cpp:49-96), computed on the device by level-synchronous
    stable partition, equals the host restatement's — ranks, node / leaf counts and depth, integer for integer."""
    if scene == "synth":
        synth = importlib.import_module("relativistic-ray-tracer_b200.synth")
        s = synth.c5_scene(load_scene("cbspheres_lambertian_96x72"), 96, 97)
    else:
        s = load_scene(scene)
    host = api.bvh_build_host(s, 0)
    dev = _device_built(api, s).bvh_info()
    assert np.array_equal(dev["dfs_rank"], host["dfs_rank"])
    assert sorted(dev["dfs_rank"].tolist()) == list(range(s.n_prims))
    assert (dev["n_nodes"], dev["n_leaves"], dev["depth"]) == (host["n_nodes"], host["n_leaves"], host["depth"])
    g = golden("kat_bvh_order")      # the reference's own tree, walked (oracle/ref_probe.cpp bvhorder)
    if scene + "_rank" in g.files:
        assert np.array_equal(dev["dfs_rank"], g[scene + "_rank"])
        assert (dev["n_nodes"], dev["n_leaves"], dev["depth"]) == tuple(int(v) for v in g[scene + "_info"])


@pytest.mark.parametrize("scene", ["cbbunny_microfacet_480x360", "cbgems_480x360"])
@pytest.mark.parametrize("mode", ["flat", "bh"])
def test_device_bvh_build_hits_match_reference(api, scene, mode):
    """The LBVH walks to the same hits: the reference's own records, bit for bit, as for the host-built trees."""
    g = golden("kat_trace_" + scene)
    c = _device_built(api, load_scene(scene))
    if mode == "flat":
        c.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_FLAT)
    rays = g["rays_" + mode]; ref = hits_from_probe(g["hit_" + mode])
    got, st = c.trace_rays(rays[:, 0:3], rays[:, 3:6], rays[:, 6], rays[:, 7])
    assert np.array_equal(got["status"] == 1, ref["hit"])
    h = ref["hit"]
    assert np.array_equal(got["prim"][h], ref["prim"][h])
    assert same(got["hit_p"][h], ref["hit_p"][h]) and same(got["w_out"][h], ref["w_out"][h]) and same(got["n"][h], ref["n"][h])
    anyh, _ = c.trace_rays(rays[:, 0:3], rays[:, 3:6], rays[:, 6], rays[:, 7], any_hit=True)
    assert np.array_equal(anyh["status"] == 1, ref["any"])


def test_device_bvh_build_renders_the_same_image(api):
    synth = importlib.import_module("relativistic-ray-tracer_b200.synth")
    s = synth.c5_scene(load_scene("cbspheres_lambertian_96x72"), 96, 97)
    p = sc.make_params(ns_aa=4, max_ray_depth=4, ns_area_light=2, seed=2)
    imgs = []
    for builder in (2, 1):
        c = api.Context(0); c.set_option(api.OPT_BVH_BUILDER, builder); c.upload_scene(s); c.set_frame(96, 72)
        hdr, cnt, st = c.render(p)
        imgs.append((hdr, cnt, st["rays"], st["geodesic_steps"]))
    assert np.array_equal(imgs[0][0], imgs[1][0]) and np.array_equal(imgs[0][1], imgs[1][1]) and imgs[0][2:] == imgs[1][2:]
    small = api.Context(0); small.set_option(api.OPT_BVH_BUILDER, 1); small.upload_scene(load_scene("cbspheres_lambertian_96x72"))   # 14 primitives: host path
    small.set_frame(96, 72)
    assert np.isfinite(small.render(p)[0]).all()


def _soup_scene(base, tri_pos, spheres=None):
    """walls of `base` + a triangle soup (flat normals) [+ spheres]"""
    tri_pos = np.asarray(tri_pos, np.float64).reshape(-1, 9)
    n = np.cross(tri_pos[:, 3:6] - tri_pos[:, 0:3], tri_pos[:, 6:9] - tri_pos[:, 0:3])
    ln = np.linalg.norm(n, axis=1, keepdims=True); ln[ln == 0] = 1.0
    nrm = np.tile(n / ln, 3)
    nb, ns = base.tri_pos.shape[0], 0 if spheres is None else len(spheres)
    nt = nb + len(tri_pos)
    kind = np.concatenate([np.zeros(nt, np.uint8), np.ones(ns, np.uint8)])
    index = np.concatenate([np.arange(nt, dtype=np.uint32), np.arange(ns, dtype=np.uint32)])
    return sc.SceneData(kind, index, np.concatenate([base.tri_pos, tri_pos]), np.concatenate([base.tri_nrm, nrm]),
                        np.concatenate([base.tri_mat, np.zeros(len(tri_pos), np.int32)]),
                        np.zeros((0, 4)) if spheres is None else np.asarray(spheres, np.float64), np.zeros(ns, np.int32), base.mats, base.lights, base.camera)


@pytest.mark.parametrize("case", ["identical", "coplanar_grid", "clustered_with_spheres", "just_above_threshold", "nested_deep"])
def test_device_bvh_build_degenerate_inputs(api, case):
    """Inputs that push the builder's corner cases — every centroid equal (index halving at every level of the reference
    tree, Morton keys distinguished by the primitive id alone), thousands of coplanar triangles, two tight clusters far
    apart plus spheres, 64 primitives — give the host's ranks and the host-built tree's hits, bit for bit."""
    base = load_scene("cbspheres_lambertian_96x72")
    rng = np.random.default_rng(7)
    spheres = None
    if case == "identical":
        tri = np.tile(np.array([[-0.2, 0.3, 0.0, 0.2, 0.3, 0.0, 0.0, 0.6, 0.1]]), (5000, 1))
    elif case == "coplanar_grid":
        g = np.stack(np.meshgrid(np.arange(70), np.arange(70)), -1).reshape(-1, 2) / 70.0 - 0.5
        a = np.concatenate([g, np.full((len(g), 1), 0.0)], 1)[:, [0, 2, 1]] + [0, 0.4, 0]
        tri = np.concatenate([a, a + [1 / 70.0, 0, 0], a + [0, 0, 1 / 70.0]], 1)
    elif case == "clustered_with_spheres":
        c = np.concatenate([rng.normal([-0.6, 0.3, -0.5], 1e-3, (3000, 3)), rng.normal([0.6, 1.2, 0.5], 1e-3, (3000, 3))])
        tri = np.concatenate([c, c + rng.normal(0, 2e-3, c.shape), c + rng.normal(0, 2e-3, c.shape)], 1)
        spheres = np.concatenate([rng.uniform([-0.8, 0.1, -0.8], [0.8, 1.3, 0.8], (40, 3)), rng.uniform(0.01, 0.05, (40, 1))], 1)
    elif case == "nested_deep":
        # clusters nested toward one corner of the scene box, then 6000 identical triangles in the innermost Morton cell: an LBVH of
        # ~30 + 13 levels, deeper than the 40-entry traversal stack => the builder must hand the scene to the host path
        corner = np.array([-0.999, 0.001, -0.999])
        c = np.concatenate([corner + 1.9 * 2.0 ** -k * np.ones((1, 3)) for k in range(1, 31)] + [np.tile(corner + 1e-9, (6000, 1))])
        tri = np.concatenate([c, c + [1e-4, 0, 0], c + [0, 1e-4, 0]], 1)
    else:
        c = rng.uniform([-0.8, 0.1, -0.8], [0.8, 1.3, 0.8], (64 - base.tri_pos.shape[0], 3))
        tri = np.concatenate([c, c + rng.normal(0, 0.05, c.shape), c + rng.normal(0, 0.05, c.shape)], 1)
    s = _soup_scene(base, tri, spheres)
    host = api.bvh_build_host(s, 0)
    dev = api.Context(0); dev.set_option(api.OPT_BVH_BUILDER, 1); dev.upload_scene(s)
    ref = api.Context(0); ref.set_option(api.OPT_BVH_BUILDER, 2); ref.upload_scene(s)
    info = dev.bvh_info()
    assert np.array_equal(info["dfs_rank"], host["dfs_rank"])
    assert (info["n_nodes"], info["n_leaves"], info["depth"]) == (host["n_nodes"], host["n_leaves"], host["depth"])
    ro, rd = dev.generate_rays(rng.uniform(0, 1, (20000, 2)))
    for mode in (sc.MODE_LITERAL, sc.MODE_FLAT):
        dev.set_blackhole((0, 1, 0), 0.1, 0.1, mode); ref.set_blackhole((0, 1, 0), 0.1, 0.1, mode)
        a, _ = dev.trace_rays(ro, rd); b, _ = ref.trace_rays(ro, rd)
        for f in ("status", "prim", "chord", "steps"):
            assert np.array_equal(a[f], b[f]), (case, f)
        assert same(a["t"], b["t"]) and same(a["hit_p"], b["hit_p"]) and same(a["n"], b["n"])
        aa, _ = dev.trace_rays(ro, rd, any_hit=True); bb, _ = ref.trace_rays(ro, rd, any_hit=True)
        assert np.array_equal(aa["status"], bb["status"])
    assert case == "nested_deep" or (a["prim"] >= base.tri_pos.shape[0]).sum() > 50          # the soup is actually hit


def test_c5_full_size_properties(api):
    """BASELINE config 5 at its full frame and triangle count (3840x2160, 1,002,540 triangles, hole on; 1 spp, depth 2 instead of
    4096 spp): the device-built and the host-built BVH give the same image bit for bit, the reference's depth-first order is
    reproduced on the device for all million primitives, the render is deterministic and tile shards add up."""
    synth = importlib.import_module("relativistic-ray-tracer_b200.synth")
    s = synth.c5_scene(load_scene("cbspheres_lambertian_480x360"))
    assert s.n_prims == 1002540
    p = sc.make_params(ns_aa=1, max_ray_depth=2, ns_area_light=1, seed=4)
    cd = api.Context(0); cd.upload_scene(s); cd.set_frame(3840, 2160)            # default: device build at this size
    hd, nd, sd = cd.render(p)
    assert np.isfinite(hd).all() and (hd >= 0).all() and sd["paths"] == 3840 * 2160
    assert (hd.reshape(-1, 3).max(1) > 0).mean() > 0.01
    h2, _, s2 = cd.render(p)
    assert np.array_equal(hd, h2) and s2["rays"] == sd["rays"]
    host = api.bvh_build_host(s, 0)
    info = cd.bvh_info()
    assert np.array_equal(info["dfs_rank"], host["dfs_rank"]) and (info["n_nodes"], info["depth"]) == (host["n_nodes"], host["depth"])
    acc = np.zeros_like(hd)
    for r in range(2):
        cd.set_tile_shard(16, r, 2)
        acc += cd.render(p)[0]
    cd.set_tile_shard(16, 0, 1)
    assert np.array_equal(acc, hd)
    cd.close()
    ch = api.Context(0); ch.set_option(api.OPT_BVH_BUILDER, 2); ch.upload_scene(s); ch.set_frame(3840, 2160)
    hh, nh, sh = ch.render(p)
    assert np.array_equal(hd, hh) and np.array_equal(nd, nh) and sh["rays"] == sd["rays"] and sh["geodesic_steps"] == sd["geodesic_steps"]
