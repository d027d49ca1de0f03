"""Pins the CPU oracle (oracle/rrt_oracle.c) against the golden vectors produced by the reference's own objects
(tests/golden/make_golden.py -> oracle/_ref/ref_probe).  CPU only.  Integer/index results must be identical;
double results are required to be BIT-identical (same IEEE operations in the same order), floats likewise."""
import numpy as np
import pytest

from util import RAND_MAX, golden, hits_from_probe, load_scene, sc
from oracle import orc


def same(a, b):
    """bitwise equality treating NaN == NaN"""
    a = np.asarray(a); b = np.asarray(b)
    return bool(np.all((a == b) | (np.isnan(a) & np.isnan(b))))


def test_philox_known_answers():
    # Random123 kat_vectors, philox4x32 10 rounds
    assert [hex(x) for x in orc.philox(0, [[0, 0, 0, 0]])[0]] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    assert [hex(x) for x in orc.philox(0xFFFFFFFFFFFFFFFF, [[0xFFFFFFFF] * 4])[0]] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    assert [hex(x) for x in orc.philox(0x299F31D0A4093822, [[0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344]])[0]] == \
        ["0xd16cfe09", "0x94fdcceb", "0x5001e420", "0x24126ea1"]


@pytest.mark.parametrize("tag", ["default", "alt", "r0"])
def test_geodesic_chords_bit_exact(tag):
    g = golden("kat_geodesic")
    o = orc.Oracle(load_scene("cbempty_480x360"))
    bh = g["bh_" + tag]
    o.set_blackhole(bh[:3], bh[3], bh[4], sc.MODE_LITERAL)
    ref = g["seg_" + tag]
    got = o.geodesic(g["o"], g["d"], ref.shape[1])
    assert same(got, ref)
    if tag == "default":
        assert np.isnan(ref).any() and (ref[..., 7] > 0).any()  # the set does contain radial (NaN) rays and captures
    assert o.chords_per_ray() == (63 if bh[4] == 0.1 else 32)


def test_geodesic_survey_appendix_e():
    o = orc.Oracle(load_scene("cbempty_480x360"))
    seg = o.geodesic([[0.5, 0, 0.5]], [[-0.25097738160425154, 0.93489074647583692, -0.25097738160425154]], 32)[0]
    assert seg[0, 3] == -0.2513010566264775 and seg[0, 6] == 0.35449775213862111
    assert seg[31, 0] == -0.50480278519061161 and seg[31, 6] == 0.99706616486007338


@pytest.mark.parametrize("scene", ["cbspheres_lambertian_480x360", "cbgems_480x360", "cbbunny_microfacet_480x360"])
@pytest.mark.parametrize("mode", ["flat", "bh", "alt"])
def test_bvh_intersect_matches_reference(scene, mode):
    g = golden("kat_trace_" + scene)
    if "hit_" + mode not in g.files:
        pytest.skip("no alt-hole vectors for this scene")
    s = load_scene(scene)
    o = orc.Oracle(s)
    if mode == "flat":
        o.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_FLAT)
    elif mode == "alt":
        bh = g["bh_alt"]; o.set_blackhole(bh[:3], bh[3], bh[4], sc.MODE_LITERAL)
    rays = g["rays_" + mode]; ref = hits_from_probe(g["hit_" + mode])
    got, _ = o.trace_rays(rays[:, 0:3], rays[:, 3:6], rays[:, 6], rays[:, 7])
    assert np.array_equal(got["status"] == 1, ref["hit"])
    assert np.array_equal(got["prim"][ref["hit"]], ref["prim"][ref["hit"]])
    h = ref["hit"]
    assert same(got["hit_p"][h], ref["hit_p"][h]) and same(got["w_out"][h], ref["w_out"][h]) and same(got["n"][h], ref["n"][h])
    anyh, _ = o.trace_rays(rays[:, 0:3], rays[:, 3:6], rays[:, 6], rays[:, 7], any_hit=True)
    assert np.array_equal(anyh["status"] == 1, ref["any"])
    assert ref["hit"].sum() > 100 and (~ref["hit"]).sum() > 10


def test_bvh_shape_matches_survey():
    assert {k: v for k, v in orc.Oracle(load_scene("cbspheres_lambertian_480x360")).bvh_info().items() if k != "dfs_rank"} == \
        dict(n_nodes=7, n_leaves=4, depth=2)
    assert {k: v for k, v in orc.Oracle(load_scene("cbbunny_microfacet_480x360")).bvh_info().items() if k != "dfs_rank"} == \
        dict(n_nodes=19103, n_leaves=9552, depth=19)


def test_camera_rays_bit_exact():
    g = golden("kat_trace_cbspheres_lambertian_480x360")
    o, d = orc.Oracle(load_scene("cbspheres_lambertian_480x360")).generate_rays(g["xy"])
    assert same(o, g["cam"][:, 0:3]) and same(d, g["cam"][:, 3:6])


@pytest.mark.parametrize("scene", ["cbspheres_lambertian_480x360", "cbspheres_480x360", "cbspheres_microfacet_480x360"])
def test_bsdf_f_and_sample_f(scene):
    g = golden("kat_bsdf_" + scene)
    s = load_scene(scene)
    o = orc.Oracle(s)
    kinds = set()
    for mi in range(len(s.mats)):
        inp, ref = g["in_%d" % mi], g["out_%d" % mi]
        kinds.add(int(s.mats[mi]["type"]))
        u = inp[:, 6:10] / RAND_MAX
        f, wi, pdf = o.bsdf_sample(mi, inp[:, 0:3], u)
        fe = o.bsdf_eval(mi, inp[:, 0:3], inp[:, 3:6])
        assert same(fe.astype(np.float64), ref[:, 0:3]), (scene, mi, "f")
        assert same(f.astype(np.float64), ref[:, 3:6]), (scene, mi, "sample_f value")
        assert same(pdf.astype(np.float64), ref[:, 9]), (scene, mi, "pdf")
        live = ref[:, 9] != 0
        assert same(wi[live], ref[live, 6:9]), (scene, mi, "wi")
    assert kinds  # materials covered across the three scenes: diffuse, emission, mirror, glass, microfacet


@pytest.mark.parametrize("scene", ["cbspheres_lambertian_480x360", "cbempty_480x360"])
def test_light_sample_L(scene):
    g = golden("kat_light_" + scene)
    s = load_scene(scene)
    inp, ref = g["inp"], g["out"]
    rad, wi, dist, pdf = orc.Oracle(s).light_sample(0, inp[:, 0:3], inp[:, 3:5] / RAND_MAX)
    assert same(rad.astype(np.float64), ref[:, 0:3]) and same(wi, ref[:, 3:6])
    assert same(dist.astype(np.float64), ref[:, 6]) and same(pdf.astype(np.float64), ref[:, 7])


def test_hemisphere_and_directional_light_sample_L():
    """InfiniteHemisphereLight and DirectionalLight (reference src/static_scene/light.cpp:34-42, 17-23): the two on-path light classes
    no Cornell box has; vectors from the reference's own objects on dae/keenan/banana.dae (<ambient> and <directional>)"""
    g = golden("kat_light_banana_480x360")
    s = load_scene("banana_480x360")
    assert [int(l["type"]) for l in s.lights] == [1, 0] and [int(l["is_delta"]) for l in s.lights] == [0, 1]   # RRT_LIGHT_HEMISPHERE, RRT_LIGHT_DIRECTIONAL
    o = orc.Oracle(s)
    inp = g["inp"]
    for k, key in ((0, "out"), (1, "out_1")):
        ref = g[key]
        rad, wi, dist, pdf = o.light_sample(k, inp[:, 0:3], inp[:, 3:5] / RAND_MAX)
        assert same(rad.astype(np.float64), ref[:, 0:3]) and same(wi, ref[:, 3:6]), k
        assert same(dist.astype(np.float64), ref[:, 6]) and same(pdf.astype(np.float64), ref[:, 7]), k
    hemi, dirl = g["out"], g["out_1"]
    assert np.isinf(hemi[:, 6]).all() and np.allclose(hemi[:, 7], 1 / (2 * np.pi)) and (hemi[:, 4] >= -1e-7).all()   # uniform over the +y hemisphere (the reference's sampleToWorld has float-rounded zeros)
    assert np.isinf(dirl[:, 6]).all() and (dirl[:, 7] == 1).all() and (dirl[:, 3:6] == dirl[0, 3:6]).all()        # one fixed direction, pdf 1


def test_environment_light_sample_L_and_sample_dir():
    """EnvironmentLight (reference src/static_scene/environment_light.cpp) as the last light of the scene: importance
    sampling through the marginal / conditional tables, and the direction lookup used for camera rays that miss"""
    g = golden("kat_envmap")
    o = orc.Oracle(load_scene(str(g["scene"])))
    o.set_envmap(g["env"])
    inp, ref = g["inp"], g["out"]
    rad, wi, dist, pdf = o.light_sample(int(g["light_index"]), inp[:, 0:3], inp[:, 3:5] / RAND_MAX)
    assert same(rad.astype(np.float64), ref[:, 0:3]) and same(wi, ref[:, 3:6])
    assert np.isinf(dist).all() and np.isinf(ref[:, 6]).all() and same(pdf.astype(np.float64), ref[:, 7])
    assert (ref[:, 8] == 2).all() and (ref[:, 9] == 0).all()      # two draws per sample, not a delta light
    assert same(o.env_lookup(g["dirs"]).astype(np.float64), g["dir_out"])


def test_thin_lens_rays_match_reference():
    """Camera::generate_ray_for_thin_lens (-b / -d; the reference compiles its call out with THIN_LENS 0 but ships the function)"""
    g = golden("kat_thin_lens")
    o, d = orc.Oracle(load_scene(str(g["scene"]))).generate_rays_thin_lens(g["inp"][:, :2], g["inp"][:, 2:])
    assert np.array_equal(o, g["out"][:, :3]) and np.array_equal(d, g["out"][:, 3:6])
