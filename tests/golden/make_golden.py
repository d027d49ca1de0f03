#!/usr/bin/env python
"""Regenerates tests/golden/* from the reference's own code (oracle/_ref/ref_probe[_flat], built by
oracle/build_ref.sh from /root/reference).  Runs only in the build container; the fixtures it writes are
committed so that the GPU box (no /root/reference) can check against them.

    python tests/golden/make_golden.py [kat] [render] [scenes] [fullsize] [bvhorder] [rng] [piz]

Inputs are seeded numpy draws; outputs are what the reference's BVHAccel::intersect, BlackHole::next_micro_ray,
Camera::generate_ray, BSDF::f/sample_f, SceneLight::sample_L and PathTracer::raytrace_pixel return for them.
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.path.join(ROOT, "oracle", "_ref")
DAE = os.path.join(REF, "dae")
PROBE = {"bh": os.path.join(REF, "ref_probe"), "flat": os.path.join(REF, "ref_probe_flat")}
RAND_MAX = 2147483647
TMP = tempfile.mkdtemp(prefix="rrt_golden_")


def run(args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    subprocess.check_call([str(a) for a in args], env=e, stdout=subprocess.DEVNULL, cwd=TMP)   # cwd: EnvironmentLight drops a debug png there


def probe(mode, cmd, *args, env=None):
    run([PROBE[mode], cmd] + list(args), env)


def tmpf(name):
    return os.path.join(TMP, name)


def wr(name, a):
    p = tmpf(name)
    np.ascontiguousarray(a, np.float64).tofile(p)
    return p


def rd(name, cols):
    return np.fromfile(tmpf(name), np.float64).reshape(-1, cols)


# ---------------------------------------------------------------- scenes
SCENES = {  # name -> (dae, W, H, screenDist scale)
    "cbspheres_lambertian_480x360": ("CBspheres_lambertian.dae", 480, 360, None),
    "cbspheres_480x360": ("CBspheres.dae", 480, 360, None),
    "cbspheres_microfacet_480x360": ("CBspheres_microfacet_al_ag.dae", 480, 360, None),
    "cbgems_480x360": ("CBgems.dae", 480, 360, None),
    "cbempty_480x360": ("CBempty.dae", 480, 360, None),
    "cbbunny_microfacet_480x360": ("CBbunny_microfacet_cu.dae", 480, 360, None),
    # small frames that still see the whole box: screenDist scaled by W/480
    "cbspheres_lambertian_96x72": ("CBspheres_lambertian.dae", 96, 72, 0.2),
    "cbspheres_96x72": ("CBspheres.dae", 96, 72, 0.2),
    "cbspheres_microfacet_96x72": ("CBspheres_microfacet_al_ag.dae", 96, 72, 0.2),
    "cbgems_96x72": ("CBgems.dae", 96, 72, 0.2),
    "cbempty_96x72": ("CBempty.dae", 96, 72, 0.2),
    "cbbunny_microfacet_96x72": ("CBbunny_microfacet_cu.dae", 96, 72, 0.2),
    # C5's generator at 4,608 triangles, exported as COLLADA (synth.write_c5_dae) and loaded by the reference's own loader
    "cbsynth_480x360": ("_generated_CBsynth_mesh.dae", 480, 360, None),
    "cbsynth_96x72": ("_generated_CBsynth_mesh.dae", 96, 72, 0.2),
    # dae/keenan/banana.dae: the bundled scene with the two light classes no Cornell box has — <ambient> = InfiniteHemisphereLight,
    # <directional> = DirectionalLight (reference src/static_scene/light.cpp:17-42)
    "banana_480x360": ("banana.dae", 480, 360, None),
    "banana_96x72": ("banana.dae", 96, 72, 0.2),
    # the frames of BASELINE.json configs 3 and 4 (camera blocks only: -r crops / extends the view, the geometry is the 480x360 dump)
    "cbbunny_microfacet_1280x720": ("CBbunny_microfacet_cu.dae", 1280, 720, None),
    "cbbunny_microfacet_1920x1080": ("CBbunny_microfacet_cu.dae", 1920, 1080, None),
}
SYNTH_MESH = (48, 49)   # n_u, n_v of the exported mesh


def ensure_generated_dae():
    sys.path.insert(0, ROOT)
    import importlib
    synth = importlib.import_module("relativistic-ray-tracer_b200.synth")
    synth.write_c5_dae(os.path.join(DAE, "_generated_CBsynth_mesh.dae"), open(os.path.join(DAE, "CBspheres_lambertian.dae")).read(), *SYNTH_MESH)


def scene_env(name):
    k = SCENES[name][3]
    return {"RRT_PROBE_SCREENDIST_SCALE": str(k)} if k else {}


def make_scenes(only=None):
    sys.path.insert(0, ROOT)
    import importlib
    sc = importlib.import_module("relativistic-ray-tracer_b200.scene")
    ensure_generated_dae()
    for name, (dae, w, h, _) in SCENES.items():
        if only and name not in only:
            continue
        out = tmpf(name + ".rrtscene")
        probe("bh", "scene", os.path.join(DAE, dae), w, h, out, env=scene_env(name))
        s = sc.SceneData.load(out)
        raw = open(out, "rb").read()
        dst = os.path.join(HERE, name + ".rrtscene")
        if not name.endswith("_480x360"):  # same geometry as the 480x360 dump: keep only the camera block (24 doubles)
            np.save(os.path.join(HERE, name + ".cam.npy"), np.frombuffer(raw, np.float64, 24, 40))
            print("scene", name, "-> camera block only")
            continue
        if len(raw) > 256 * 1024:  # big meshes: compressed container of the same bytes
            np.savez_compressed(dst + ".npz", raw=np.frombuffer(raw, np.uint8))
            print("scene", name, s.n_prims, "prims ->", os.path.getsize(dst + ".npz"), "bytes (npz)")
        else:
            open(dst, "wb").write(raw)
            print("scene", name, s.n_prims, "prims ->", len(raw), "bytes")


# ---------------------------------------------------------------- KATs
def unit(v):
    return v / np.linalg.norm(v, axis=-1, keepdims=True)


def kat_geodesic():
    rng = np.random.default_rng(184)
    n = 192
    o = np.stack([rng.uniform(-1, 1, n), rng.uniform(0, 1.5, n), rng.uniform(-1, 5, n)], 1)
    d = unit(rng.normal(size=(n, 3)))
    c = np.array([0.0, 1.0, 0.0])
    # rays aimed close to the hole (strong bending / capture), exactly radial rays (dy = 0 -> NaN chords), App. E ray
    aim = c + rng.normal(scale=0.15, size=(48, 3))
    o2 = np.stack([rng.uniform(-1, 1, 48), rng.uniform(0, 1.5, 48), rng.uniform(1, 5, 48)], 1)
    d2 = unit(aim - o2)
    o3 = np.array([[0.0, 1.0, 3.0], [0.0, 0.0, 0.0], [1.0, 1.0, 0.0]])
    d3 = unit(c - o3)
    o4 = np.array([[0.5, 0.0, 0.5]])
    d4 = unit(np.array([[0.1, 1.49, 0.1]]) - o4)
    O = np.concatenate([o, o2, o3, o4]); D = np.concatenate([d, d2, d3, d4])
    rays = wr("geo_in.bin", np.concatenate([O, D], 1))
    out = {}
    for tag, bh, nsteps in (("default", None, 63), ("alt", (0.2, 0.8, -0.1, 0.05, 0.2), 32), ("r0", (0, 1, 0, 0, 0.1), 63)):
        args = ["geodesic", rays, nsteps, tmpf("geo_out.bin")]
        if bh:
            args += ["-B"] + list(bh)
        probe("bh", *args)
        out["seg_" + tag] = rd("geo_out.bin", 8).reshape(len(O), nsteps, 8)
        out["bh_" + tag] = np.array(bh if bh else (0, 1, 0, 0.1, 0.1), np.float64)
    np.savez_compressed(os.path.join(HERE, "kat_geodesic.npz"), o=O, d=D, **out)
    print("kat_geodesic", len(O), "rays")


def kat_trace(scene, gw, gh, n_extra, seed, bh_alt=None):
    dae, w, h, _ = SCENES[scene]
    dae = os.path.join(DAE, dae)
    env = scene_env(scene)
    rng = np.random.default_rng(seed)
    # camera rays through the reference's own generate_ray (pixel centres of a gw x gh grid + random jitter set)
    gx, gy = np.meshgrid((np.arange(gw) + 0.5) / gw, (np.arange(gh) + 0.5) / gh)
    xy = np.concatenate([np.stack([gx.ravel(), gy.ravel()], 1), rng.uniform(0, 1, (n_extra, 2))])
    probe("bh", "camrays", dae, w, h, wr("xy.bin", xy), tmpf("cam.bin"), env=env)
    cam = rd("cam.bin", 8)
    res = {"xy": xy, "cam": cam}
    for mode in ("flat", "bh"):
        rays = cam.copy()
        probe(mode, "trace", dae, w, h, wr("rays.bin", rays), tmpf("hit.bin"), env=env)
        hit = rd("hit.bin", 16)
        # secondary rays from the hit points: shadow-style (toward the light quad, max_t = float dist) and random bounces
        ok = hit[:, 0] > 0
        p = hit[ok, 2:5]
        m = min(len(p), 3000)
        sel = rng.choice(len(p), m, replace=False)
        p = p[sel]
        lp = np.stack([rng.uniform(-0.3, 0.3, m), np.full(m, 1.49), rng.uniform(-0.4, 0.4, m)], 1)
        dl = lp - p
        dist = np.sqrt(np.float32((dl * dl).sum(1))).astype(np.float32)
        wi = dl / dist[:, None].astype(np.float64)
        sh = np.concatenate([p + 1e-11 * wi, wi, np.zeros((m, 1)), dist[:, None].astype(np.float64)], 1)
        wb = unit(rng.normal(size=(m, 3)))
        bo = np.concatenate([p + 1e-11 * wb, wb, np.zeros((m, 1)), np.full((m, 1), np.inf)], 1)
        sec = np.concatenate([sh, bo])
        probe(mode, "trace", dae, w, h, wr("rays2.bin", sec), tmpf("hit2.bin"), env=env)
        res["rays_" + mode] = np.concatenate([rays, sec])
        res["hit_" + mode] = np.concatenate([hit, rd("hit2.bin", 16)])
    if bh_alt:
        probe("bh", "trace", dae, w, h, wr("rays.bin", cam), tmpf("hit.bin"), "-B", *bh_alt, env=env)
        res["rays_alt"] = cam
        res["hit_alt"] = rd("hit.bin", 16)
        res["bh_alt"] = np.array(bh_alt, np.float64)
    np.savez_compressed(os.path.join(HERE, "kat_trace_%s.npz" % scene), **res)
    print("kat_trace", scene, {k: v.shape for k, v in res.items() if k.startswith("hit")})


def kat_bsdf(scene, seed):
    dae, w, h, _ = SCENES[scene]
    dae = os.path.join(DAE, dae)
    rng = np.random.default_rng(seed)
    sys.path.insert(0, ROOT)
    import importlib
    sc = importlib.import_module("relativistic-ray-tracer_b200.scene")
    s = sc.SceneData.load(os.path.join(HERE, scene + ".rrtscene"))
    n = 512
    out = {}
    for mi in range(len(s.mats)):
        wo = unit(rng.normal(size=(n, 3)))
        wo[: n // 4, 2] = np.abs(wo[: n // 4, 2])  # more upper-hemisphere
        wi = unit(rng.normal(size=(n, 3)))
        wi[: n // 2, 2] = np.abs(wi[: n // 2, 2])
        wo[: n // 2, 2] = np.abs(wo[: n // 2, 2])
        r = rng.integers(0, RAND_MAX, (n, 4), endpoint=True).astype(np.float64)
        r[0] = [0, 0, 0, 0]; r[1] = [RAND_MAX] * 4; r[2] = [0, RAND_MAX, 0, RAND_MAX]; r[3] = [RAND_MAX, 0, RAND_MAX, 0]
        inp = np.concatenate([wo, wi, r], 1)
        probe("bh", "bsdf", dae, w, h, mi, wr("bsdf_in.bin", inp), tmpf("bsdf_out.bin"))
        out["in_%d" % mi] = inp
        out["out_%d" % mi] = rd("bsdf_out.bin", 12)
    np.savez_compressed(os.path.join(HERE, "kat_bsdf_%s.npz" % scene), **out)
    print("kat_bsdf", scene, len(s.mats), "materials")


def kat_light(scene, seed, lights=(0,)):
    """SceneLight::sample_L of the scene's lights `lights` for seeded points and uniforms: `out` = light lights[0], `out_<k>` = light k"""
    dae, w, h, _ = SCENES[scene]
    dae = os.path.join(DAE, dae)
    rng = np.random.default_rng(seed)
    n = 512
    p = np.stack([rng.uniform(-1, 1, n), rng.uniform(0, 1.6, n), rng.uniform(-1, 1, n)], 1)
    r = rng.integers(0, RAND_MAX, (n, 2), endpoint=True).astype(np.float64)
    r[0] = [0, 0]; r[1] = [RAND_MAX, RAND_MAX]
    inp = np.concatenate([p, r], 1)
    outs = {}
    for k in lights:
        probe("bh", "light", dae, w, h, k, wr("light_in.bin", inp), tmpf("light_out.bin"))
        outs["out" if k == lights[0] else "out_%d" % k] = rd("light_out.bin", 10)
    np.savez_compressed(os.path.join(HERE, "kat_light_%s.npz" % scene), inp=inp, lights=np.array(lights), **outs)
    print("kat_light", scene, lights)


# ---------------------------------------------------------------- environment map (the reference ships none)
def synthetic_envmap(w=64, h=32, seed=30):
    """smooth sky gradient + a bright sun + seeded texture, strictly positive (a black row would put 0/0 into the
    reference's conditional table); [h, w, 3] float32, row 0 = top"""
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    sky = np.stack([0.25 + 0.5 * y / h, 0.35 + 0.3 * (1 - y / h), 0.6 + 0.3 * np.cos(2 * np.pi * x / w)], -1)
    sun = 40.0 * np.exp(-(((x - 0.7 * w) / 2.5) ** 2 + ((y - 0.3 * h) / 2.0) ** 2))
    tex = rng.uniform(0.0, 0.15, (h, w, 3))
    return (sky + sun[..., None] * np.array([1.0, 0.9, 0.7]) + tex).astype(np.float32)


def envmap_file(env):
    p = tmpf("env.bin")
    with open(p, "wb") as f:
        f.write(np.array([env.shape[1], env.shape[0]], np.int32).tobytes()); f.write(np.ascontiguousarray(env, np.float32).tobytes())
    return p


def kat_envmap(scene, seed):
    """EnvironmentLight::sample_L (as the last light of the scene) and ::sample_dir for seeded inputs"""
    sys.path.insert(0, ROOT)
    import importlib
    sc = importlib.import_module("relativistic-ray-tracer_b200.scene")
    dae, w, h, _ = SCENES[scene]
    n_lights = len(sc.SceneData.load(os.path.join(HERE, scene + ".rrtscene")).lights)
    env = synthetic_envmap()
    e = {"RRT_PROBE_ENVMAP": envmap_file(env)}
    rng = np.random.default_rng(seed)
    n = 2048
    p = np.stack([rng.uniform(-1, 1, n), rng.uniform(0, 1.6, n), rng.uniform(-1, 1, n)], 1)
    r = rng.integers(0, RAND_MAX, (n, 2), endpoint=True).astype(np.float64)
    r[0] = [0, 0]; r[1] = [RAND_MAX - 1, RAND_MAX - 1]; r[2] = [0, RAND_MAX - 1]; r[3] = [RAND_MAX - 1, 0]
    inp = np.concatenate([p, r], 1)
    probe("bh", "light", os.path.join(DAE, dae), w, h, n_lights, wr("env_in.bin", inp), tmpf("env_out.bin"), env=e)
    d = unit(rng.normal(size=(n, 3)))
    d[:6] = [[0, 1, 0], [0, -1, 0], [1, 0, 0], [-1, 0, 0], [0, 0, 1], [0, 0, -1]]
    probe("bh", "envdir", os.path.join(DAE, dae), w, h, wr("envdir_in.bin", d), tmpf("envdir_out.bin"), env=e)
    np.savez_compressed(os.path.join(HERE, "kat_envmap.npz"), env=env, scene=scene, light_index=n_lights, inp=inp, out=rd("env_out.bin", 10),
                        dirs=d, dir_out=rd("envdir_out.bin", 3))
    print("kat_envmap", scene)


def kat_thin_lens(scene, seed):
    """Camera::generate_ray_for_thin_lens for seeded (x, y, rndR, rndTheta); lens parameters are the AppConfig defaults the fixture carries"""
    dae, w, h, _ = SCENES[scene]
    rng = np.random.default_rng(seed)
    n = 2048
    inp = np.stack([rng.uniform(0, 1, n), rng.uniform(0, 1, n), rng.uniform(0, 1, n), rng.uniform(0, 2 * np.pi, n)], 1)
    inp[0] = [0.5, 0.5, 0.0, 0.0]; inp[1] = [0.0, 1.0, 1.0, 2 * np.pi]; inp[2] = [1.0, 0.0, 1.0, np.pi]
    probe("bh", "lensrays", os.path.join(DAE, dae), w, h, wr("lens_in.bin", inp), tmpf("lens_out.bin"), env=scene_env(scene))
    np.savez_compressed(os.path.join(HERE, "kat_thin_lens.npz"), scene=scene, inp=inp, out=rd("lens_out.bin", 8))
    print("kat_thin_lens", scene)


def make_kat(only=None):
    if only and "thin_lens" in only:
        kat_thin_lens("cbspheres_lambertian_480x360", 23)
        return
    if only and "keenan" in only:
        kat_light("banana_480x360", 25, lights=(0, 1))
        return
    kat_thin_lens("cbspheres_lambertian_480x360", 23)
    kat_geodesic()
    kat_trace("cbspheres_lambertian_480x360", 96, 72, 2000, 1, bh_alt=(0.2, 0.8, -0.1, 0.05, 0.2))
    kat_trace("cbgems_480x360", 64, 48, 1000, 2)
    kat_trace("cbbunny_microfacet_480x360", 96, 72, 2000, 3)
    for i, sname in enumerate(("cbspheres_lambertian_480x360", "cbspheres_480x360", "cbspheres_microfacet_480x360")):
        kat_bsdf(sname, 10 + i)
    kat_light("cbspheres_lambertian_480x360", 20)
    kat_light("cbempty_480x360", 21)
    kat_light("banana_480x360", 25, lights=(0, 1))
    kat_envmap("cbspheres_lambertian_480x360", 22)


# ---------------------------------------------------------------- renders (reference raytrace_pixel, Philox-fed rand)
RENDERS = {  # name -> (scene, mode, spp, l, m, batch, tol, seed, extra args)
    "c1_flat_lambertian": ("cbspheres_lambertian_96x72", "flat", 16, 4, 5, 32, 0.05, 1, []),
    "c2_bh_lambertian": ("cbspheres_lambertian_96x72", "bh", 8, 4, 5, 32, 0.05, 2, []),
    "bh_glass_mirror": ("cbspheres_96x72", "bh", 8, 2, 6, 32, 0.05, 3, []),
    "flat_glass_mirror": ("cbspheres_96x72", "flat", 16, 2, 8, 32, 0.05, 4, []),
    "flat_microfacet": ("cbspheres_microfacet_96x72", "flat", 16, 2, 5, 32, 0.05, 5, []),
    "bh_microfacet": ("cbspheres_microfacet_96x72", "bh", 4, 1, 4, 32, 0.05, 6, []),
    "flat_gems": ("cbgems_96x72", "flat", 8, 2, 8, 32, 0.05, 7, []),
    "flat_bunny_microfacet": ("cbbunny_microfacet_96x72", "flat", 8, 2, 5, 32, 0.05, 8, []),
    "bh_bunny_microfacet": ("cbbunny_microfacet_96x72", "bh", 2, 1, 3, 32, 0.05, 9, []),
    "flat_adaptive": ("cbspheres_lambertian_96x72", "flat", 64, 2, 4, 8, 0.2, 10, []),
    "bh_adaptive": ("cbspheres_lambertian_96x72", "bh", 32, 1, 3, 8, 0.3, 11, []),
    "flat_depth0": ("cbspheres_lambertian_96x72", "flat", 4, 1, 0, 32, 0.05, 12, []),
    "flat_depth1": ("cbspheres_lambertian_96x72", "flat", 4, 4, 1, 32, 0.05, 13, []),
    "bh_depth1_spp1": ("cbspheres_lambertian_96x72", "bh", 1, 4, 1, 32, 0.05, 14, []),
    "flat_pointlight": ("cbempty_96x72", "flat", 8, 1, 4, 32, 0.05, 15, []),
    "flat_hemisphere": ("cbspheres_lambertian_96x72", "flat", 8, 4, 3, 32, 0.05, 16, ["hemi"]),
    "bh_alt_hole": ("cbspheres_lambertian_96x72", "bh", 4, 2, 4, 32, 0.05, 17, ["-B", 0.2, 0.8, -0.1, 0.05, 0.2]),
    # environment light (-e): synthetic_envmap() handed to the reference through RRT_PROBE_ENVMAP
    "flat_envmap": ("cbspheres_lambertian_96x72", "flat", 8, 2, 4, 32, 0.05, 18, ["env"]),
    "bh_envmap": ("cbspheres_lambertian_96x72", "bh", 4, 2, 3, 32, 0.05, 19, ["env"]),
    "bh_envmap_gems": ("cbgems_96x72", "bh", 4, 1, 5, 32, 0.05, 20, ["env"]),
    # thin lens (-b / -d): raytrace_pixel(x, y, useThinLens = true), which the reference's THIN_LENS switch compiles out of its tile loop
    "flat_thin_lens": ("cbspheres_lambertian_96x72", "flat", 8, 2, 4, 32, 0.05, 21, ["lens"]),
    "bh_thin_lens": ("cbspheres_lambertian_96x72", "bh", 4, 1, 3, 32, 0.05, 22, ["lens"]),
    # the procedural mesh (C5's generator), through the reference's loader and raytrace_pixel
    "flat_synth_mesh": ("cbsynth_96x72", "flat", 8, 2, 4, 32, 0.05, 23, []),
    "bh_synth_mesh": ("cbsynth_96x72", "bh", 4, 2, 4, 32, 0.05, 24, []),
    # InfiniteHemisphereLight + DirectionalLight (dae/keenan/banana.dae)
    "flat_keenan_lights": ("banana_96x72", "flat", 16, 2, 4, 32, 0.05, 26, []),
    "bh_keenan_lights": ("banana_96x72", "bh", 8, 2, 3, 32, 0.05, 27, []),
}


def make_render(only=None):
    ensure_generated_dae()
    for name, (scene, mode, spp, l, m, batch, tol, seed, extra) in RENDERS.items():
        if only and name not in only:
            continue
        dae, w, h, _ = SCENES[scene]
        out = tmpf("render.bin")
        env = scene_env(scene)
        if "env" in extra:
            env = dict(env); env["RRT_PROBE_ENVMAP"] = envmap_file(synthetic_envmap())
        probe(mode, "render", os.path.join(DAE, dae), w, h, spp, l, m, batch, tol, seed, out, *[x for x in extra if x != "env"], env=env)
        a = np.fromfile(out, np.float64)
        img = a[: w * h * 4].reshape(h, w, 4)
        tail = a[w * h * 4:]
        np.savez_compressed(os.path.join(HERE, "render_%s.npz" % name), hdr=img[..., :3].astype(np.float32),
                            counts=img[..., 3].astype(np.int32), scene=scene, mode=mode,
                            params=np.array([spp, l, m, batch, tol, seed], np.float64),
                            hemi=np.array("hemi" in extra), envmap=np.array("env" in extra), lens=np.array("lens" in extra), bh=np.array([x for x in extra if not isinstance(x, str)], np.float64),
                            draws=tail[2])
        print("render", name, "mean radiance %.5f  paths %d" % (img[..., :3].mean(), img[..., 3].sum()))


# ---------------------------------------------------------------- the reference's own BVH, walked (tie-break ranks)
def kat_bvh_order():
    out = {}
    for scene in ("cbspheres_lambertian_480x360", "cbgems_480x360", "cbbunny_microfacet_480x360", "banana_480x360", "cbsynth_480x360"):
        dae, w, h, _ = SCENES[scene]
        probe("bh", "bvhorder", os.path.join(DAE, dae), w, h, tmpf("order.bin"))
        a = np.fromfile(tmpf("order.bin"), np.float64)
        out[scene + "_info"] = a[:3].astype(np.int64)          # nodes, leaves, depth of BVHAccel::get_root()'s tree
        out[scene + "_rank"] = a[3:].astype(np.uint32)         # per primitive (reference order): position in the depth-first leaf order
        print("kat_bvh_order", scene, a[:3], len(a) - 3)
    np.savez_compressed(os.path.join(HERE, "kat_bvh_order.npz"), **out)


# ---------------------------------------------------------------- independent-RNG reference image (statistical parity)
def render_independent_rng():
    """The reference's raytrace_pixel driven by glibc's OWN generator (random_r, the algorithm behind rand(); nothing shared with the product's
    Philox streams), adaptive stop off, 96x72, 4096 spp, hole on: what a converged run of the unmodified binary estimates, kept as HDR."""
    scene, spp, l, m = "cbspheres_lambertian_96x72", 4096, 2, 4
    dae, w, h, _ = SCENES[scene]
    env = dict(scene_env(scene)); env["RRT_PROBE_GLIBC_RAND"] = "1"
    out = tmpf("render_rng.bin")
    probe("bh", "render", os.path.join(DAE, dae), w, h, spp, l, m, spp + 1, 0.0, 12345, out, env=env)
    a = np.fromfile(out, np.float64)
    img = a[: w * h * 4].reshape(h, w, 4)
    np.savez_compressed(os.path.join(HERE, "render_independent_rng.npz"), hdr=img[..., :3].astype(np.float32), counts=img[..., 3].astype(np.int32),
                        scene=scene, mode="bh", params=np.array([spp, l, m, spp + 1, 0.0, 0], np.float64))
    print("render_independent_rng mean radiance %.5f  paths %d  %.1f s" % (img[..., :3].mean(), img[..., 3].sum(), a[w * h * 4 + 3]))


# ---------------------------------------------------------------- full-size frames of BASELINE.json's own configs
FULLSIZE = {  # name -> (scene, mode, spp, l, m, batch, tol, seed, cell (x0, y0, w, h) or None = whole frame)
    "c1": ("cbspheres_lambertian_480x360", "flat", 16, 4, 5, 32, 0.05, 1, None),
    "c2": ("cbspheres_lambertian_480x360", "bh", 16, 4, 5, 32, 0.05, 1, None),
    # C3: a 256x144 cell of the 1280x720 frame over the bunny (flat, 256 spp; the adaptive rule acts from 32 spp on)
    "c3_cell": ("cbbunny_microfacet_1280x720", "flat", 256, 4, 5, 32, 0.05, 1, (512, 216, 256, 144)),
    # C4: a 256x144 cell of the 1920x1080 hole frame, left of the hole's image: mesh, walls, lensed region (64 spp, two adaptive batches)
    "c4_cell": ("cbbunny_microfacet_1920x1080", "bh", 64, 4, 5, 32, 0.05, 1, (736, 468, 256, 144)),
}


def make_fullsize(only=None):
    """Rendered by the reference's own raytrace_pixel (Philox-fed rand()) at the configs' real frame sizes; minutes of CPU each."""
    for name, (scene, mode, spp, l, m, batch, tol, seed, cell) in FULLSIZE.items():
        if only and name not in only:
            continue
        dae, w, h, _ = SCENES[scene]
        out = tmpf("render_full.bin")
        env = dict(scene_env(scene))
        x0, y0, cw, ch = cell if cell else (0, 0, w, h)
        env["RRT_PROBE_ROWS"] = "%d:%d" % (y0, y0 + ch); env["RRT_PROBE_COLS"] = "%d:%d" % (x0, x0 + cw)
        probe(mode, "render", os.path.join(DAE, dae), w, h, spp, l, m, batch, tol, seed, out, env=env)
        a = np.fromfile(out, np.float64)
        img = a[: w * h * 4].reshape(h, w, 4)[y0:y0 + ch, x0:x0 + cw]
        np.savez_compressed(os.path.join(HERE, "render_full_%s.npz" % name), hdr=img[..., :3].astype(np.float32), counts=img[..., 3].astype(np.int32),
                            scene=scene, mode=mode, params=np.array([spp, l, m, batch, tol, seed], np.float64), frame=np.array([w, h]),
                            cell=np.array([x0, y0, cw, ch]), seconds=a[w * h * 4 + 3])
        print("render_full", name, "mean radiance %.5f  paths %d  %.1f s" % (img[..., :3].mean(), img[..., 3].sum(), a[w * h * 4 + 3]), flush=True)


PIZ_FILES = {   # name -> (pixel type: 1 half / 2 float, what it exercises)
    "envmap_piz_half_96x70": (1, "sun + sky + noise: three chunks (32, 32, 6 lines), 14-bit wavelet, Huffman codes well past 12 bits, runs"),
    "envmap_piz_float_20x35": (2, "32-bit samples (two 16-bit planes per channel), noise: the chunks come out larger than their data"),
    "envmap_piz_half_wide_200x37": (1, "19 200 distinct values in the first chunk: more than 2^14 numbers, the modulo-2^16 wavelet"),
}


def piz_image(name):
    rng = np.random.default_rng(40)
    if name == "envmap_piz_half_96x70":
        yy, xx = np.mgrid[0:70, 0:96]
        sun = np.exp(-((yy - 12) ** 2 + (xx - 60) ** 2) / 40.0)[..., None] * np.array([900.0, 700.0, 400.0])
        sky = (0.2 + 0.6 * (1 - yy / 70.0))[..., None] * np.array([0.4, 0.6, 1.0])
        return (sun + sky + rng.uniform(0, 0.02, (70, 96, 3))).astype(np.float32)
    if name == "envmap_piz_float_20x35":
        return rng.uniform(0, 4, (35, 20, 3)).astype(np.float32)
    pat = rng.choice(0x7C00, size=200 * 32 * 3, replace=False).astype(np.uint16)         # distinct finite non-negative halves
    top = pat.view(np.float16).astype(np.float32).reshape(32, 200, 3)
    return np.concatenate([top, rng.uniform(0, 1, (5, 200, 3)).astype(np.float32)], axis=0)


def make_piz():
    """PIZ-compressed OpenEXR maps: written by tests/exr_piz.py (the reference's tinyexr has its PIZ encoder compiled out), read back by the
    reference's own `-e` code path (oracle/exr_probe.cpp on the reference's vendored tinyexr); what THAT returns is the expected output."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import struct
    import exr_piz
    for name, (pt, _) in PIZ_FILES.items():
        img = piz_image(name)
        path = os.path.join(HERE, name + ".exr")
        exr_piz.write_exr_piz(path, img, pt)
        out = tmpf(name + ".bin")
        run([os.path.join(REF, "exr_probe"), path, out])
        raw = open(out, "rb").read()
        w, h = struct.unpack("<ii", raw[:8])
        got = np.frombuffer(raw[8:], np.float32).reshape(h, w, 3)
        want = img.astype(np.float16).astype(np.float32) if pt == 1 else img
        assert np.array_equal(got, want), name        # lossless codec: the reference decodes what was written
        store = got.astype(np.float16) if pt == 1 else got
        assert np.array_equal(store.astype(np.float32), got)
        np.savez_compressed(os.path.join(HERE, name + ".expected.npz"), rgb=store)
        print("%s: %d x %d, %d bytes, reference reader agrees" % (name, w, h, os.path.getsize(path)))


if __name__ == "__main__":
    what = sys.argv[1:] or ["scenes", "kat", "render"]
    if "piz" in what:
        make_piz()
    if "scenes" in what:
        make_scenes([w for w in what if w in SCENES] or None)
    if "kat" in what:
        make_kat([w for w in what if w in ("thin_lens", "keenan")])
    if "render" in what:
        make_render([w for w in what if w in RENDERS] or None)
    if "bvhorder" in what:
        kat_bvh_order()
    if "rng" in what:
        render_independent_rng()
    if "fullsize" in what:
        make_fullsize([w for w in what if w in FULLSIZE] or None)
