"""A/B of run-time options on a GPU box: frame time (counters off, best of 3 after a warm-up) and the instrumented counts,
for the C2 frame and the bunny at 480x360, hole on.  Images must be identical across every bit-exact option set.
usage: ab_options.py [spp]"""
import importlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from util import load_scene, sc
api = importlib.import_module("relativistic-ray-tracer_b200.api")
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 16
SETS = [("default", {}), ("fp32 stepper", {api.OPT_STEPPER: 1}), ("binary nodes", {api.OPT_BVH_WIDTH: 2}),
        ("binary nodes, fp32 stepper", {api.OPT_BVH_WIDTH: 2, api.OPT_STEPPER: 1})]
for scene, W, H in (("cbspheres_lambertian_480x360", 480, 360), ("cbbunny_microfacet_480x360", 480, 360)):
    base = None
    for name, opts in SETS:
        c = api.Context(0)
        for k, v in opts.items():
            c.set_option(k, v)
        c.upload_scene(load_scene(scene)); c.set_frame(W, H); c.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_LITERAL)
        p = sc.make_params(ns_aa=spp, max_ray_depth=5, ns_area_light=4, seed=3, samples_per_batch=spp + 1)
        hdr, _, st = c.render(p)
        c.set_option(api.OPT_COUNTERS, 0)
        ms = min(c.render(p)[2]["seconds"] for _ in range(3)) * 1e3
        exact = api.OPT_STEPPER not in opts
        if exact:
            if base is None:
                base = hdr
            assert np.array_equal(base, hdr), "image changed under " + name
        print(json.dumps(dict(scene=scene, options=name, ms=round(ms, 3), rays=st["rays"], chords=st["geodesic_steps"], nodes=st["node_visits"],
                              tri_records=st["tri_tests"], sphere_tests=st["sphere_tests"], alg_bytes=st["algorithmic_bytes"],
                              identical_to_default=bool(exact), rmse_over_mean=float(np.sqrt(np.mean((hdr - base) ** 2)) / base.mean()))), flush=True)
        c.close()
