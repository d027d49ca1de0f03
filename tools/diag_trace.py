"""Diagnostic: time rrt_trace_rays on the camera rays of a frame for several scenes (GPU box)."""
import importlib, sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from util import load_scene, sc
api = importlib.import_module("relativistic-ray-tracer_b200.api")
W, H = int(sys.argv[1]) if len(sys.argv) > 1 else 1920, int(sys.argv[2]) if len(sys.argv) > 2 else 1080
for name in ("cbspheres_lambertian_480x360", "cbbunny_microfacet_480x360", "cbempty_480x360"):
    s = load_scene(name); s.set_screen_size(W, H)
    for topo in (1, 0):
        c = api.Context(0); c.set_option(api.OPT_BVH_TOPOLOGY, topo); c.upload_scene(s)
        rng = np.random.default_rng(1)
        n = 1 << 20
        xy = rng.uniform(0, 1, (n, 2))
        o, d = c.generate_rays(xy)
        for rep in range(2):
            hits, st = c.trace_rays(o, d)
        status = np.bincount(hits["status"], minlength=3)
        print("%-32s topo %d  %7.2f ms  %6.3f Gsteps/s  steps/ray %5.1f  nodes/step %5.2f  tris/step %5.2f  status %s" % (
            name, topo, st["trace_seconds"] * 1e3, st["geodesic_steps"] / st["trace_seconds"] / 1e9, st["geodesic_steps"] / n,
            st["node_visits"] / st["geodesic_steps"], st["tri_tests"] / st["geodesic_steps"], status.tolist()), flush=True)
        # per-status step histogram
        for k in range(3):
            m = hits["status"] == k
            if m.any():
                print("      status %d: mean steps %.1f" % (k, hits["steps"][m].mean()))
