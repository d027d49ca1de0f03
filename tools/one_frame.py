"""Renders N frames of a bench workload with counters off (for ncu captures).  usage: one_frame.py workload [fp64|fp32] [frames]"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench
from util import sc
api = importlib.import_module("relativistic-ray-tracer_b200.api")
wl = sys.argv[1]; stepper = sys.argv[2] if len(sys.argv) > 2 else "fp64"; frames = int(sys.argv[3]) if len(sys.argv) > 3 else 2
fixture, _, W, H, spp, l, m, mode, desc = bench.WORKLOADS[wl]
c = api.Context(0)
c.set_option(api.OPT_STEPPER, 1 if stepper == "fp32" else 0); c.set_option(api.OPT_COUNTERS, 0)
c.upload_scene(bench.workload_scene(wl)); c.set_frame(W, H)
c.set_blackhole((0.0, 1.0, 0.0), 0.1, 0.1, sc.MODE_LITERAL if mode == "bh" else sc.MODE_FLAT)
ad = bench.ADAPTIVE.get(wl)
p = sc.make_params(ns_aa=spp, max_ray_depth=m, ns_area_light=l, samples_per_batch=ad["batch"] if ad else max(32, spp + 1), max_tolerance=ad["tol"] if ad else 0.05, seed=1)
for _ in range(frames):
    hdr, cnt, st = c.render(p)
print(desc, stepper, "ms", st["seconds"] * 1e3, "trace launches", st["trace_launches"])
