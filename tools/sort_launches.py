"""One frame of a workload with RRT_OPT_SORT_SHADOW off, then on, counters off: run under `ncu --metrics gpu__time_duration.sum` for the per-launch
durations of the trace kernels and of the three sort kernels (ncu serialises the launches: durations are comparable one to one).
usage: sort_launches.py [workload, default bunny_box]"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench
from util import sc
api = importlib.import_module("relativistic-ray-tracer_b200.api")
wl = sys.argv[1] if len(sys.argv) > 1 else "bunny_box"
fixture, _, W, H, spp, l, m, mode, desc = bench.WORKLOADS[wl]
scene = bench.workload_scene(wl)
for sort in (0, 1):
    c = api.Context(0)
    c.set_option(api.OPT_SORT_SHADOW, sort); c.set_option(api.OPT_COUNTERS, 0)
    c.upload_scene(scene); c.set_frame(W, H)
    c.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_LITERAL if mode == "bh" else sc.MODE_FLAT)
    p = sc.make_params(ns_aa=spp, max_ray_depth=m, ns_area_light=l, seed=3, samples_per_batch=spp + 1)
    _, _, st = c.render(p)
    print("sort", sort, "frame ms", round(st["seconds"] * 1e3, 2), flush=True)
    c.close()
