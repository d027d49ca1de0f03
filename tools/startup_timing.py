"""Where a cold start goes: library load, rrt_create (CUDA context), rrt_upload_scene, first and second rrt_render of a tiny frame."""
import importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
t0 = time.perf_counter()
from util import load_scene, sc
api = importlib.import_module("relativistic-ray-tracer_b200.api")
api.lib(); t1 = time.perf_counter()
c = api.Context(0); t2 = time.perf_counter()
s = load_scene("cbspheres_lambertian_96x72")
c.upload_scene(s); t3 = time.perf_counter()
c.set_frame(96, 72)
p = sc.make_params(ns_aa=4, max_ray_depth=5, ns_area_light=4, seed=1)
c.render(p); t4 = time.perf_counter()
c.render(p); t5 = time.perf_counter()
c.set_frame(3840, 2160); c.render(p); t6 = time.perf_counter()
c.render(p); t7 = time.perf_counter()
print("imports + dlopen %.0f ms | rrt_create %.0f ms | upload %.0f ms | first render 96x72 %.0f ms | second %.0f ms | first 4K render %.0f ms | second %.0f ms" %
      tuple(1e3 * x for x in (t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4, t6 - t5, t7 - t6)))
