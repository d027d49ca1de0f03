"""Wall-clock time of rrt_upload_scene (BVH build + uploads) for the C5 scene and the bunny, host vs device builder."""
import importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench
api = importlib.import_module("relativistic-ray-tracer_b200.api")
for wl in ("c5", "bunny_bh"):
    s = bench.workload_scene(wl)
    for name, b in (("host", 2), ("device", 1)):
        c = api.Context(0); c.set_option(api.OPT_BVH_BUILDER, b)
        ts = []
        for _ in range(3):
            t = time.perf_counter(); c.upload_scene(s); ts.append(time.perf_counter() - t)
        print("%-9s %7d primitives  builder %-6s  rrt_upload_scene %.1f ms (first call %.1f ms)  reference tree: %s" %
              (wl, s.n_prims, name, min(ts) * 1e3, ts[0] * 1e3, {k: v for k, v in c.bvh_info().items() if k != "dfs_rank"}), flush=True)
        c.close()
