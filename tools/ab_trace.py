"""A/B of trace-kernel options on a GPU box: per (workload, option set) the frame time with counters off (best of 3 after a
warm-up), trace time, and the instrumented counts.  Images must be identical across every option set of a workload.
usage: ab_trace.py workload[,workload...] "name:opt=val,opt=val;name2:..."     (opt = kernel | grid | width | builder | leaf | topology | streams | overlap | top | escape | sort | shard=N: rank 0 of N | m=N: max ray depth | inflight=N: paths per chunk | workgb=N: work-buffer budget)
workloads: c2, bunny_box (bunny at 480x360: the in-box pixels of the 1080p frame), bunny_bh, c5, or any bench.py workload; [@spp]"""
import hashlib, importlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench
from util import sc
api = importlib.import_module("relativistic-ray-tracer_b200.api")
OPT = dict(kernel=api.OPT_TRACE_KERNEL, grid=api.OPT_CLEARANCE_GRID, width=api.OPT_BVH_WIDTH, builder=api.OPT_BVH_BUILDER, leaf=api.OPT_LEAF_SIZE,
           topology=api.OPT_BVH_TOPOLOGY, stepper=api.OPT_STEPPER, streams=api.OPT_STREAMS, escape=api.OPT_ESCAPE, top=api.OPT_TOP_NODE, overlap=api.OPT_OVERLAP, sort=api.OPT_SORT_SHADOW)
sets = []
for part in (sys.argv[2] if len(sys.argv) > 2 else "default:").split(";"):
    name, _, kv = part.partition(":")
    kvs = [x.split("=") for x in kv.split(",") if x]
    sets.append((name, {OPT[k]: int(v) for k, v in kvs if k in OPT}, {k: int(v) for k, v in kvs if k not in OPT}))
for wl in sys.argv[1].split(","):
    wl, _, spp_s = wl.partition("@")
    fixture, _, W, H, spp, l, m, mode, desc = bench.WORKLOADS[wl]
    spp = int(spp_s) if spp_s else spp
    scene = bench.workload_scene(wl)
    base = None
    for name, opts, extra in sets:
        c = api.Context(0)
        for k, v in opts.items():
            c.set_option(k, v)
        import time
        t0 = time.perf_counter(); c.upload_scene(scene); up = time.perf_counter() - t0
        c.set_frame(W, H)
        if extra.get("workgb"):
            c.set_option(api.OPT_WORK_BYTES, extra["workgb"] << 30)
        if extra.get("shard"):
            c.set_tile_shard(8, 0, extra["shard"])     # rank 0's share of an N-GPU strong-scaling run, on one GPU
        c.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_LITERAL if mode == "bh" else sc.MODE_FLAT)
        ad = bench.ADAPTIVE.get(wl) if not spp_s else None
        p = sc.make_params(ns_aa=spp, max_ray_depth=extra.get("m", m), ns_area_light=l, seed=3, samples_per_batch=ad["batch"] if ad else spp + 1, max_tolerance=ad["tol"] if ad else 0.05,
                           max_paths_in_flight=extra.get("inflight", 0))
        hdr, _, st = c.render(p)
        c.set_option(api.OPT_COUNTERS, 0)
        best = min((c.render(p)[2] for _ in range(3)), key=lambda s: s["seconds"])
        exact = api.OPT_STEPPER not in opts
        exact = exact and "m" not in extra
        if exact:
            if base is None:
                base = hdr
            assert np.array_equal(base, hdr), "image changed under " + name
        print(json.dumps(dict(workload=wl, spp=spp, options=name, ms=round(best["seconds"] * 1e3, 3), trace_ms=round(best["trace_seconds"] * 1e3, 3),
                              mpaths_s=round(st["paths"] / best["seconds"] / 1e6, 1), upload_s=round(up, 3), rays=st["rays"], chords=st["geodesic_steps"],
                              nodes=st["node_visits"], tri_records=st["tri_tests"], sphere_tests=st["sphere_tests"],
                              identical=bool(exact), sha=hashlib.sha256(np.ascontiguousarray(hdr).tobytes()).hexdigest()[:16])), flush=True)
        c.close()
