"""What would sorting the rays of a launch be worth?    Builds the ray populations of the deeper launches of a
frame by hand, through the parity hooks — camera rays (32 samples per pixel, pixel-major, like the path slots), their first hits, one
cosine-weighted bounce, the hits of that, and 4 shadow rays per second vertex towards the area light — and traces the SAME rays through
rrt_trace_rays in several orders: as the wavefront would queue them, and sorted by Morton code of the origin at several cell sizes, with and
without the direction octant in front.  The sort itself is done on the host and costs nothing here: the times are upper bounds of the gain.
usage: sort_probe.py [workload, default bunny_box] [pixels, default 32768]"""
import importlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench
from util import sc
api = importlib.import_module("relativistic-ray-tracer_b200.api")

wl = sys.argv[1] if len(sys.argv) > 1 else "bunny_box"
npx = int(sys.argv[2]) if len(sys.argv) > 2 else 32768
fixture, _, W, H, spp, l, m, mode, desc = bench.WORKLOADS[wl]
scene = bench.workload_scene(wl)
c = api.Context(0); c.upload_scene(scene); c.set_frame(W, H)
c.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_LITERAL if mode == "bh" else sc.MODE_FLAT)
rng = np.random.default_rng(7)


def unit(v):
    return v / np.linalg.norm(v, axis=1, keepdims=True)


def cosine_about(n, w_out):
    """cosine-weighted direction about the normal, on w_out's side"""
    n = unit(n); n = np.where((np.sum(n * w_out, axis=1) < 0)[:, None], -n, n)
    a = np.where(np.abs(n[:, :1]) > 0.9, np.array([[0.0, 1.0, 0.0]]), np.array([[1.0, 0.0, 0.0]]))
    t = unit(np.cross(n, a)); b = np.cross(n, t)
    u1 = rng.uniform(0, 1, n.shape[0]); u2 = rng.uniform(0, 2 * np.pi, n.shape[0])
    r = np.sqrt(u1)
    return unit(t * (r * np.cos(u2))[:, None] + b * (r * np.sin(u2))[:, None] + n * np.sqrt(1 - u1)[:, None])


def morton(p, bits, lo, hi):
    q = np.clip(((p - lo) / (hi - lo) * (1 << bits)).astype(np.int64), 0, (1 << bits) - 1)
    key = np.zeros(p.shape[0], np.int64)
    for b in range(bits):
        for a in range(3):
            key |= ((q[:, a] >> b) & 1) << (3 * b + a)
    return key


def octant(d):
    return ((d[:, 0] < 0).astype(np.int64) | ((d[:, 1] < 0).astype(np.int64) << 1) | ((d[:, 2] < 0).astype(np.int64) << 2))


def timed(o, d, any_hit, order):
    oo = np.ascontiguousarray(o[order]); dd = np.ascontiguousarray(d[order])
    best = None
    for _ in range(3):
        h, st = c.trace_rays(oo, dd, any_hit=any_hit)
        if best is None or st["trace_seconds"] < best[1]["trace_seconds"]:
            best = (h, st)
    back = np.empty_like(best[0]); back[order] = best[0]
    return back, best[1]


# camera rays: npx pixels, 32 samples each, pixel-major
pix = rng.choice(W * H, size=min(npx, W * H), replace=False)
px = np.repeat(pix % W, 32); py = np.repeat(pix // W, 32)
xy = np.stack([(px + rng.uniform(0, 1, px.size)) / W, (py + rng.uniform(0, 1, py.size)) / H], axis=1)
o0, d0 = c.generate_rays(xy)
h0, _ = c.trace_rays(o0, d0)
k0 = h0["status"] == 1
w1 = cosine_about(h0["n"][k0], h0["w_out"][k0])
o1 = h0["hit_p"][k0] + 1e-11 * w1                      # extension rays of vertex 0: origins coherent (same pixel), directions random
h1, _ = c.trace_rays(o1, w1)
k1 = h1["status"] == 1
w2 = cosine_about(h1["n"][k1], h1["w_out"][k1])
o2 = h1["hit_p"][k1] + 1e-11 * w2                      # extension rays of vertex 1: origins and directions random
L = [x for x in scene.lights if int(x["type"]) == sc.LIGHT_AREA][0] if hasattr(sc, "LIGHT_AREA") else scene.lights[0]
P1 = np.repeat(h1["hit_p"][k1], 4, axis=0)             # 4 shadow rays per second vertex, vertex-major
uv = rng.uniform(-0.5, 0.5, (P1.shape[0], 2))
tgt = np.asarray(L["position"])[None, :] + uv[:, :1] * np.asarray(L["dim_x"])[None, :] + uv[:, 1:] * np.asarray(L["dim_y"])[None, :]
ws = unit(tgt - P1); os_ = P1 + 1e-11 * ws

P0 = np.repeat(h0["hit_p"][k0], 4, axis=0)             # and per first vertex: coherent as queued (the samples of a pixel sit side by side)
uv0 = rng.uniform(-0.5, 0.5, (P0.shape[0], 2))
tgt0 = np.asarray(L["position"])[None, :] + uv0[:, :1] * np.asarray(L["dim_x"])[None, :] + uv0[:, 1:] * np.asarray(L["dim_y"])[None, :]
ws0 = unit(tgt0 - P0); os0 = P0 + 1e-11 * ws0

lo = scene.tri_pos.reshape(-1, 3).min(axis=0) - 1e-3; hi = scene.tri_pos.reshape(-1, 3).max(axis=0) + 1e-3   # the scene box: 2 x 1.5 x 2 m, cells of 8 mm at 8 bits
for name, o, d, any_hit in (("shadow rays of vertex 0", os0, ws0, True), ("extension rays of vertex 0", o1, w1, False), ("extension rays of vertex 1", o2, w2, False),
                            ("shadow rays of vertex 1", os_, ws, True)):
    n = o.shape[0]
    orders = [("as queued", np.arange(n)), ("shuffled", rng.permutation(n))]
    for bits in (4, 6, 8):
        orders.append(("origin morton %d bits/axis" % bits, np.argsort(morton(o, bits, lo, hi), kind="stable")))
    orders.append(("octant, then origin morton 8", np.argsort((octant(d) << 24) | morton(o, 8, lo, hi), kind="stable")))
    orders.append(("origin morton 4, then octant, then morton 8", np.argsort((morton(o, 4, lo, hi) << 27) | (octant(d) << 24) | morton(o, 8, lo, hi), kind="stable")))
    ref = None
    for oname, order in orders:
        h, st = timed(o, d, any_hit, order)
        key = h["status"].tobytes() + (b"" if any_hit else h["prim"].tobytes() + h["hit_p"].tobytes())
        if ref is None:
            ref = key
        print(json.dumps(dict(workload=wl, rays=name, n=int(n), order=oname, trace_ms=round(st["trace_seconds"] * 1e3, 3), chords=st["geodesic_steps"],
                              gchords_s=round(st["geodesic_steps"] / st["trace_seconds"] / 1e9, 2), nodes=st["node_visits"], same_results=bool(key == ref))), flush=True)
