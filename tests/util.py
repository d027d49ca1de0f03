"""Shared helpers for the tests: golden fixtures, scene loading, the two bindings."""
import importlib
import io
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

pkg = importlib.import_module("relativistic-ray-tracer_b200")
sc = importlib.import_module("relativistic-ray-tracer_b200.scene")
RAND_MAX = 2147483647


def load_scene(name):
    """name like 'cbspheres_lambertian_480x360' or '..._96x72' (camera block over the 480x360 geometry)."""
    p = os.path.join(GOLDEN, name + ".rrtscene")
    if os.path.exists(p):
        return sc.SceneData.load(p)
    if os.path.exists(p + ".npz"):
        raw = np.load(p + ".npz")["raw"].tobytes()
        tmp = os.path.join("/tmp", "rrt_%d_%s.rrtscene" % (os.getpid(), name))
        with open(tmp, "wb") as f:
            f.write(raw)
        s = sc.SceneData.load(tmp)
        os.remove(tmp)
        return s
    cam = os.path.join(GOLDEN, name + ".cam.npy")
    if os.path.exists(cam):
        base = name.rsplit("_", 1)[0] + "_480x360"
        s = load_scene(base)
        v = np.load(cam)
        s.camera = dict(pos=v[0:3].tolist(), c2w=v[3:12].tolist(), hFov=float(v[12]), vFov=float(v[13]), nClip=float(v[14]),
                        fClip=float(v[15]), lensRadius=float(v[16]), focalDistance=float(v[17]), screenW=float(v[18]),
                        screenH=float(v[19]), screenDist=float(v[20]))
        return s
    raise FileNotFoundError(name)


def golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)


def hits_from_probe(rows):
    """ref_probe trace rows [hit, prim, hit_p3, w_out3, n3, anyhit, ...] -> dict of arrays"""
    return dict(hit=rows[:, 0] > 0, prim=rows[:, 1].astype(np.int64), hit_p=rows[:, 2:5], w_out=rows[:, 5:8], n=rows[:, 8:11],
                any=rows[:, 11] > 0)


def rel_err(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return np.abs(a - b) / np.maximum(np.abs(b), 1e-300)


def grazing_rays(scene, n, seed):
    """Rays aimed at the places where a conservative FP32 slab test and the reference's double slab test (src/bbox.cpp:10-25) could
    disagree: every face of every BVH node box is a vertex coordinate, so rays are sent (A) through triangle vertices, (B) through points
    of triangle edges, (C) INSIDE the axis-aligned plane of a vertex coordinate (origin coordinate == vertex coordinate exactly, that
    direction component exactly 0: the reference's slab quotient is 0/0 there), (D) from origins exactly on a vertex, (E) exactly
    axis-parallel through a vertex.  Flat mode, [0, inf).  Returns (o, d) float64 [n, 3]."""
    rng = np.random.default_rng(seed)
    tri = scene.tri_pos.reshape(-1, 3, 3)
    nt = tri.shape[0]
    m = n // 5
    def origins(k):
        return np.stack([rng.uniform(-0.95, 0.95, k), rng.uniform(0.05, 1.9, k), rng.uniform(-0.9, 3.0, k)], 1)
    def unit(v):
        return v / np.linalg.norm(v, axis=1, keepdims=True)
    def verts(k):
        return tri[rng.integers(0, nt, k), rng.integers(0, 3, k)]
    O, D = [], []
    v = verts(m); o = origins(m); O.append(o); D.append(unit(v - o))                                   # A
    t = tri[rng.integers(0, nt, m)]; a = rng.integers(0, 3, m); s = rng.uniform(0, 1, (m, 1))
    pa = t[np.arange(m), a]; pb = t[np.arange(m), (a + 1) % 3]
    v = pa + s * (pb - pa); o = origins(m); O.append(o); D.append(unit(v - o))                          # B
    v = verts(m); o = origins(m); k = rng.integers(0, 3, m)
    o[np.arange(m), k] = v[np.arange(m), k]
    d = v - o; d[np.arange(m), k] = 0.0
    O.append(o); D.append(unit(d))                                                                      # C
    v = verts(m); O.append(v.copy()); D.append(unit(rng.normal(size=(m, 3))))                           # D
    r = n - 4 * m
    v = verts(r); k = rng.integers(0, 3, r); sgn = rng.choice([-1.0, 1.0], r)
    d = np.zeros((r, 3)); d[np.arange(r), k] = sgn
    o = v.copy(); o[np.arange(r), k] -= sgn * rng.uniform(0.1, 2.0, r)
    O.append(o); D.append(d)                                                                            # E
    return np.ascontiguousarray(np.concatenate(O)), np.ascontiguousarray(np.concatenate(D))
