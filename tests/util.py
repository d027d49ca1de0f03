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
