"""ctypes binding of oracle/librrt_oracle.so — TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module.
It mirrors the call shapes of the product binding (relativistic-ray-tracer_b200/api.py) so parity tests can run
the same inputs through both.
"""
import ctypes as C
import importlib
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_sc = importlib.import_module("relativistic-ray-tracer_b200.scene")
_LIB = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "librrt_oracle.so"])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "librrt_oracle.so")
        if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(os.path.join(_HERE, "rrt_oracle.c")):
            build()
        L = C.CDLL(path)
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.POINTER(_sc.Scene), C.POINTER(_sc.Camera)]
        for name in ("orc_destroy", "orc_set_camera", "orc_set_blackhole", "orc_set_frame", "orc_set_cell", "orc_set_tile_shard",
                     "orc_bvh_info", "orc_geodesic", "orc_trace_rays", "orc_generate_rays", "orc_generate_rays_thin_lens", "orc_philox", "orc_tonemap", "orc_set_envmap"):
            getattr(L, name).restype = None
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    def __init__(self, scene):
        self.scene = scene
        self._cs, self._cc = scene.c_scene(), scene.c_camera()
        self.h = C.c_void_p(lib().orc_create(C.byref(self._cs), C.byref(self._cc)))
        self.W = self.H = 0

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_destroy(self.h)
            self.h = None

    def set_blackhole(self, centre=(0.0, 1.0, 0.0), r=0.1, dtheta=0.1, mode=_sc.MODE_LITERAL):
        lib().orc_set_blackhole(self.h, (C.c_double * 3)(*centre), C.c_double(r), C.c_double(dtheta), C.c_int(mode))

    def set_frame(self, w, h):
        self.W, self.H = int(w), int(h)
        lib().orc_set_frame(self.h, C.c_int(w), C.c_int(h))

    def set_cell(self, x, y, w, h):
        lib().orc_set_cell(self.h, C.c_int(x), C.c_int(y), C.c_int(w), C.c_int(h))

    def set_tile_shard(self, tile, rank, world):
        lib().orc_set_tile_shard(self.h, C.c_int(tile), C.c_int(rank), C.c_int(world))

    def set_envmap(self, rgb):
        """rgb: [h, w, 3] float32, row 0 = top (None clears)"""
        if rgb is None:
            lib().orc_set_envmap(self.h, C.c_int(0), C.c_int(0), None)
            return
        a = np.ascontiguousarray(rgb, np.float32)
        lib().orc_set_envmap(self.h, C.c_int(a.shape[1]), C.c_int(a.shape[0]), _p(a))

    def env_lookup(self, d):
        d = np.ascontiguousarray(d, np.float64).reshape(-1, 3)
        out = np.zeros((d.shape[0], 3), np.float32)
        assert lib().orc_env_lookup(self.h, C.c_int64(d.shape[0]), _p(d), _p(out)) == 0
        return out

    def chords_per_ray(self):
        return int(lib().orc_chords_per_ray(self.h))

    def bvh_info(self):
        nn, nl, dp = C.c_int32(), C.c_int32(), C.c_int32()
        rank = np.zeros(self.scene.n_prims, np.uint32)
        lib().orc_bvh_info(self.h, C.byref(nn), C.byref(nl), C.byref(dp), _p(rank))
        return dict(n_nodes=nn.value, n_leaves=nl.value, depth=dp.value, dfs_rank=rank)

    def geodesic(self, o, d, nsteps):
        o = np.ascontiguousarray(o, np.float64).reshape(-1, 3); d = np.ascontiguousarray(d, np.float64).reshape(-1, 3)
        out = np.zeros((o.shape[0], nsteps, 8), np.float64)
        lib().orc_geodesic(self.h, C.c_int64(o.shape[0]), _p(o), _p(d), C.c_int(nsteps), _p(out))
        return out

    def trace_rays(self, o, d, min_t=None, max_t=None, any_hit=False):
        o = np.ascontiguousarray(o, np.float64).reshape(-1, 3); d = np.ascontiguousarray(d, np.float64).reshape(-1, 3)
        n = o.shape[0]
        mn = None if min_t is None else np.ascontiguousarray(min_t, np.float64)
        mx = None if max_t is None else np.ascontiguousarray(max_t, np.float64)
        out = np.zeros(n, _sc.HIT_DTYPE)
        st = _sc.Stats()
        lib().orc_trace_rays(self.h, C.c_int64(n), _p(o), _p(d), None if mn is None else _p(mn), None if mx is None else _p(mx),
                             C.c_int(int(any_hit)), _p(out), C.byref(st))
        return out, st.as_dict()

    def generate_rays(self, xy):
        xy = np.ascontiguousarray(xy, np.float64).reshape(-1, 2)
        o = np.zeros((xy.shape[0], 3)); d = np.zeros((xy.shape[0], 3))
        lib().orc_generate_rays(self.h, C.c_int64(xy.shape[0]), _p(xy), _p(o), _p(d))
        return o, d

    def generate_rays_thin_lens(self, xy, lens):
        xy = np.ascontiguousarray(xy, np.float64).reshape(-1, 2); lens = np.ascontiguousarray(lens, np.float64).reshape(-1, 2)
        o = np.zeros((xy.shape[0], 3)); d = np.zeros((xy.shape[0], 3))
        lib().orc_generate_rays_thin_lens(self.h, C.c_int64(xy.shape[0]), _p(xy), _p(lens), _p(o), _p(d))
        return o, d

    def bsdf_sample(self, mat, wo, u):
        wo = np.ascontiguousarray(wo, np.float64).reshape(-1, 3); u = np.ascontiguousarray(u, np.float64).reshape(-1, 4)
        n = wo.shape[0]
        f = np.zeros((n, 3), np.float32); wi = np.zeros((n, 3)); pdf = np.zeros(n, np.float32)
        rc = lib().orc_bsdf_sample(self.h, C.c_int(mat), C.c_int64(n), _p(wo), _p(u), _p(f), _p(wi), _p(pdf))
        assert rc == 0
        return f, wi, pdf

    def bsdf_eval(self, mat, wo, wi):
        wo = np.ascontiguousarray(wo, np.float64).reshape(-1, 3); wi = np.ascontiguousarray(wi, np.float64).reshape(-1, 3)
        f = np.zeros((wo.shape[0], 3), np.float32)
        rc = lib().orc_bsdf_eval(self.h, C.c_int(mat), C.c_int64(wo.shape[0]), _p(wo), _p(wi), _p(f))
        assert rc == 0
        return f

    def light_sample(self, light, p, u):
        p = np.ascontiguousarray(p, np.float64).reshape(-1, 3); u = np.ascontiguousarray(u, np.float64).reshape(-1, 2)
        n = p.shape[0]
        rad = np.zeros((n, 3), np.float32); wi = np.zeros((n, 3)); dist = np.zeros(n, np.float32); pdf = np.zeros(n, np.float32)
        rc = lib().orc_light_sample(self.h, C.c_int(light), C.c_int64(n), _p(p), _p(u), _p(rad), _p(wi), _p(dist), _p(pdf))
        assert rc == 0
        return rad, wi, dist, pdf

    def render(self, params):
        assert self.W and self.H
        hdr = np.zeros((self.H, self.W, 3), np.float32); cnt = np.zeros((self.H, self.W), np.int32)
        st = _sc.Stats()
        rc = lib().orc_render(self.h, C.byref(params), _p(hdr), _p(cnt), C.byref(st))
        assert rc == 0, rc
        return hdr, cnt, st.as_dict()


def philox(seed, counters):
    counters = np.ascontiguousarray(counters, np.uint32).reshape(-1, 4)
    out = np.zeros_like(counters)
    lib().orc_philox(C.c_uint64(seed), C.c_int64(counters.shape[0]), _p(counters), _p(out))
    return out


def tonemap(hdr):
    hdr = np.ascontiguousarray(hdr, np.float32).reshape(-1, 3)
    out = np.zeros(hdr.shape[0], np.uint32)
    lib().orc_tonemap(C.c_int64(hdr.shape[0]), _p(hdr), _p(out))
    return out
