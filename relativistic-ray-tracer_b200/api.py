"""ctypes binding of csrc/librrt_b200.so (include/rrt_b200.h) and a Python mirror of the reference's
PathTracer / BVHAccel interface for the hot path.

This module is plumbing around the C-ABI: no arithmetic of the path happens here.  There is no CPU fallback —
if the CUDA library is missing or no device is present every entry point raises.

Reference interface mirrored (paths relative to /root/reference/pathtracer):
  PathTracer   src/pathtracer.h:64-167   (ctor, set_scene, set_camera, set_frame_size, start_raytracing,
                                          render_to_file, stop, clear, save_image, state machine :229-235)
  BVHAccel     src/bvh.h:50-144          (ctor from the primitive list, get_bbox, intersect x2, total_rays/total_isects)
"""
import ctypes as C
import os
import struct
import threading
import zlib

import numpy as np

from . import scene as _sc

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RRT_B200_LIB") or os.path.join(_HERE, "csrc", "librrt_b200.so")  # env override: A/B builds only
_LIB = None

OPT_BVH_TOPOLOGY, OPT_COUNTERS, OPT_WORK_BYTES, OPT_TRACE_KERNEL, OPT_LEAF_SIZE, OPT_CLEARANCE_GRID, OPT_BVH_WIDTH, OPT_SHADE_SPLIT = 0, 1, 2, 3, 4, 5, 6, 7
OPT_BVH_BUILDER = 9  # 0 = auto (device for >= 4096 primitives), 1 = device (reference ranks + LBVH), 2 = host
OPT_ESCAPE = 11   # escape hand-off: rays outside the scene box continue in a lean kernel (default 1)
OPT_TOP_NODE = 12  # 8-wide record for the top of a binary tree (default 1)
OPT_OVERLAP = 13  # shadow rays of depth i beside extension rays of depth i + 1 (default 1)
OPT_SORT_SHADOW = 14  # hole on, meshes: shadow rays handed to the trace kernel cell by cell of their origin (1 = chunks < 2 M paths, default; 2 = always; 0 = never)
OPT_STREAMS = 10  # wavefront chunks of a round dealt over this many CUDA streams (default 2)
OPT_STEPPER = 8   # 0 = FP64 bit-exact stepper (default), 1 = FP32 state-carrying stepper (include/rrt_b200.h)

EXPORTS = ["rrt_abi_version", "rrt_device_count", "rrt_create", "rrt_destroy", "rrt_last_error", "rrt_set_option", "rrt_upload_scene",
           "rrt_set_camera", "rrt_set_blackhole", "rrt_set_frame", "rrt_set_cell", "rrt_set_tile_shard", "rrt_render", "rrt_render_device",
           "rrt_trace_rays", "rrt_geodesic", "rrt_generate_rays", "rrt_generate_rays_thin_lens", "rrt_set_progress", "rrt_snapshot", "rrt_bsdf_sample", "rrt_bsdf_eval", "rrt_light_sample", "rrt_philox",
           "rrt_tonemap", "rrt_bvh_info", "rrt_bvh_build_host", "rrt_chords_per_ray", "rrt_set_envmap", "rrt_env_lookup", "rrt_trace_rays_static", "rrt_create_multi", "rrt_gpu_count"]
STATIC_MAX_RAYS = 65536  # RRT_STATIC_MAX_RAYS


PROGRESS_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int32, C.c_int32, C.c_int64)


class RrtError(RuntimeError):
    pass


def lib():
    """Loads the product library.  Raises (never falls back) when it has not been built."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RrtError("%s is missing: build it with `make -C %s` (or __graft_entry__.build()); there is no CPU fallback"
                           % (LIB_PATH, os.path.dirname(LIB_PATH)))
        L = C.CDLL(LIB_PATH)
        L.rrt_last_error.restype = C.c_char_p
        L.rrt_last_error.argtypes = [C.c_void_p]
        L.rrt_create.argtypes = [C.POINTER(C.c_void_p), C.c_int]
        L.rrt_destroy.argtypes = [C.c_void_p]
        L.rrt_create_multi.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.POINTER(C.c_int)]
        L.rrt_gpu_count.argtypes = [C.c_void_p]
        L.rrt_destroy.restype = None
        L.rrt_set_option.argtypes = [C.c_void_p, C.c_int, C.c_int64]
        L.rrt_upload_scene.argtypes = [C.c_void_p, C.POINTER(_sc.Scene)]
        L.rrt_set_camera.argtypes = [C.c_void_p, C.POINTER(_sc.Camera)]
        L.rrt_set_blackhole.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.c_double, C.c_double, C.c_int]
        L.rrt_set_frame.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.rrt_set_cell.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
        L.rrt_set_tile_shard.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.rrt_render.argtypes = [C.c_void_p, C.POINTER(_sc.Params), C.c_void_p, C.c_void_p, C.POINTER(_sc.Stats)]
        L.rrt_render_device.argtypes = [C.c_void_p, C.POINTER(_sc.Params), C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(_sc.Stats)]
        L.rrt_trace_rays.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                     C.POINTER(_sc.Stats)]
        L.rrt_trace_rays_static.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.rrt_geodesic.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.rrt_generate_rays.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.rrt_set_progress.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.rrt_snapshot.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.rrt_generate_rays_thin_lens.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.rrt_bsdf_sample.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.rrt_bsdf_eval.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.rrt_light_sample.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.rrt_philox.argtypes = [C.c_void_p, C.c_uint64, C.c_int64, C.c_void_p, C.c_void_p]
        L.rrt_set_envmap.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.rrt_env_lookup.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        L.rrt_tonemap.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        L.rrt_bvh_info.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_void_p]
        L.rrt_chords_per_ray.argtypes = [C.c_void_p]
        L.rrt_bvh_build_host.argtypes = [C.POINTER(_sc.Scene), C.c_int, C.POINTER(C.c_int32), C.c_void_p]
        _LIB = L
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def bvh_build_host(scene, topology=0):
    """rrt_bvh_build_host: the host-side BVH construction, no device needed."""
    info = (C.c_int32 * 8)()
    rank = np.zeros(max(scene.n_prims, 1), np.uint32)
    cs = scene.c_scene()
    rc = lib().rrt_bvh_build_host(C.byref(cs), int(topology), info, _p(rank))
    if rc != 0:
        raise RrtError("rrt_bvh_build_host failed (%d)" % rc)
    return dict(n_nodes=info[0], n_leaves=info[1], depth=info[2], flat_nodes=info[3], leaf_records=info[4], root=info[5], flat_depth=info[6], stack=info[7],
                dfs_rank=rank[:scene.n_prims])


_NCCL = None


def _preload_nccl():
    """rrt_create_multi dlopen()s libnccl.so.2.  In a Python process that (later) imports torch, the NCCL torch ships (site-packages/nvidia/nccl) must be
    the one the process holds: the loader identifies libraries by SONAME, and an older system libnccl.so.2 loaded first would be handed to
    libtorch_cuda.so in place of its own (ImportError: undefined symbol ncclDevCommCreate).  So that copy, when present, is loaded first; the library
    then finds it already there.  RRT_NCCL_LIB overrides."""
    global _NCCL
    if _NCCL is not None or os.environ.get("RRT_NCCL_LIB"):
        return
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for base in (list(spec.submodule_search_locations) if spec and spec.submodule_search_locations else []):
            cand = os.path.join(base, "lib", "libnccl.so.2")
            if os.path.exists(cand):
                _NCCL = C.CDLL(cand, mode=os.RTLD_LOCAL)
                return
    except Exception:
        pass
    _NCCL = False


class Context:
    """One rrt_ctx, one host thread at a time: one GPU (device = int) or, with a list of devices, a multi-GPU context
    (rrt_create_multi: the frame's interleaved tiles are dealt over the GPUs, one NCCL reduce per frame onto the first)."""

    def __init__(self, device=0):
        self.h = C.c_void_p()
        if isinstance(device, (list, tuple)):
            _preload_nccl()
            devs = (C.c_int * len(device))(*[int(d) for d in device])
            rc = lib().rrt_create_multi(C.byref(self.h), len(device), devs)
        else:
            rc = lib().rrt_create(C.byref(self.h), int(device))
        if rc != 0:
            raise RrtError("rrt_create failed (%d): %s" % (rc, lib().rrt_last_error(None).decode()))
        self.W = self.H = 0
        self.scene = None

    def gpu_count(self):
        return int(lib().rrt_gpu_count(self.h))

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            lib().rrt_destroy(self.h)
            self.h = C.c_void_p()

    __del__ = close

    def _ck(self, rc):
        if rc != 0:
            raise RrtError("rrt call failed (%d): %s" % (rc, lib().rrt_last_error(self.h).decode()))

    def set_option(self, opt, value):
        self._ck(lib().rrt_set_option(self.h, int(opt), int(value)))

    def upload_scene(self, scene):
        self.scene = scene
        cs = scene.c_scene()
        self._ck(lib().rrt_upload_scene(self.h, C.byref(cs)))
        self.set_camera(scene)

    def set_camera(self, scene_or_camera):
        cc = scene_or_camera.c_camera() if hasattr(scene_or_camera, "c_camera") else scene_or_camera
        self._ck(lib().rrt_set_camera(self.h, C.byref(cc)))

    def set_blackhole(self, centre=(0.0, 1.0, 0.0), r=0.1, dtheta=0.1, mode=_sc.MODE_LITERAL):
        self._ck(lib().rrt_set_blackhole(self.h, (C.c_double * 3)(*[float(x) for x in centre]), float(r), float(dtheta), int(mode)))

    def set_frame(self, w, h):
        self.W, self.H = int(w), int(h)
        self._ck(lib().rrt_set_frame(self.h, self.W, self.H))

    def set_cell(self, x, y, w, h):
        self._ck(lib().rrt_set_cell(self.h, int(x), int(y), int(w), int(h)))

    def set_progress(self, fn):
        """fn(samples_done, samples_total, active_pixels) -> truthy to stop; called between adaptive batches on the
        rendering thread, may call self.snapshot().  None removes it."""
        if fn is None:
            self._progress = None
            self._ck(lib().rrt_set_progress(self.h, None, None))
            return
        def tramp(_user, done, total, active):
            try:
                return 1 if fn(int(done), int(total), int(active)) else 0
            except Exception:          # an exception cannot cross the C frame: stop the render instead
                import traceback; traceback.print_exc()
                return 1
        self._progress = PROGRESS_FN(tramp)
        self._ck(lib().rrt_set_progress(self.h, C.cast(self._progress, C.c_void_p), None))

    def snapshot(self):
        """running estimate (hdr [H, W, 3] float32, counts [H, W] int32); only inside a progress callback"""
        hdr = np.zeros((self.H, self.W, 3), np.float32); cnt = np.zeros((self.H, self.W), np.int32)
        self._ck(lib().rrt_snapshot(self.h, _p(hdr), _p(cnt)))
        return hdr, cnt

    def set_tile_shard(self, tile, rank, world):
        self._ck(lib().rrt_set_tile_shard(self.h, int(tile), int(rank), int(world)))

    def chords_per_ray(self):
        return int(lib().rrt_chords_per_ray(self.h))

    def bvh_info(self):
        nn, nl, dp = C.c_int32(), C.c_int32(), C.c_int32()
        rank = np.zeros(self.scene.n_prims, np.uint32)
        self._ck(lib().rrt_bvh_info(self.h, C.byref(nn), C.byref(nl), C.byref(dp), _p(rank)))
        return dict(n_nodes=nn.value, n_leaves=nl.value, depth=dp.value, dfs_rank=rank)

    # ---- the hot path
    def render(self, params, hdr=None, counts=None, want_stats=True):
        """rrt_render with HOST buffers (numpy, optionally pinned by the caller)."""
        if hdr is None:
            hdr = np.empty((self.H, self.W, 3), np.float32)
        if counts is None:
            counts = np.empty((self.H, self.W), np.int32)
        st = _sc.Stats()
        self._ck(lib().rrt_render(self.h, C.byref(params), _p(hdr), _p(counts), C.byref(st) if want_stats else None))
        return hdr, counts, st.as_dict()

    def render_device(self, params, d_rgbn_ptr, d_counts_ptr=0, stream=0, want_stats=True):
        """rrt_render_device: results stay in HBM (float4 rgbn per pixel at d_rgbn_ptr)."""
        st = _sc.Stats()
        self._ck(lib().rrt_render_device(self.h, C.byref(params), C.c_void_p(d_rgbn_ptr), C.c_void_p(d_counts_ptr or None),
                                         C.c_void_p(stream or None), C.byref(st) if want_stats else None))
        return st.as_dict()

    # ---- parity hooks
    def trace_rays(self, o, d, min_t=None, max_t=None, any_hit=False):
        o = np.ascontiguousarray(o, np.float64).reshape(-1, 3); d = np.ascontiguousarray(d, np.float64).reshape(-1, 3)
        n = o.shape[0]
        mn = None if min_t is None else np.ascontiguousarray(min_t, np.float64)
        mx = None if max_t is None else np.ascontiguousarray(max_t, np.float64)
        out = np.zeros(n, _sc.HIT_DTYPE)
        st = _sc.Stats()
        self._ck(lib().rrt_trace_rays(self.h, n, _p(o), _p(d), _p(mn), _p(mx), int(bool(any_hit)), _p(out), C.byref(st)))
        return out, st.as_dict()

    def trace_rays_static(self, o, d, min_t=None, max_t=None, any_hit=False):
        """rrt_trace_rays_static: the same kernel with all device scratch in the library's static array (no allocation in the call)."""
        o = np.ascontiguousarray(o, np.float64).reshape(-1, 3); d = np.ascontiguousarray(d, np.float64).reshape(-1, 3)
        n = o.shape[0]
        mn = None if min_t is None else np.ascontiguousarray(min_t, np.float64)
        mx = None if max_t is None else np.ascontiguousarray(max_t, np.float64)
        out = np.zeros(n, _sc.HIT_DTYPE)
        self._ck(lib().rrt_trace_rays_static(self.h, n, _p(o), _p(d), _p(mn), _p(mx), int(bool(any_hit)), _p(out)))
        return out

    def geodesic(self, o, d, nsteps):
        o = np.ascontiguousarray(o, np.float64).reshape(-1, 3); d = np.ascontiguousarray(d, np.float64).reshape(-1, 3)
        out = np.zeros((o.shape[0], nsteps, 8), np.float64)
        self._ck(lib().rrt_geodesic(self.h, o.shape[0], _p(o), _p(d), int(nsteps), _p(out)))
        return out

    def generate_rays(self, xy):
        xy = np.ascontiguousarray(xy, np.float64).reshape(-1, 2)
        o = np.zeros((xy.shape[0], 3)); d = np.zeros((xy.shape[0], 3))
        self._ck(lib().rrt_generate_rays(self.h, xy.shape[0], _p(xy), _p(o), _p(d)))
        return o, d

    def generate_rays_thin_lens(self, xy, lens):
        """Camera::generate_ray_for_thin_lens; lens [n, 2] = (rndR, rndTheta)"""
        xy = np.ascontiguousarray(xy, np.float64).reshape(-1, 2); lens = np.ascontiguousarray(lens, np.float64).reshape(-1, 2)
        o = np.zeros((xy.shape[0], 3)); d = np.zeros((xy.shape[0], 3))
        self._ck(lib().rrt_generate_rays_thin_lens(self.h, xy.shape[0], _p(xy), _p(lens), _p(o), _p(d)))
        return o, d

    def bsdf_sample(self, mat, wo, u):
        wo = np.ascontiguousarray(wo, np.float64).reshape(-1, 3); u = np.ascontiguousarray(u, np.float64).reshape(-1, 4)
        n = wo.shape[0]
        f = np.zeros((n, 3), np.float32); wi = np.zeros((n, 3)); pdf = np.zeros(n, np.float32)
        self._ck(lib().rrt_bsdf_sample(self.h, int(mat), n, _p(wo), _p(u), _p(f), _p(wi), _p(pdf)))
        return f, wi, pdf

    def bsdf_eval(self, mat, wo, wi):
        wo = np.ascontiguousarray(wo, np.float64).reshape(-1, 3); wi = np.ascontiguousarray(wi, np.float64).reshape(-1, 3)
        f = np.zeros((wo.shape[0], 3), np.float32)
        self._ck(lib().rrt_bsdf_eval(self.h, int(mat), wo.shape[0], _p(wo), _p(wi), _p(f)))
        return f

    def light_sample(self, light, p, u):
        p = np.ascontiguousarray(p, np.float64).reshape(-1, 3); u = np.ascontiguousarray(u, np.float64).reshape(-1, 2)
        n = p.shape[0]
        rad = np.zeros((n, 3), np.float32); wi = np.zeros((n, 3)); dist = np.zeros(n, np.float32); pdf = np.zeros(n, np.float32)
        self._ck(lib().rrt_light_sample(self.h, int(light), n, _p(p), _p(u), _p(rad), _p(wi), _p(dist), _p(pdf)))
        return rad, wi, dist, pdf

    def set_envmap(self, rgb):
        """environment light (-e): rgb [h, w, 3] float32, row 0 = top; None removes it"""
        if rgb is None:
            self._ck(lib().rrt_set_envmap(self.h, 0, 0, None))
            return
        a = np.ascontiguousarray(rgb, np.float32)
        if a.ndim != 3 or a.shape[2] != 3:
            raise ValueError("set_envmap: expected an [h, w, 3] array")
        self._ck(lib().rrt_set_envmap(self.h, a.shape[1], a.shape[0], _p(a)))

    def env_lookup(self, d):
        d = np.ascontiguousarray(d, np.float64).reshape(-1, 3)
        out = np.zeros((d.shape[0], 3), np.float32)
        self._ck(lib().rrt_env_lookup(self.h, d.shape[0], _p(d), _p(out)))
        return out

    def philox(self, seed, counters):
        counters = np.ascontiguousarray(counters, np.uint32).reshape(-1, 4)
        out = np.zeros_like(counters)
        self._ck(lib().rrt_philox(self.h, int(seed), counters.shape[0], _p(counters), _p(out)))
        return out

    def tonemap(self, hdr):
        hdr = np.ascontiguousarray(hdr, np.float32).reshape(-1, 3)
        out = np.zeros(hdr.shape[0], np.uint32)
        self._ck(lib().rrt_tonemap(self.h, hdr.shape[0], _p(hdr), _p(out)))
        return out


# --------------------------------------------------------------------------------------------------------------
# Mirror of the reference classes.  Same names, argument meaning and error behaviour (silent early return in the
# wrong state, reference src/pathtracer.cpp:97-99,121-123,205,218-220,225).

class BVHAccel:
    """reference src/bvh.h:50-144.  Built from a flat scene (the primitive list in reference order)."""

    def __init__(self, scene, max_leaf_size=4, ctx=None):
        if max_leaf_size != 4:
            raise ValueError("the device BVH uses the reference's default max_leaf_size = 4")
        self.ctx = ctx or Context()
        self.ctx.upload_scene(scene)
        self.scene = scene
        self.total_rays = 0
        self.total_isects = 0

    def get_bbox(self):
        t = self.scene.tri_pos.reshape(-1, 3)
        lo = [t.min(0)] if len(t) else []
        hi = [t.max(0)] if len(t) else []
        if len(self.scene.sph):
            lo.append((self.scene.sph[:, :3] - self.scene.sph[:, 3:4]).min(0)); hi.append((self.scene.sph[:, :3] + self.scene.sph[:, 3:4]).max(0))
        return np.min(lo, 0), np.max(hi, 0)

    def intersect(self, o, d, min_t=None, max_t=None, want_isect=True):
        """Batched BVHAccel::intersect(ray[, isect]).  Returns the rrt_hit records."""
        hits, st = self.ctx.trace_rays(o, d, min_t, max_t, any_hit=not want_isect)
        self.total_rays += st["rays"]
        self.total_isects += st["tri_tests"] + st["sphere_tests"]
        return hits


class PathTracer:
    """reference src/pathtracer.h:64-167.  States INIT -> READY -> RENDERING -> DONE (:229-235)."""
    INIT, READY, VISUALIZE, RENDERING, DONE = range(5)

    def __init__(self, ns_aa=1, max_ray_depth=1, ns_area_light=1, ns_diff=1, ns_glsy=1, ns_refr=1, num_threads=1,
                 samples_per_batch=32, max_tolerance=0.05, envmap=None, direct_hemisphere_sample=False, filename="",
                 lensRadius=0.25, focalDistance=4.7, device=0, seed=0):
        self.envmap = envmap      # [h, w, 3] float array or None (reference: HDRImageBuffer*, src/pathtracer.cpp:61-65)
        self.ns_aa, self.max_ray_depth, self.ns_area_light = ns_aa, max_ray_depth, ns_area_light
        self.ns_diff, self.ns_glsy, self.ns_refr = ns_diff, ns_diff, ns_refr  # reference bug kept: ns_glsy = ns_diff (src/pathtracer.cpp:51)
        self.numWorkerThreads = num_threads  # accepted and ignored: the device replaces the worker pool
        self.samplesPerBatch, self.maxTolerance = samples_per_batch, max_tolerance
        self.direct_hemisphere_sample = direct_hemisphere_sample
        self.filename, self.lensRadius, self.focalDistance = filename, lensRadius, focalDistance
        self.seed = seed
        self.state = PathTracer.INIT
        self.ctx = Context(device)
        if self.envmap is not None:
            self.ctx.set_envmap(self.envmap)
        self.bvh = None
        self.scene = None
        self.camera_set = False
        self.sampleBuffer = None       # HDR float [H, W, 3], row 0 = image bottom (reference src/pathtracer.cpp:575)
        self.sampleCountBuffer = None  # int [H, W]
        self.stats = None
        self._thread = None
        self._cv = threading.Condition()
        self.render_cell = False
        self.cell_tl = (0, 0); self.cell_br = (0, 0)

    # reference src/pathtracer.cpp:95-117
    def set_scene(self, scene):
        if self.state != PathTracer.INIT:
            return
        self.scene = scene
        self.bvh = BVHAccel(scene, ctx=self.ctx)
        if self.camera_set:
            self.state = PathTracer.READY

    # reference src/pathtracer.cpp:119-134
    def set_camera(self, camera):
        if self.state != PathTracer.INIT:
            return
        cam = camera.c_camera() if hasattr(camera, "c_camera") else camera
        cam.lensRadius, cam.focalDistance = self.lensRadius, self.focalDistance
        self.ctx.set_camera(cam)
        self.camera_set = True
        if self.bvh is not None:
            self.state = PathTracer.READY

    # reference src/pathtracer.cpp:136-149
    def set_frame_size(self, width, height):
        if self.state not in (PathTracer.INIT, PathTracer.READY):
            self.stop()
        self.ctx.set_frame(width, height)
        self.sampleBuffer = np.zeros((height, width, 3), np.float32)
        self.sampleCountBuffer = np.zeros((height, width), np.int32)

    def set_blackhole(self, centre=(0.0, 1.0, 0.0), r=0.1, dtheta=0.1, mode=_sc.MODE_LITERAL):
        """-B x y z r dtheta (reference src/main.cpp:139-145) + the flat-space switch of DESIGN.md."""
        self.ctx.set_blackhole(centre, r, dtheta, mode)

    def _params(self):
        return _sc.make_params(self.ns_aa, self.max_ray_depth, self.ns_area_light, self.samplesPerBatch, self.maxTolerance,
                               self.direct_hemisphere_sample, False, self.seed)

    def _work(self):
        if self.render_cell:
            x0, y0 = self.cell_tl; x1, y1 = self.cell_br
            self.ctx.set_cell(x0, y0, x1 - x0, y1 - y0)
        else:
            self.ctx.set_cell(0, 0, 0, 0)
        _, _, self.stats = self.ctx.render(self._params(), self.sampleBuffer, self.sampleCountBuffer)
        with self._cv:
            self.state = PathTracer.DONE
            self._cv.notify_all()

    # reference src/pathtracer.cpp:224-282: asynchronous, returns at once
    def start_raytracing(self):
        if self.state != PathTracer.READY:
            return
        self.state = PathTracer.RENDERING
        self._thread = threading.Thread(target=self._work)
        self._thread.start()

    # reference src/pathtracer.cpp:180-202
    def stop(self):
        if self._thread is not None:
            self._thread.join()
            self._thread = None
        if self.state in (PathTracer.RENDERING, PathTracer.DONE, PathTracer.VISUALIZE):
            self.state = PathTracer.READY

    def clear(self):
        if self.state != PathTracer.READY:
            return
        self.bvh = None; self.scene = None; self.camera_set = False
        self.state = PathTracer.INIT

    # reference src/pathtracer.cpp:284-301
    def render_to_file(self, filename, x=-1, y=-1, dx=0, dy=0):
        if x == -1:
            self.render_cell = False
        else:
            self.render_cell = True
            self.cell_tl = (x, y); self.cell_br = (x + dx, y + dy)
        self.start_raytracing()
        with self._cv:
            while self.state != PathTracer.DONE:
                self._cv.wait()
        self._thread.join(); self._thread = None
        self.save_image(filename)

    # reference src/pathtracer.cpp:646-717: tone map on the device, vertical flip, PNG + the _rate.png heat map
    def save_image(self, filename=None):
        if self.state != PathTracer.DONE:
            return
        filename = filename or self.filename
        h, w = self.sampleCountBuffer.shape
        rgba = self.ctx.tonemap(self.sampleBuffer).reshape(h, w)
        write_png(filename, np.ascontiguousarray(rgba[::-1]).view(np.uint8).reshape(h, w, 4))
        self.save_sampling_rate_image(filename)

    def save_sampling_rate_image(self, filename):
        """reference src/pathtracer.cpp:686-717: rate = samples taken / ns_aa; blue (few) -> green (half) -> red (all)."""
        h, w = self.sampleCountBuffer.shape
        rate = (self.sampleCountBuffer.astype(np.float32) * np.float32(1.0)) / np.float32(self.ns_aa)
        lo = rate <= 0.5
        r_lo = ((0.5 - rate.astype(np.float64)) / 0.5).astype(np.float32)
        r_hi = ((1.0 - rate.astype(np.float64)) / 0.5).astype(np.float32)
        red = np.where(lo, np.float32(0), (1.0 - r_hi.astype(np.float64)).astype(np.float32))
        green = np.where(lo, (1.0 - r_lo.astype(np.float64)).astype(np.float32), r_hi)
        blue = np.where(lo, r_lo, np.float32(0))
        chan = lambda c: (np.minimum(np.float32(1), c) * np.float32(255)).astype(np.uint32).astype(np.uint8)   # ImageBuffer::update_pixel
        out = np.zeros((h, w, 4), np.uint8)
        out[..., 0], out[..., 1], out[..., 2], out[..., 3] = chan(red), chan(green), chan(blue), 255
        base, ext = os.path.splitext(filename)
        write_png(base + "_rate" + (ext or ".png"), np.ascontiguousarray(out[::-1]))


def write_png(path, rgba):
    """Minimal RGBA8 PNG writer (the reference uses the vendored lodepng, src/pathtracer.cpp:676)."""
    h, w, _ = rgba.shape
    raw = b"".join(b"\x00" + rgba[y].tobytes() for y in range(h))

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 6, 0, 0, 0)) +
                chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))
