"""Flat scene container + ctypes mirrors of include/rrt_b200.h.

Host-side plumbing only (no compute).  A scene is the reference's StaticScene flattened in the order
PathTracer::build_accel collects primitives (reference src/pathtracer.cpp:311-316); the on-disk form is the
"RRTSCN01" file written by oracle/ref_probe.cpp (from the reference's COLLADA loader) or by the procedural
generator in synth.py.  Layout: DESIGN.md, "Data formats".
"""
import ctypes as C
import math
import struct

import numpy as np

MAT_DIFFUSE, MAT_MIRROR, MAT_GLASS, MAT_MICROFACET, MAT_EMISSION, MAT_REFRACTION = range(6)
LIGHT_DIRECTIONAL, LIGHT_HEMISPHERE, LIGHT_POINT, LIGHT_AREA, LIGHT_NULL = range(5)
MODE_FLAT, MODE_LITERAL = 0, 1


class Material(C.Structure):
    _fields_ = [("type", C.c_int32), ("a", C.c_float * 3), ("b", C.c_float * 3), ("alpha", C.c_float), ("ior", C.c_float)]


class Light(C.Structure):
    _fields_ = [("type", C.c_int32), ("radiance", C.c_float * 3), ("position", C.c_double * 3), ("direction", C.c_double * 3),
                ("dim_x", C.c_double * 3), ("dim_y", C.c_double * 3), ("area", C.c_float), ("is_delta", C.c_int32)]


class Scene(C.Structure):
    _fields_ = [("n_prims", C.c_uint32), ("prim_kind", C.c_void_p), ("prim_index", C.c_void_p),
                ("n_tris", C.c_uint32), ("tri_pos", C.c_void_p), ("tri_nrm", C.c_void_p), ("tri_mat", C.c_void_p),
                ("n_spheres", C.c_uint32), ("sph", C.c_void_p), ("sph_mat", C.c_void_p),
                ("n_mats", C.c_uint32), ("mats", C.c_void_p), ("n_lights", C.c_uint32), ("lights", C.c_void_p)]


class Camera(C.Structure):
    _fields_ = [("pos", C.c_double * 3), ("c2w", C.c_double * 9), ("hFov", C.c_double), ("vFov", C.c_double),
                ("nClip", C.c_double), ("fClip", C.c_double), ("lensRadius", C.c_double), ("focalDistance", C.c_double)]


class Params(C.Structure):
    _fields_ = [("ns_aa", C.c_int32), ("max_ray_depth", C.c_int32), ("ns_area_light", C.c_int32),
                ("samples_per_batch", C.c_int32), ("max_tolerance", C.c_float), ("direct_hemisphere_sample", C.c_int32),
                ("thin_lens", C.c_int32), ("seed", C.c_uint64), ("max_paths_in_flight", C.c_int32), ("reserved", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("seconds", C.c_double), ("trace_seconds", C.c_double), ("paths", C.c_uint64), ("rays", C.c_uint64),
                ("geodesic_steps", C.c_uint64), ("node_visits", C.c_uint64), ("tri_tests", C.c_uint64),
                ("sphere_tests", C.c_uint64), ("kernel_launches", C.c_uint64), ("trace_launches", C.c_uint64),
                ("algorithmic_bytes", C.c_uint64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class Hit(C.Structure):
    _fields_ = [("status", C.c_int32), ("prim", C.c_int32), ("chord", C.c_int32), ("steps", C.c_int32), ("t", C.c_double),
                ("hit_p", C.c_double * 3), ("w_out", C.c_double * 3), ("n", C.c_double * 3)]


HIT_DTYPE = np.dtype([("status", "<i4"), ("prim", "<i4"), ("chord", "<i4"), ("steps", "<i4"), ("t", "<f8"),
                      ("hit_p", "<f8", 3), ("w_out", "<f8", 3), ("n", "<f8", 3)])
MAT_DTYPE = np.dtype([("type", "<i4"), ("a", "<f4", 3), ("b", "<f4", 3), ("alpha", "<f4"), ("ior", "<f4")])
LIGHT_DTYPE = np.dtype([("type", "<i4"), ("radiance", "<f4", 3), ("position", "<f8", 3), ("direction", "<f8", 3),
                        ("dim_x", "<f8", 3), ("dim_y", "<f8", 3), ("area", "<f4"), ("is_delta", "<i4")])
assert C.sizeof(Material) == MAT_DTYPE.itemsize == 36
assert C.sizeof(Light) == LIGHT_DTYPE.itemsize == 120
assert C.sizeof(Hit) == HIT_DTYPE.itemsize == 96


def make_params(ns_aa=1, max_ray_depth=1, ns_area_light=1, samples_per_batch=32, max_tolerance=0.05,
                direct_hemisphere_sample=False, thin_lens=False, seed=0, max_paths_in_flight=0):
    """Defaults are the reference's AppConfig (src/application.h:43-64)."""
    return Params(int(ns_aa), int(max_ray_depth), int(ns_area_light), int(samples_per_batch), float(max_tolerance),
                  int(bool(direct_hemisphere_sample)), int(bool(thin_lens)), int(seed), int(max_paths_in_flight), 0)


class SceneData:
    """Numpy-backed scene + camera.  Keeps the arrays alive for the ctypes view."""

    def __init__(self, prim_kind, prim_index, tri_pos, tri_nrm, tri_mat, sph, sph_mat, mats, lights, camera):
        self.prim_kind = np.ascontiguousarray(prim_kind, np.uint8)
        self.prim_index = np.ascontiguousarray(prim_index, np.uint32)
        self.tri_pos = np.ascontiguousarray(tri_pos, np.float64).reshape(-1, 9)
        self.tri_nrm = np.ascontiguousarray(tri_nrm, np.float64).reshape(-1, 9)
        self.tri_mat = np.ascontiguousarray(tri_mat, np.int32)
        self.sph = np.ascontiguousarray(sph, np.float64).reshape(-1, 4)
        self.sph_mat = np.ascontiguousarray(sph_mat, np.int32)
        self.mats = np.ascontiguousarray(mats, MAT_DTYPE)
        self.lights = np.ascontiguousarray(lights, LIGHT_DTYPE)
        self.camera = dict(camera)  # pos[3], c2w[9], hFov, vFov, nClip, fClip, lensRadius, focalDistance, screenW, screenH, screenDist

    @property
    def n_prims(self):
        return int(self.prim_kind.shape[0])

    def c_scene(self):
        p = lambda a: a.ctypes.data_as(C.c_void_p) if a.size else None
        return Scene(self.n_prims, p(self.prim_kind), p(self.prim_index), self.tri_pos.shape[0], p(self.tri_pos), p(self.tri_nrm),
                     p(self.tri_mat), self.sph.shape[0], p(self.sph), p(self.sph_mat), self.mats.shape[0], p(self.mats),
                     self.lights.shape[0], p(self.lights))

    def c_camera(self):
        c = self.camera
        return Camera((C.c_double * 3)(*c["pos"]), (C.c_double * 9)(*c["c2w"]), c["hFov"], c["vFov"], c["nClip"], c["fClip"],
                      c["lensRadius"], c["focalDistance"])

    def set_screen_size(self, w, h):
        """Camera::set_screen_size (reference src/camera.cpp:68-74): screenDist stays, the FOV follows the frame."""
        c = self.camera
        c["screenW"], c["screenH"] = float(w), float(h)
        deg = lambda rad: rad * (180 / 3.14159265358979323)
        c["hFov"] = 2 * deg(math.atan(float(w) / (2 * c["screenDist"])))
        c["vFov"] = 2 * deg(math.atan(float(h) / (2 * c["screenDist"])))
        return self

    # ---- "RRTSCN01" ----
    def save(self, path):
        c = self.camera
        camv = list(c["pos"]) + list(c["c2w"]) + [c["hFov"], c["vFov"], c["nClip"], c["fClip"], c["lensRadius"], c["focalDistance"],
                                                   c["screenW"], c["screenH"], c["screenDist"], 0.0, 0.0, 0.0]
        with open(path, "wb") as f:
            f.write(b"RRTSCN01")
            f.write(struct.pack("<8I", self.n_prims, self.tri_pos.shape[0], self.sph.shape[0], self.mats.shape[0],
                                self.lights.shape[0], 0, 0, 0))
            f.write(np.asarray(camv, np.float64).tobytes())
            for a in (self.prim_kind, self.prim_index, self.tri_pos, self.tri_nrm, self.tri_mat, self.sph, self.sph_mat, self.mats,
                      self.lights):
                b = a.tobytes()
                f.write(b + b"\0" * ((8 - len(b) % 8) % 8))

    @staticmethod
    def load(path):
        with open(path, "rb") as f:
            buf = f.read()
        if buf[:8] != b"RRTSCN01":
            raise ValueError("%s: not an RRTSCN01 scene file" % path)
        n_prims, n_tris, n_sph, n_mats, n_lights = struct.unpack_from("<5I", buf, 8)
        off = 8 + 32
        camv = np.frombuffer(buf, np.float64, 24, off)
        off += 24 * 8

        def take(dtype, count):
            nonlocal off
            a = np.frombuffer(buf, dtype, count, off).copy()
            nbytes = a.nbytes
            off += nbytes + (8 - nbytes % 8) % 8
            return a

        kind = take(np.uint8, n_prims)
        index = take(np.uint32, n_prims)
        tpos = take(np.float64, n_tris * 9)
        tnrm = take(np.float64, n_tris * 9)
        tmat = take(np.int32, n_tris)
        sph = take(np.float64, n_sph * 4)
        smat = take(np.int32, n_sph)
        mats = take(MAT_DTYPE, n_mats)
        lights = take(LIGHT_DTYPE, n_lights)
        cam = dict(pos=camv[0:3].tolist(), c2w=camv[3:12].tolist(), hFov=float(camv[12]), vFov=float(camv[13]), nClip=float(camv[14]),
                   fClip=float(camv[15]), lensRadius=float(camv[16]), focalDistance=float(camv[17]), screenW=float(camv[18]),
                   screenH=float(camv[19]), screenDist=float(camv[20]))
        return SceneData(kind, index, tpos, tnrm, tmat, sph, smat, mats, lights, cam)
