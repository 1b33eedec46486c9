"""ctypes binding of oracle/liboracle.so — TEST INFRASTRUCTURE ONLY.

Loaded by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs; the
product package (ray-tracer_b200/) never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class OrcPrimInfo(C.Structure):
    _fields_ = [("kind", C.c_int32), ("material_kind", C.c_int32), ("world_from_local", C.c_double * 16),
                ("param", C.c_double * 4)]


class OrcCameraArgs(C.Structure):
    _fields_ = [("eye", C.c_double * 3), ("center", C.c_double * 3), ("up", C.c_double * 3), ("fov", C.c_double),
                ("aspect", C.c_double), ("focusDistance", C.c_double), ("lensRadius", C.c_double)]


class OrcRenderParams(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("spp", C.c_uint32), ("max_depth", C.c_uint32),
                ("seed", C.c_uint64), ("row_begin", C.c_uint32), ("row_stride", C.c_uint32),
                ("sampler_mode", C.c_int32), ("threads", C.c_int32)]


class OrcStats(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("segments", C.c_uint64), ("box_tests", C.c_uint64), ("prim_tests", C.c_uint64),
                ("medium_draws", C.c_uint64), ("seconds", C.c_double), ("threads", C.c_int32)]


class OrcHit(C.Structure):
    _fields_ = [("hit", C.c_int32), ("prim", C.c_int32), ("stable", C.c_int32), ("material_kind", C.c_int32),
                ("t", C.c_double), ("p", C.c_double * 3), ("n", C.c_double * 3), ("u", C.c_double), ("v", C.c_double),
                ("t_second", C.c_double), ("t_sens", C.c_double), ("p_sens", C.c_double), ("n_sens", C.c_double)]


HIT_DTYPE = np.dtype([("hit", "<i4"), ("prim", "<i4"), ("stable", "<i4"), ("material_kind", "<i4"), ("t", "<f8"),
                      ("p", "<f8", 3), ("n", "<f8", 3), ("u", "<f8"), ("v", "<f8"), ("t_second", "<f8"), ("t_sens", "<f8"),
                      ("p_sens", "<f8"), ("n_sens", "<f8")])
assert HIT_DTYPE.itemsize == C.sizeof(OrcHit)
PRIM_DTYPE = np.dtype([("kind", "<i4"), ("material_kind", "<i4"), ("world_from_local", "<f8", 16), ("param", "<f8", 4)])
FAITHFUL, PAIRED = 0, 1


def build(force=False):
    """Compile liboracle.so with oracle/Makefile (g++)."""
    so = os.path.join(_HERE, "liboracle.so")
    if force and os.path.exists(so):
        os.remove(so)
    subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    so = os.path.join(_HERE, "liboracle.so")
    if not os.path.exists(so):
        build()
    L = C.CDLL(so)
    L.orc_scene_create.restype = C.c_void_p
    L.orc_scene_create.argtypes = [C.c_char_p, C.c_uint64]
    L.orc_scene_destroy.argtypes = [C.c_void_p]
    L.orc_scene_camera.argtypes = [C.c_void_p, C.c_double, C.POINTER(OrcCameraArgs)]
    L.orc_camera_frame.argtypes = [C.POINTER(OrcCameraArgs), C.c_void_p]
    L.orc_scene_prim_count.restype = C.c_uint32
    L.orc_scene_prim_count.argtypes = [C.c_void_p]
    L.orc_scene_prims.argtypes = [C.c_void_p, C.c_void_p]
    L.orc_render.argtypes = [C.c_void_p, C.POINTER(OrcCameraArgs), C.POINTER(OrcRenderParams), C.c_void_p, C.c_void_p,
                             C.POINTER(OrcStats)]
    L.orc_paths.argtypes = [C.c_void_p, C.POINTER(OrcCameraArgs), C.POINTER(OrcRenderParams), C.c_void_p, C.c_void_p,
                            C.c_uint64, C.c_void_p, C.c_void_p]
    L.orc_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_double, C.c_double, C.c_void_p]
    L.orc_resolve_u8.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
    L.orc_philox.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.orc_mat4_translation.argtypes = [C.c_void_p, C.c_void_p]
    L.orc_mat4_rotation.argtypes = [C.c_double, C.c_void_p, C.c_void_p]
    L.orc_mat4_multiplied.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.orc_mat4_inversed.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_double)]
    L.orc_vec4_transformed.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.orc_cube_faces.argtypes = [C.c_double, C.c_double, C.c_double, C.c_void_p, C.c_void_p]
    L.orc_prim_hit.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.POINTER(OrcHit)]
    L.orc_medium_hit.argtypes = [C.c_double, C.c_double, C.c_void_p, C.c_int, C.c_uint64, C.c_uint32, C.POINTER(OrcHit)]
    L.orc_scatter.argtypes = [C.c_int, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_uint64,
                              C.c_uint32, C.c_void_p, C.c_void_p]
    L.orc_camera_ray.argtypes = [C.POINTER(OrcCameraArgs), C.c_double, C.c_double, C.c_int, C.c_uint64, C.c_uint32,
                                 C.c_uint32, C.c_void_p]
    _LIB = L
    return L


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _f8(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64))


class Scene:
    """An oracle scene graph built by the shared example builders (ray-tracer_b200/examples/scenes.hpp)."""

    def __init__(self, name, seed=1):
        self.name, self.seed = name, seed
        self._h = lib().orc_scene_create(name.encode(), seed)
        if not self._h:
            raise ValueError(f"unknown scene {name!r}")

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_scene_destroy(self._h)
            self._h = None

    def camera(self, aspect):
        cam = OrcCameraArgs()
        lib().orc_scene_camera(self._h, float(aspect), C.byref(cam))
        return cam

    def prims(self):
        n = lib().orc_scene_prim_count(self._h)
        out = np.zeros(n, dtype=PRIM_DTYPE)
        lib().orc_scene_prims(self._h, _p(out))
        return out

    def _params(self, width, height, spp, max_depth, seed, row_begin, row_stride, mode, threads):
        return OrcRenderParams(width, height, spp, max_depth, seed, row_begin, row_stride, mode, threads)

    def render(self, width, height, spp, max_depth, seed=1, row_begin=0, row_stride=1, mode=PAIRED, threads=0, cam=None):
        """Returns (sum[rows,W,3], sumsq[rows,W,3], stats) — f64 radiance sums over spp; row k is y = row_begin + k*row_stride."""
        cam = cam or self.camera(width / height)
        prm = self._params(width, height, spp, max_depth, seed, row_begin, row_stride, mode, threads)
        rows = len(range(row_begin, height, row_stride))
        s = np.zeros((rows, width, 3), dtype=np.float64)
        q = np.zeros((rows, width, 3), dtype=np.float64)
        st = OrcStats()
        rc = lib().orc_render(self._h, C.byref(cam), C.byref(prm), _p(s), _p(q), C.byref(st))
        assert rc == 0
        return s, q, st

    def paths(self, width, height, max_depth, xy, sample, seed=1, mode=PAIRED, cam=None):
        cam = cam or self.camera(width / height)
        prm = self._params(width, height, 1, max_depth, seed, 0, 1, mode, 1)
        xy = np.ascontiguousarray(np.asarray(xy, dtype=np.uint32).reshape(-1, 2))
        sample = np.ascontiguousarray(np.asarray(sample, dtype=np.uint32))
        n = len(sample)
        rgb = np.zeros((n, 3), dtype=np.float64)
        seg = np.zeros(n, dtype=np.uint32)
        rc = lib().orc_paths(self._h, C.byref(cam), C.byref(prm), _p(xy), _p(sample), n, _p(rgb), _p(seg))
        assert rc == 0
        return rgb, seg

    def color_rays(self, rays, spp, max_depth, seed=1, mode=PAIRED):
        """render::color on caller rays [n,6]: (sum[n,3] over spp evaluations keyed (ray, sample), segments[n])."""
        rays = _f8(rays).reshape(-1, 6)
        out = np.zeros((len(rays), 3), dtype=np.float64)
        seg = np.zeros(len(rays), dtype=np.uint64)
        L = lib()
        L.orc_color_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p]
        rc = L.orc_color_rays(self._h, _p(rays), len(rays), int(spp), int(max_depth), int(seed), int(mode), _p(out), _p(seg))
        assert rc == 0
        return out, seg

    def trace(self, rays, seed=1, rel_perturb=1e-6, tie_rel=1e-4):
        """Closest hit per ray plus the margin diagnostics (stable, t_second, *_sens) of orc_trace."""
        rays = _f8(rays).reshape(-1, 6)
        out = np.zeros(len(rays), dtype=HIT_DTYPE)
        rc = lib().orc_trace(self._h, _p(rays), len(rays), seed, float(rel_perturb), float(tie_rel), _p(out))
        assert rc == 0
        return out


def set_earth_image(rgb_top_down):
    """orc_set_earth_image: the image book-two maps onto its earth sphere (examples/main.rs:266-281); None = synthetic stand-in."""
    L = lib()
    L.orc_set_earth_image.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
    L.orc_set_earth_image.restype = None
    if rgb_top_down is None:
        L.orc_set_earth_image(None, 0, 0)
        return
    img = np.ascontiguousarray(rgb_top_down, dtype=np.uint8)
    L.orc_set_earth_image(img.ctypes.data_as(C.c_void_p), img.shape[1], img.shape[0])


def camera_args(eye, center, up, fov, aspect, focus, lens):
    cam = OrcCameraArgs()
    cam.eye[:], cam.center[:], cam.up[:] = eye, center, up
    cam.fov, cam.aspect, cam.focusDistance, cam.lensRadius = fov, aspect, focus, lens
    return cam


def camera_frame(cam):
    out = np.zeros(9)
    lib().orc_camera_frame(C.byref(cam), _p(out))
    return out.reshape(3, 3)  # lowerLeft, horizontal, vertical


def camera_ray(cam, u, v, mode=PAIRED, seed=1, key0=0, key1=0):
    out = np.zeros(6)
    lib().orc_camera_ray(C.byref(cam), u, v, mode, seed, key0, key1, _p(out))
    return out


def camera_rays(cam, width, height, jitter=0.5):
    """Pixel-centre camera rays (lens draw of sample 0), [H*W, 6]."""
    out = np.zeros((height * width, 6))
    for y in range(height):
        for x in range(width):
            out[y * width + x] = camera_ray(cam, (x + jitter) / width, (y + jitter) / height, PAIRED, 1, y * width + x, 0)
    return out


def resolve_u8(sum_rgb, spp):
    h, w, _ = sum_rgb.shape
    s = _f8(sum_rgb)
    out = np.zeros((h, w, 3), dtype=np.uint8)
    lib().orc_resolve_u8(_p(s), w, h, spp, _p(out))
    return out


def philox(ctr, key):
    c = np.asarray(ctr, dtype=np.uint32)
    k = np.asarray(key, dtype=np.uint32)
    o = np.zeros(4, dtype=np.uint32)
    lib().orc_philox(_p(c), _p(k), _p(o))
    return o


def mat4_translation(v):
    o = np.zeros(16)
    lib().orc_mat4_translation(_p(_f8(v)), _p(o))
    return o


def mat4_rotation(radians, axis):
    o = np.zeros(16)
    lib().orc_mat4_rotation(float(radians), _p(_f8(axis)), _p(o))
    return o


def mat4_multiplied(a, b):
    o = np.zeros(16)
    lib().orc_mat4_multiplied(_p(_f8(a)), _p(_f8(b)), _p(o))
    return o


def mat4_inversed(a):
    o = np.zeros(16)
    det = C.c_double()
    ok = lib().orc_mat4_inversed(_p(_f8(a)), _p(o), C.byref(det))
    return (o if ok else None), det.value


def vec4_transformed(v, m):
    o = np.zeros(4)
    lib().orc_vec4_transformed(_p(_f8(v)), _p(_f8(m)), _p(o))
    return o


def cube_faces(w, h, d):
    n = np.zeros(18)
    c = np.zeros(18)
    lib().orc_cube_faces(w, h, d, _p(n), _p(c))
    return n.reshape(6, 3), c.reshape(6, 3)


def prim_hit(kind, param, ray):
    h = OrcHit()
    lib().orc_prim_hit(kind, _p(_f8(param)), _p(_f8(ray)), C.byref(h))
    return h


def medium_hit(r, density, ray, mode=PAIRED, seed=1, key=0):
    h = OrcHit()
    lib().orc_medium_hit(float(r), float(density), _p(_f8(ray)), mode, seed, key, C.byref(h))
    return h


def scatter(kind, param, albedo, ray, p, n, mode=PAIRED, seed=1, key=0):
    out = np.zeros(6)
    att = np.zeros(3)
    rc = lib().orc_scatter(kind, float(param), _p(_f8(albedo)), _p(_f8(ray)), _p(_f8(p)), _p(_f8(n)), mode, seed, key,
                           _p(out), _p(att))
    return rc, out, att
