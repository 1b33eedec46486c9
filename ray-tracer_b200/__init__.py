"""ray-tracer_b200 — ctypes binding of librt_b200.so (include/rt_b200.h, include/rt_b200_host.h).

This is harness plumbing over the C ABI (the product is the shared library): it loads the in-tree
`librt_b200.so`, mirrors the POD structs, and offers `Device`, `FlatScene`, `Scene` wrappers used by
tests/, bench.py and __graft_entry__.py.  There is no CPU fallback: if the library is missing the
import raises, and `Device()` raises without a CUDA device.

Import with `importlib.import_module("ray-tracer_b200")` (the directory name is not an identifier).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RT_B200_LIB") or os.path.join(_HERE, "librt_b200.so")  # override: tuning variants only

RENDER_STAGE_TIMING, RENDER_COUNT_WORK, RENDER_SAMPLE_MAJOR, RENDER_EXTEND_PER_THREAD, RENDER_EXTEND_PERSISTENT, RENDER_TIME_EVERY_ITERATION = 1, 2, 4, 8, 16, 32
CTX_NO_ACCUMULATE_OVERLAP, CTX_STAGE_TIMING, CTX_SHADE_ONE_RAY, CTX_NO_FUSED_RECTS = 1, 2, 4, 8
BVH_AUTO, BVH_HOST_SAH, BVH_DEVICE_LBVH = 0, 1, 2
SCENE_NO_FRAME_RECTS, SCENE_NO_STAGED_PRIMS, SCENE_NO_SMALL_LIST, SCENE_NO_FRAME_BOXES, SCENE_NO_BOX_PRIMS, SCENE_NO_WHOLE_TREE = 1, 2, 4, 8, 16, 32
RT_OK, RT_ERR_INVALID, RT_ERR_UNSUPPORTED, RT_ERR_CUDA, RT_ERR_NO_DEVICE, RT_ERR_OOM, RT_ERR_NCCL = 0, -1, -2, -3, -4, -5, -6
PRIM_SPHERE, PRIM_RECT, PRIM_MEDIUM_SPHERE, PRIM_MEDIUM = 0, 1, 2, 3
NO_PRIM = 0xFFFFFFFF
NOALLOC_MAX_PIXELS, NOALLOC_QUEUE_RAYS, NOALLOC_SCRATCH_BYTES = 65536, 32768, 9 << 20
MAT_LAMBERTIAN, MAT_METAL, MAT_DIELECTRIC, MAT_DIFFUSE_LIGHT, MAT_ISOTROPIC = 0, 1, 2, 3, 4
TEX_SOLID, TEX_CHECKER, TEX_IMAGE = 0, 1, 2

PRIM_DTYPE = np.dtype([("kind", "<u4"), ("material", "<i4"), ("order", "<u4"), ("aux", "<u4"), ("param", "<f4", 4),
                       ("local_from_world", "<f4", 12), ("world_from_local", "<f4", 12)])
MATERIAL_DTYPE = np.dtype([("kind", "<u4"), ("texture", "<i4"), ("fuzz", "<f4"), ("ior", "<f4")])
TEXTURE_DTYPE = np.dtype([("kind", "<u4"), ("a", "<i4"), ("b", "<i4"), ("width", "<u4"), ("height", "<u4"), ("reserved", "<u4"),
                          ("texel_offset", "<u8"), ("rgb", "<f4", 4)])
RAY_DTYPE = np.dtype([("origin", "<f4", 3), ("direction", "<f4", 3)])
HIT_DTYPE = np.dtype([("hit", "<i4"), ("prim", "<i4"), ("t", "<f4"), ("p", "<f4", 3), ("n", "<f4", 3), ("uv", "<f4", 2), ("material", "<i4")])
assert PRIM_DTYPE.itemsize == 128 and MATERIAL_DTYPE.itemsize == 16 and TEXTURE_DTYPE.itemsize == 48
assert RAY_DTYPE.itemsize == 24 and HIT_DTYPE.itemsize == 48


class RtSceneDesc(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("n_prims", C.c_uint32), ("n_materials", C.c_uint32), ("n_textures", C.c_uint32),
                ("prims", C.c_void_p), ("materials", C.c_void_p), ("textures", C.c_void_p), ("texels", C.c_void_p), ("n_texels", C.c_uint64),
                ("boundary_prims", C.c_void_p), ("n_boundary_prims", C.c_uint32), ("reserved", C.c_uint32)]


class RtContextOptions(C.Structure):
    _fields_ = [("lanes", C.c_uint32), ("queue_capacity", C.c_uint32), ("flags", C.c_uint32), ("gen_blocks", C.c_uint32), ("acc_blocks", C.c_uint32),
                ("reserved", C.c_uint32 * 3)]


class RtSceneOptions(C.Structure):
    _fields_ = [("bvh_builder", C.c_uint32), ("bvh_max_leaf", C.c_uint32), ("bvh_treelet_nodes", C.c_uint32), ("staged_nodes", C.c_uint32),
                ("flags", C.c_uint32), ("bvh_threads", C.c_uint32), ("reserved", C.c_uint32 * 2)]


class RtCamera(C.Structure):
    _fields_ = [("eye", C.c_float * 3), ("lower_left", C.c_float * 3), ("horizontal", C.c_float * 3), ("vertical", C.c_float * 3),
                ("lens_radius", C.c_float), ("reserved", C.c_float * 3)]


class RtRenderParams(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("spp", C.c_uint32), ("max_depth", C.c_uint32), ("seed", C.c_uint64),
                ("row_begin", C.c_uint32), ("row_stride", C.c_uint32), ("queue_capacity", C.c_uint32), ("flags", C.c_uint32)]


class RtStats(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("rays", C.c_uint64), ("iterations", C.c_uint64), ("launches", C.c_uint64), ("ms_total", C.c_float),
                ("ms_generate", C.c_float), ("ms_extend", C.c_float), ("ms_shade", C.c_float), ("ms_accumulate", C.c_float),
                ("bvh_nodes", C.c_uint32), ("bvh_always", C.c_uint32), ("box_tests", C.c_uint64), ("prim_tests", C.c_uint64),
                ("timed_iterations", C.c_uint64)]


class RtScenePlan(C.Structure):
    _fields_ = [("n_box_prims", C.c_uint32), ("n_leaf_items", C.c_uint32), ("n_nodes", C.c_uint32), ("bvh_depth", C.c_uint32), ("n_always", C.c_uint32),
                ("whole_tree_threads", C.c_uint32), ("whole_tree_smem", C.c_uint64)]


class RthCameraArgs(C.Structure):
    _fields_ = [("eye", C.c_double * 3), ("center", C.c_double * 3), ("up", C.c_double * 3), ("fov", C.c_double), ("aspect", C.c_double),
                ("focus_distance", C.c_double), ("lens_radius", C.c_double)]


# every symbol include/rt_b200.h and include/rt_b200_host.h declare
ABI_SYMBOLS = ["rt_init", "rt_init_ex", "rt_options_from_env", "rt_shutdown", "rt_scene_create", "rt_scene_create_ex", "rt_scene_destroy", "rt_render", "rt_render_device",
               "rt_trace", "rt_trace_from", "rt_color_rays", "rt_resolve_u8", "rt_init_multi", "rt_shutdown_multi", "rt_multi_device_count", "rt_multi_context",
               "rt_multi_scene_create", "rt_multi_scene_destroy", "rt_render_multi", "rt_render_noalloc",
               "rt_deinterleave_rows", "rt_scene_bvh_info", "rt_scene_frame_rect_info", "rt_scene_extend_info", "rt_scene_plan", "rt_scene_device_bytes", "rt_last_error", "rt_abi_version", "rth_example_scene", "rth_flat_destroy", "rth_flat_camera_args",
               "rth_flat_desc", "rth_scene_upload", "rth_camera", "rth_encode_image", "rth_decode_image", "rth_set_earth_image"]


class BackendError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"rt_b200 error {code}: {message}")
        self.code = code


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `make -C {_HERE}` (or __graft_entry__.build()); "
                          "this backend has no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, u32, u64 = C.c_void_p, C.c_uint32, C.c_uint64
    L.rt_init.argtypes = [C.c_int, C.POINTER(vp)]
    L.rt_init_ex.argtypes = [C.c_int, C.POINTER(RtContextOptions), C.POINTER(vp)]
    L.rt_options_from_env.argtypes = [C.POINTER(RtContextOptions), C.POINTER(RtSceneOptions)]
    L.rt_options_from_env.restype = None
    L.rt_scene_create_ex.argtypes = [vp, C.POINTER(RtSceneDesc), C.POINTER(RtSceneOptions), C.POINTER(vp)]
    L.rt_trace_from.argtypes = [vp, vp, vp, u64, u64, vp]
    L.rt_init_multi.argtypes = [C.c_int, C.POINTER(C.c_int), C.POINTER(RtContextOptions), C.POINTER(vp)]
    L.rt_shutdown_multi.argtypes = [vp]
    L.rt_shutdown_multi.restype = None
    L.rt_multi_device_count.argtypes = [vp]
    L.rt_multi_context.argtypes = [vp, C.c_int]
    L.rt_multi_context.restype = vp
    L.rt_multi_scene_create.argtypes = [vp, C.POINTER(RtSceneDesc), C.POINTER(RtSceneOptions), C.POINTER(vp)]
    L.rt_multi_scene_destroy.argtypes = [vp]
    L.rt_multi_scene_destroy.restype = None
    L.rt_render_multi.argtypes = [vp, C.POINTER(RtCamera), C.POINTER(RtRenderParams), vp, vp, C.POINTER(RtStats)]
    L.rt_shutdown.argtypes = [vp]
    L.rt_shutdown.restype = None
    L.rt_scene_create.argtypes = [vp, C.POINTER(RtSceneDesc), C.POINTER(vp)]
    L.rt_scene_destroy.argtypes = [vp]
    L.rt_scene_destroy.restype = None
    L.rt_render.argtypes = [vp, C.POINTER(RtCamera), C.POINTER(RtRenderParams), vp, vp, C.POINTER(RtStats)]
    L.rt_render_noalloc.argtypes = [vp, C.POINTER(RtCamera), C.POINTER(RtRenderParams), vp, vp, C.POINTER(RtStats)]
    L.rt_render_device.argtypes = [vp, C.POINTER(RtCamera), C.POINTER(RtRenderParams), vp, vp, vp, C.POINTER(RtStats)]
    L.rt_trace.argtypes = [vp, vp, u64, u64, vp]
    L.rt_color_rays.argtypes = [vp, vp, u64, u32, u32, u64, vp, vp, C.POINTER(RtStats)]
    L.rt_resolve_u8.argtypes = [vp, vp, u32, u32, u32, vp]
    L.rt_deinterleave_rows.argtypes = [vp, vp, vp, u32, u32, u32, vp]
    L.rt_scene_device_bytes.argtypes = [vp]
    L.rt_scene_device_bytes.restype = u64
    L.rt_scene_bvh_info.argtypes = [vp, C.POINTER(u32), C.POINTER(u32), C.POINTER(u32)]
    L.rt_scene_frame_rect_info.argtypes = [vp, C.POINTER(u32), C.POINTER(u32), C.POINTER(u32)]
    L.rt_scene_extend_info.argtypes = [vp, C.POINTER(u32), C.POINTER(u32), C.POINTER(u32)]
    L.rt_scene_plan.argtypes = [C.POINTER(RtSceneDesc), C.POINTER(RtSceneOptions), C.POINTER(RtScenePlan)]
    L.rt_last_error.restype = C.c_char_p
    L.rt_abi_version.restype = u32
    L.rth_example_scene.argtypes = [C.c_char_p, u64, C.POINTER(vp)]
    L.rth_flat_destroy.argtypes = [vp]
    L.rth_flat_destroy.restype = None
    L.rth_flat_camera_args.argtypes = [vp, C.c_double, C.POINTER(RthCameraArgs)]
    L.rth_flat_desc.argtypes = [vp, C.POINTER(RtSceneDesc)]
    L.rth_scene_upload.argtypes = [vp, vp, C.POINTER(vp)]
    L.rth_camera.argtypes = [C.POINTER(RthCameraArgs), C.POINTER(RtCamera)]
    L.rth_encode_image.argtypes = [C.c_int, vp, u32, u32, vp, u64, C.POINTER(u64)]
    L.rth_decode_image.argtypes = [vp, u64, C.POINTER(u32), C.POINTER(u32), vp, u64, C.POINTER(u64)]
    L.rth_set_earth_image.argtypes = [vp, u32, u32]
    return L


lib = _load()


def _check(rc):
    if rc != RT_OK:
        raise BackendError(rc, (lib.rt_last_error() or b"").decode())


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def context_options(**kw):
    """RtContextOptions from the environment defaults (rt_options_from_env), overridden by keyword: lanes, queue_capacity, flags, gen_blocks, acc_blocks."""
    o = RtContextOptions()
    lib.rt_options_from_env(C.byref(o), None)
    for k, v in kw.items():
        setattr(o, k, v)
    return o


def scene_options(**kw):
    """RtSceneOptions from the environment defaults, overridden by keyword: bvh_builder, bvh_max_leaf, bvh_treelet_nodes, staged_nodes (+1), flags."""
    o = RtSceneOptions()
    lib.rt_options_from_env(None, C.byref(o))
    for k, v in kw.items():
        setattr(o, k, v)
    return o


class Device:
    """RtContext bound to one CUDA device.  `options`: keyword overrides of RtContextOptions (None: rt_init, environment defaults)."""

    def __init__(self, device_id=0, options=None, _handle=None):
        self._owned = _handle is None
        self._h = C.c_void_p(_handle) if _handle is not None else C.c_void_p()
        if _handle is None:
            if options is None:
                _check(lib.rt_init(int(device_id), C.byref(self._h)))
            else:
                o = context_options(**options)
                _check(lib.rt_init_ex(int(device_id), C.byref(o), C.byref(self._h)))
        self.device_id = device_id

    def close(self):
        if getattr(self, "_h", None):
            if self._owned:
                lib.rt_shutdown(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def deinterleave_rows(self, d_gathered_ptr, d_full_ptr, width, height, n_shards, stream=None):
        """rt_deinterleave_rows on raw device pointers (multi-GPU framebuffer reassembly)."""
        _check(lib.rt_deinterleave_rows(self._h, C.c_void_p(d_gathered_ptr), C.c_void_p(d_full_ptr), width, height, n_shards,
                                        C.c_void_p(stream) if stream else None))

    def resolve_u8(self, sum_rgb, spp):
        s = np.ascontiguousarray(sum_rgb, dtype=np.float32)
        h, w, _ = s.shape
        out = np.zeros((h, w, 3), dtype=np.uint8)
        _check(lib.rt_resolve_u8(self._h, _ptr(s), w, h, int(spp), _ptr(out)))
        return out


IMAGE_PPM_P3, IMAGE_PNG_RGBA8 = 0, 1


def encode_image(rgb_top_down, fmt):
    """rth_encode_image: the example drivers' PPM P3 (examples/book-one.rs:28-30,90-100) or RGBA8 PNG (examples/main.rs:105-135)
    bytes of a top-down [H, W, 3] uint8 image (the output of Device.resolve_u8).  Host-only: no GPU involved."""
    img = np.ascontiguousarray(rgb_top_down, dtype=np.uint8)
    if img.ndim != 3 or img.shape[2] != 3:
        raise ValueError("expected [H, W, 3] uint8")
    h, w = img.shape[:2]
    need = C.c_uint64(0)
    _check(lib.rth_encode_image(fmt, _ptr(img), w, h, None, 0, C.byref(need)))
    out = np.empty(need.value, dtype=np.uint8)
    _check(lib.rth_encode_image(fmt, _ptr(img), w, h, _ptr(out), need.value, C.byref(need)))
    return out.tobytes()


def decode_image(data):
    """rth_decode_image: PNG (8-bit, non-interlaced) or PPM (P3 / P6) bytes -> top-down [H, W, 3] uint8 (examples/main.rs:266-268)."""
    buf = np.frombuffer(bytes(data), dtype=np.uint8)
    w, h, need = C.c_uint32(0), C.c_uint32(0), C.c_uint64(0)
    _check(lib.rth_decode_image(_ptr(buf), len(buf), C.byref(w), C.byref(h), None, 0, C.byref(need)))
    out = np.empty(need.value, dtype=np.uint8)
    _check(lib.rth_decode_image(_ptr(buf), len(buf), C.byref(w), C.byref(h), _ptr(out), need.value, C.byref(need)))
    return out.reshape(h.value, w.value, 3)


def set_earth_image(rgb_top_down):
    """rth_set_earth_image: the image the "book-two" example maps onto its earth sphere (None = the synthetic stand-in)."""
    if rgb_top_down is None:
        _check(lib.rth_set_earth_image(None, 0, 0))
        return
    img = np.ascontiguousarray(rgb_top_down, dtype=np.uint8)
    _check(lib.rth_set_earth_image(_ptr(img), img.shape[1], img.shape[0]))


def camera_args(eye, center, up, fov, aspect, focus, lens):
    a = RthCameraArgs()
    a.eye[:], a.center[:], a.up[:] = eye, center, up
    a.fov, a.aspect, a.focus_distance, a.lens_radius = fov, aspect, focus, lens
    return a


def make_camera(args):
    """PerspectiveCamera::new (src/camera.rs:25-59) -> RtCamera constants."""
    cam = RtCamera()
    rc = lib.rth_camera(C.byref(args), C.byref(cam))
    if rc != RT_OK:
        raise BackendError(rc, "rth_camera")
    return cam


class FlatScene:
    """An example scene built through the host mirror of the reference API and flattened (CPU only)."""

    def __init__(self, name, seed=1):
        self.name, self.seed = name, seed
        self._h = C.c_void_p()
        rc = lib.rth_example_scene(name.encode(), int(seed), C.byref(self._h))
        if rc != RT_OK:
            raise BackendError(rc, f"cannot build example scene {name!r}")

    def __del__(self):
        if getattr(self, "_h", None):
            lib.rth_flat_destroy(self._h)
            self._h = None

    def desc(self):
        d = RtSceneDesc()
        lib.rth_flat_desc(self._h, C.byref(d))
        return d

    def _table(self, ptr, n, dtype):
        if not n:
            return np.zeros(0, dtype=dtype)
        buf = (C.c_char * (n * dtype.itemsize)).from_address(ptr)
        return np.frombuffer(buf, dtype=dtype).copy()

    def prims(self):
        d = self.desc()
        return self._table(d.prims, d.n_prims, PRIM_DTYPE)

    def boundary_prims(self):
        d = self.desc()
        return self._table(d.boundary_prims, d.n_boundary_prims, PRIM_DTYPE)

    def materials(self):
        d = self.desc()
        return self._table(d.materials, d.n_materials, MATERIAL_DTYPE)

    def textures(self):
        d = self.desc()
        return self._table(d.textures, d.n_textures, TEXTURE_DTYPE)

    def plan(self, options=None):
        """rt_scene_plan: what rt_scene_create_ex would hand to extend (box primitives, leaf items, tree size, whole-tree form) — no device needed."""
        d, out = self.desc(), RtScenePlan()
        _check(lib.rt_scene_plan(C.byref(d), C.byref(options) if options is not None else None, C.byref(out)))
        return {k: int(getattr(out, k)) for k, _ in RtScenePlan._fields_}

    def camera_args(self, aspect):
        a = RthCameraArgs()
        lib.rth_flat_camera_args(self._h, float(aspect), C.byref(a))
        return a

    def camera(self, aspect):
        return make_camera(self.camera_args(aspect))


class Scene:
    """A device-resident scene (RtScene): flattened tables + the rebuilt device BVH."""

    def __init__(self, device, flat=None, desc=None, keepalive=None, options=None):
        """`options`: keyword overrides of RtSceneOptions (None: the context's defaults)."""
        self.device = device
        self.flat = flat
        self._keep = keepalive
        self._h = C.c_void_p()
        if flat is not None:
            desc = flat.desc()
        if options is None:
            _check(lib.rt_scene_create(device._h, C.byref(desc), C.byref(self._h)))
        else:
            o = scene_options(**options)
            _check(lib.rt_scene_create_ex(device._h, C.byref(desc), C.byref(o), C.byref(self._h)))

    @classmethod
    def example(cls, device, name, seed=1, options=None):
        return cls(device, flat=FlatScene(name, seed), options=options)

    @staticmethod
    def tables_desc(prims, materials, textures, texels=None, boundary_prims=None):
        prims = np.ascontiguousarray(prims, dtype=PRIM_DTYPE)
        materials = np.ascontiguousarray(materials, dtype=MATERIAL_DTYPE)
        textures = np.ascontiguousarray(textures, dtype=TEXTURE_DTYPE)
        texels = np.ascontiguousarray(texels if texels is not None else np.zeros((0, 4)), dtype=np.float32)
        bnd = np.ascontiguousarray(boundary_prims if boundary_prims is not None else np.zeros(0, dtype=PRIM_DTYPE), dtype=PRIM_DTYPE)
        d = RtSceneDesc(lib.rt_abi_version(), len(prims), len(materials), len(textures), prims.ctypes.data, materials.ctypes.data,
                        textures.ctypes.data, texels.ctypes.data, texels.size // 4, bnd.ctypes.data, len(bnd), 0)
        return d, (prims, materials, textures, texels, bnd)

    @staticmethod
    def plan_tables(prims, materials, textures, texels=None, boundary_prims=None, options=None):
        """rt_scene_plan on flat tables (host only, no device): what extend would see."""
        d, keep = Scene.tables_desc(prims, materials, textures, texels, boundary_prims)
        out = RtScenePlan()
        _check(lib.rt_scene_plan(C.byref(d), C.byref(options) if options is not None else None, C.byref(out)))
        return {k: int(getattr(out, k)) for k, _ in RtScenePlan._fields_}

    @classmethod
    def from_tables(cls, device, prims, materials, textures, texels=None, boundary_prims=None, options=None):
        d, keep = cls.tables_desc(prims, materials, textures, texels, boundary_prims)
        return cls(device, desc=d, keepalive=keep, options=options)

    def close(self):
        if getattr(self, "_h", None):
            lib.rt_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def bvh_info(self):
        n, a, d = C.c_uint32(), C.c_uint32(), C.c_uint32()
        _check(lib.rt_scene_bvh_info(self._h, C.byref(n), C.byref(a), C.byref(d)))
        return {"nodes": n.value, "always": a.value, "depth": d.value}

    def frame_rect_info(self):
        f, b, l = C.c_uint32(), C.c_uint32(), C.c_uint32()
        _check(lib.rt_scene_frame_rect_info(self._h, C.byref(f), C.byref(b), C.byref(l)))
        return {"frames": f.value, "boxes": b.value, "loose": l.value}

    def extend_info(self):
        b, n, t = C.c_uint32(), C.c_uint32(), C.c_uint32()
        _check(lib.rt_scene_extend_info(self._h, C.byref(b), C.byref(n), C.byref(t)))
        return {"box_prims": b.value, "leaf_items": n.value, "whole_tree_threads": t.value}

    def device_bytes(self):
        return int(lib.rt_scene_device_bytes(self._h))

    @staticmethod
    def params(width, height, spp, max_depth, seed=1, row_begin=0, row_stride=1, queue_capacity=0, flags=0):
        return RtRenderParams(width, height, spp, max_depth, seed, row_begin, row_stride, queue_capacity, flags)

    def render(self, cam, width, height, spp, max_depth, seed=1, row_begin=0, row_stride=1, queue_capacity=0, want_sumsq=True, flags=0, out=None):
        """rt_render with host buffers: returns (sum[rows,W,3], sumsq or None, RtStats); rows y = row_begin + k*row_stride.
        `out`: a caller-owned C-contiguous float32 [rows, W, 3] buffer to receive the sums."""
        prm = self.params(width, height, spp, max_depth, seed, row_begin, row_stride, queue_capacity, flags)
        rows = len(range(row_begin, height, row_stride)) if row_stride > 0 else 0  # the library rejects stride 0 / empty shards
        if out is not None:
            assert out.dtype == np.float32 and out.flags["C_CONTIGUOUS"] and out.shape == (rows, width, 3)
        s = out if out is not None else np.zeros((max(rows, 1), width, 3), dtype=np.float32)[:rows]
        q = np.zeros((max(rows, 1), width, 3), dtype=np.float32)[:rows] if want_sumsq else None
        st = RtStats()
        _check(lib.rt_render(self._h, C.byref(cam), C.byref(prm), _ptr(s), _ptr(q) if want_sumsq else None, C.byref(st)))
        return s, q, st

    def render_noalloc(self, cam, width, height, spp, max_depth, seed=1, row_begin=0, row_stride=1, want_sumsq=True, flags=0):
        """rt_render_noalloc: rt_render out of the library's static device scratch (no allocation; width x rows <= NOALLOC_MAX_PIXELS)."""
        prm = self.params(width, height, spp, max_depth, seed, row_begin, row_stride, 0, flags)
        rows = len(range(row_begin, height, row_stride)) if row_stride > 0 else 0
        s = np.zeros((max(rows, 1), width, 3), dtype=np.float32)[:rows]
        q = np.zeros((max(rows, 1), width, 3), dtype=np.float32)[:rows] if want_sumsq else None
        st = RtStats()
        _check(lib.rt_render_noalloc(self._h, C.byref(cam), C.byref(prm), _ptr(s), _ptr(q) if want_sumsq else None, C.byref(st)))
        return s, q, st

    def render_device(self, cam, prm, d_sum_ptr, d_sumsq_ptr=None, stream=None):
        """rt_render_device: outputs stay in HBM (raw device pointers, e.g. torch tensor .data_ptr())."""
        st = RtStats()
        _check(lib.rt_render_device(self._h, C.byref(cam), C.byref(prm), C.c_void_p(d_sum_ptr), C.c_void_p(d_sumsq_ptr) if d_sumsq_ptr else None,
                                    C.c_void_p(stream) if stream else None, C.byref(st)))
        return st

    def trace(self, rays, seed=1, self_order=None):
        """rt_trace / rt_trace_from: rays [n,6] (origin, direction) f32 -> structured array HIT_DTYPE.  self_order [n] u32: DFS rank of
        the primitive each ray starts on (NO_PRIM: none)."""
        r = np.ascontiguousarray(np.asarray(rays, dtype=np.float32).reshape(-1, 6))
        out = np.zeros(len(r), dtype=HIT_DTYPE)
        if self_order is None:
            _check(lib.rt_trace(self._h, _ptr(r), len(r), int(seed), _ptr(out)))
        else:
            so = np.ascontiguousarray(np.asarray(self_order, dtype=np.uint32))
            assert len(so) == len(r)
            _check(lib.rt_trace_from(self._h, _ptr(r), _ptr(so), len(r), int(seed), _ptr(out)))
        return out

    def color_rays(self, rays, spp, max_depth, seed=1, want_sumsq=False):
        """rt_color_rays: batched render::color on rays [n,6] f32 -> (sum[n,3], sumsq[n,3] or None, RtStats)."""
        r = np.ascontiguousarray(np.asarray(rays, dtype=np.float32).reshape(-1, 6))
        s = np.zeros((len(r), 3), dtype=np.float32)
        q = np.zeros((len(r), 3), dtype=np.float32) if want_sumsq else None
        st = RtStats()
        _check(lib.rt_color_rays(self._h, _ptr(r), len(r), int(spp), int(max_depth), int(seed), _ptr(s), _ptr(q) if want_sumsq else None, C.byref(st)))
        return s, q, st


class MultiDevice:
    """RtMulti: several GPUs driven by ONE process through the C ABI (rt_init_multi / rt_render_multi: one host thread per device,
    rows y % n == device, one NCCL gather to device 0)."""

    def __init__(self, device_ids, options=None):
        ids = (C.c_int * len(device_ids))(*device_ids)
        self._h = C.c_void_p()
        o = context_options(**options) if options is not None else None
        _check(lib.rt_init_multi(len(device_ids), ids, C.byref(o) if o is not None else None, C.byref(self._h)))
        self.device_ids = list(device_ids)

    def context(self, index):
        return Device(self.device_ids[index], _handle=lib.rt_multi_context(self._h, index))

    def close(self):
        if getattr(self, "_h", None):
            lib.rt_shutdown_multi(self._h)
            self._h = None

    def __del__(self):
        self.close()


class MultiScene:
    def __init__(self, multi, flat=None, desc=None, keepalive=None, options=None):
        self.multi, self.flat, self._keep = multi, flat, keepalive
        self._h = C.c_void_p()
        if flat is not None:
            desc = flat.desc()
        o = scene_options(**options) if options is not None else None
        _check(lib.rt_multi_scene_create(multi._h, C.byref(desc), C.byref(o) if o is not None else None, C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None):
            lib.rt_multi_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def render(self, cam, width, height, spp, max_depth, seed=1, queue_capacity=0, want_sumsq=True, flags=0, out=None):
        """rt_render_multi with host buffers: (sum[H,W,3], sumsq or None, RtStats)."""
        prm = RtRenderParams(width, height, spp, max_depth, seed, 0, 1, queue_capacity, flags)
        s = out if out is not None else np.zeros((height, width, 3), dtype=np.float32)
        q = np.zeros((height, width, 3), dtype=np.float32) if want_sumsq else None
        st = RtStats()
        _check(lib.rt_render_multi(self._h, C.byref(cam), C.byref(prm), _ptr(s), _ptr(q) if want_sumsq else None, C.byref(st)))
        return s, q, st
