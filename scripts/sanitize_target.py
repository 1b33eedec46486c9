"""Tiny renders through every kernel form, as a compute-sanitizer target:
    compute-sanitizer --tool memcheck|racecheck|initcheck|synccheck python scripts/sanitize_target.py
Covers k_generate, k_extend<., false / true>, k_extend_bvh (cp.async ray feed, per-warp shared buffers), k_extend_rects (frame boxes and
loose rectangles), k_shade<1 / 2> (block-wide slot allocation), k_accumulate (segmented warp reduction + f64 atomics), k_finish,
k_trace (incl. surface origins), k_resolve_u8, k_deinterleave (wavefront lanes), the LBVH build kernels, the general records and
both medium forms — each with the default 4 M-ray queue and with a queue of 1056 rays (not a multiple of a block)."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rt = importlib.import_module("ray-tracer_b200")

dev = rt.Device(0, options={"lanes": 2})
cases = [("book-one", None, 0), ("cornell-box", None, 0), ("cornell-box", {"flags": rt.SCENE_NO_FRAME_BOXES}, 0), ("cornell-box", {"flags": rt.SCENE_NO_FRAME_RECTS}, 0),
         ("book-two", None, rt.RENDER_EXTEND_PERSISTENT), ("book-two", None, 0), ("checker-lists", None, 0), ("checker-lists", {"flags": rt.SCENE_NO_SMALL_LIST}, 0),
         ("stress:70000", None, 0), ("stress:70000", {"bvh_builder": rt.BVH_HOST_SAH}, rt.RENDER_EXTEND_PER_THREAD), ("book-one", None, rt.RENDER_COUNT_WORK),
         ("book-one", None, rt.RENDER_SAMPLE_MAJOR | rt.RENDER_STAGE_TIMING)]
for name, opts, flags in cases:
    sc = rt.Scene.example(dev, name, 1, options=opts)
    cam = sc.flat.camera(16 / 9)
    for cap in (0, 1056):
        s, q, st = sc.render(cam, 64, 36, 2, 50, seed=1, queue_capacity=cap, flags=flags)
    rng = np.random.default_rng(1)
    rays = rng.normal(size=(999, 6)).astype(np.float32)
    h = sc.trace(rays)
    on = h["hit"] == 1
    if on.any():  # scattered rays that start on the primitives just hit
        d = rng.normal(size=(int(on.sum()), 3)).astype(np.float32)
        sc.trace(np.concatenate([h["p"][on], d], axis=1), self_order=h["prim"][on].astype(np.uint32))
    c, _, _ = sc.color_rays(rays, 2, 8)
    print(name, opts, flags, sc.bvh_info(), sc.frame_rect_info(), "rays", st.rays, "mean", float(s.mean()), flush=True)
    sc.close()
# enough paths for two wavefront lanes (render_lanes + k_deinterleave), ragged rows
sc = rt.Scene.example(dev, "cornell-box", 1)
s, q, st = sc.render(sc.flat.camera(1.0), 301, 203, 36, 50, seed=3)
print("lanes", st.launches, float(s.mean()))
# the same scene with extend as its own kernel (k_extend_rects + k_shade<2>) and with two rays per round in the fused kernel; rt_render_noalloc
for o in ({"lanes": 2, "flags": rt.CTX_NO_FUSED_RECTS}, {"lanes": 1, "flags": rt.CTX_SHADE_ONE_RAY | rt.CTX_NO_FUSED_RECTS}):
    dev2 = rt.Device(0, options=o)
    sc2 = rt.Scene.example(dev2, "cornell-box", 1)
    for cap in (0, 1056):
        s2, _, st2 = sc2.render(sc2.flat.camera(1.0), 301, 203, 36, 50, seed=3, queue_capacity=cap)
    print("unfused", o, st2.launches, float(s2.mean()), st2.rays == st.rays)
    s3, _, st3 = sc2.render_noalloc(sc2.flat.camera(1.0), 128, 64, 9, 50, seed=3)
    print("noalloc", st3.launches, float(s3.mean()))
    sc2.close()
    dev2.close()
img = dev.resolve_u8(s, 36)
print("ok", img.shape)
