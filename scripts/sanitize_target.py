"""Tiny renders through every kernel form, as a compute-sanitizer target:
    compute-sanitizer --tool memcheck python scripts/sanitize_target.py"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rt = importlib.import_module("ray-tracer_b200")

dev = rt.Device(0)
cases = [("book-one", {}), ("cornell-box", {}), ("cornell-box", {"RT_B200_RECT_FRAMES": "0"}), ("book-two", {"RT_B200_PERSISTENT": "1"}),
         ("book-one", {"RT_B200_SHADE_PIPE": "1"}), ("stress:70000", {}), ("stress:70000", {"RT_B200_BVH_BUILDER": "host", "RT_B200_PERSISTENT": "0"})]
for name, env in cases:
    os.environ.update(env)
    sc = rt.Scene.example(dev, name, 1)
    cam = sc.flat.camera(16 / 9)
    for cap in (0, 1056):
        s, q, st = sc.render(cam, 64, 36, 2, 50, seed=1, queue_capacity=cap)
    rays = np.random.default_rng(1).normal(size=(999, 6)).astype(np.float32)
    sc.trace(rays)
    c, _, _ = sc.color_rays(rays, 2, 8)
    print(name, env, sc.bvh_info(), "rays", st.rays, "mean", float(s.mean()), flush=True)
    sc.close()
    for k in env:
        os.environ.pop(k, None)
img = dev.resolve_u8(s, 2)
print("ok", img.shape)
