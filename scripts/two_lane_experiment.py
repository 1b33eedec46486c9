"""Would two interleaved wavefronts on one GPU fill each other's kernel tails?  Two contexts (two streams) render the two row
parities of one image from two host threads; compare with one context rendering the whole image."""
import importlib
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rt = importlib.import_module("ray-tracer_b200")

name, w, h, spp, d = (sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])) if len(sys.argv) > 5 else ("cornell-box", 600, 600, 256, 50)
flat = rt.FlatScene(name, 1)
cam = flat.camera(w / h)
devs = [rt.Device(0) for _ in range(3)]
scenes = [rt.Scene(dv, flat=flat) for dv in devs]


def one(cap):
    t0 = time.perf_counter()
    _, _, st = scenes[0].render(cam, w, h, spp, d, seed=1, want_sumsq=False, queue_capacity=cap)
    return time.perf_counter() - t0, st.paths


def two(cap):
    out = [None, None]

    def work(i):
        out[i] = scenes[1 + i].render(cam, w, h, spp, d, seed=1, row_begin=i, row_stride=2, want_sumsq=False, queue_capacity=cap)
    ts = [threading.Thread(target=work, args=(i,)) for i in range(2)]
    t0 = time.perf_counter()
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    return time.perf_counter() - t0, sum(o[2].paths for o in out)


for cap1, cap2 in [(1 << 22, 1 << 21), (1 << 22, 1 << 22), (1 << 23, 1 << 22)]:
    one(cap1); two(cap2)
    a = min(one(cap1) for _ in range(3))
    b = min(two(cap2) for _ in range(3))
    print(f"{name} {w}x{h}x{spp}: one wavefront cap {cap1}: {a[1] / a[0] / 1e6:8.1f} Mpaths/s   two wavefronts cap {cap2} each: {b[1] / b[0] / 1e6:8.1f} Mpaths/s", flush=True)
