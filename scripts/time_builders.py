"""rt_scene_create (BVH build + upload) time and render rate, host SAH vs device LBVH:  python scripts/time_builders.py [scene ...]"""
import importlib
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rt = importlib.import_module("ray-tracer_b200")

names = sys.argv[1:] or ["stress:1000000", "book-two", "book-one"]
dev = rt.Device(0)
for name in names:
    flat = rt.FlatScene(name, 1)
    for builder in ["host", "device"]:
        os.environ["RT_B200_BVH_BUILDER"] = builder
        rt.Scene(dev, flat=flat).close()  # warm-up (allocator, cub temp sizes)
        best = 1e9
        for _ in range(3):
            t0 = time.perf_counter()
            sc = rt.Scene(dev, flat=flat)
            best = min(best, time.perf_counter() - t0)
            if _ < 2:
                sc.close()
        w, h, spp, d = (1920, 1080, 4, 16) if name.startswith("stress") else ((800, 800, 16, 50) if name == "book-two" else (1920, 1080, 16, 50))
        cam = flat.camera(w / h)
        sc.render(cam, w, h, 1, d, want_sumsq=False)
        _, _, st = sc.render(cam, w, h, spp, d, flags=rt.RENDER_STAGE_TIMING | 0, want_sumsq=False)
        _, _, stw = sc.render(cam, w, h, 1, d, flags=rt.RENDER_COUNT_WORK, want_sumsq=False)
        print(f"{name:16s} builder={builder:6s} rt_scene_create {best * 1e3:8.1f} ms  bvh {sc.bvh_info()}  render {st.paths / st.ms_total / 1e3:7.1f} Mpaths/s "
              f"extend {st.rays / st.ms_extend / 1e6:5.2f} Grays/s  box tests/ray {stw.box_tests / stw.rays:5.1f} prim tests/ray {stw.prim_tests / stw.rays:4.1f}", flush=True)
        sc.close()
os.environ.pop("RT_B200_BVH_BUILDER", None)
