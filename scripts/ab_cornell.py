"""A/B of the cornell-box extend forms:  python scripts/ab_cornell.py   (frame boxes / loose frame rects / general records)"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rt = importlib.import_module("ray-tracer_b200")
dev = rt.Device(0)
w = h = 600
for label, opts in [("frame boxes", None), ("loose frame rects", {"flags": rt.SCENE_NO_FRAME_BOXES}), ("general records", {"flags": rt.SCENE_NO_FRAME_RECTS})]:
    sc = rt.Scene.example(dev, "cornell-box", 1, options=opts)
    cam = sc.flat.camera(1.0)
    sc.render(cam, w, h, 4, 50, want_sumsq=False)
    best = None
    for rep in range(3):
        _, _, st = sc.render(cam, w, h, 128, 50, flags=rt.RENDER_STAGE_TIMING, want_sumsq=False)
        if best is None or st.ms_extend < best.ms_extend:
            best = st
    st = best
    _, _, full = sc.render(cam, w, h, 256, 50, want_sumsq=False)
    print(f"{label:18s} {sc.frame_rect_info()} extend {st.rays / st.ms_extend / 1e6:6.2f} Grays/s ({st.ms_extend:.2f} ms) shade {st.rays / st.ms_shade / 1e6:6.2f} "
          f"single-lane {st.paths / st.ms_total / 1e3:7.1f} Mpaths/s | two lanes {full.paths / full.ms_total / 1e3:7.1f} Mpaths/s rays {st.rays}", flush=True)
