"""Per-stage timing of the wavefront on the four scenes:  python scripts/time_scenes.py [scene ...]
(RT_B200_LIB selects a tuning variant of the library.)"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rt = importlib.import_module("ray-tracer_b200")

CASES = {"book-one": ("book-one", 1920, 1080, 16, 50), "cornell-box": ("cornell-box", 600, 600, 64, 50),
         "book-two": ("book-two", 800, 800, 16, 50), "stress": ("stress:1000000", 1920, 1080, 4, 16)}
names = sys.argv[1:] or list(CASES)
dev = rt.Device(0)
tag = os.path.basename(os.environ.get("RT_B200_LIB", "default"))
for nm in names:
    name, w, h, spp, d = CASES[nm]
    sc = rt.Scene.example(dev, name, 1)
    cam = sc.flat.camera(w / h)
    sc.render(cam, w, h, 1, d, want_sumsq=False)
    best = None
    for rep in range(2):
        _, _, st = sc.render(cam, w, h, spp, d, flags=rt.RENDER_STAGE_TIMING, want_sumsq=False)
        if best is None or st.ms_total < best.ms_total:
            best = st
    st = best
    print(f"[{tag:14s}] {nm:12s} Mpaths/s {st.paths / st.ms_total / 1e3:8.1f} Grays/s {st.rays / st.ms_total / 1e6:6.2f} | extend {st.rays / st.ms_extend / 1e6:6.2f} "
          f"shade {st.rays / st.ms_shade / 1e6:6.2f} Grays/s | ms tot {st.ms_total:7.2f} g {st.ms_generate:6.2f} e {st.ms_extend:7.2f} s {st.ms_shade:6.2f} "
          f"a {st.ms_accumulate:5.2f} iters {st.iterations}", flush=True)
