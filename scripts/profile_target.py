"""Short render used as the ncu target:  python scripts/profile_target.py <scene> <w> <h> <spp> <depth> [capacity]"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rt = importlib.import_module("ray-tracer_b200")

name, w, h, spp, depth = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
cap = int(sys.argv[6]) if len(sys.argv) > 6 else 0
dev = rt.Device(0)
scene = rt.Scene.example(dev, name, 1)
cam = scene.flat.camera(w / h)
fl = int(os.environ.get("AB_FLAGS", "0"))  # extra RtRenderParams.flags (8 = extend per thread, 16 = persistent / refill)
s, _, st = scene.render(cam, w, h, spp, depth, seed=1, queue_capacity=cap, want_sumsq=False, flags=fl)
print(f"{name} {w}x{h}x{spp} d{depth}: paths={st.paths} rays={st.rays} iters={st.iterations} launches={st.launches} dev_ms={st.ms_total:.2f} "
      f"Mpaths/s={st.paths / st.ms_total / 1e3:.1f} Grays/s={st.rays / st.ms_total / 1e6:.3f} mean={s.mean() / spp:.4f}")
