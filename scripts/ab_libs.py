"""A/B of library variants on one workload, single lane, per-stage:  python scripts/ab_libs.py <scene> <w> <h> <spp> <depth> <lib.so> [<lib.so> ...]
Each variant runs in its own subprocess (RT_B200_LIB), interleaved over 3 rounds; best stage times per variant are printed."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import importlib, json, sys
sys.path.insert(0, %r)
rt = importlib.import_module("ray-tracer_b200")
name, w, h, spp, depth = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
dev = rt.Device(0)
sc = rt.Scene.example(dev, name, 1)
cam = sc.flat.camera(w / h)
import os
fl = int(os.environ.get("AB_FLAGS", "0"))  # extra RtRenderParams.flags, e.g. 16 = RT_RENDER_EXTEND_PERSISTENT, 8 = .._PER_THREAD
sc.render(cam, w, h, 8, depth, want_sumsq=False, flags=fl)
_, _, st = sc.render(cam, w, h, spp, depth, flags=rt.RENDER_STAGE_TIMING | fl, want_sumsq=False)
_, _, two = sc.render(cam, w, h, spp, depth, want_sumsq=False, flags=fl)
_, _, two = sc.render(cam, w, h, spp, depth, want_sumsq=False, flags=fl)
print(json.dumps({"g": st.ms_generate, "e": st.ms_extend, "s": st.ms_shade, "a": st.ms_accumulate, "one": st.ms_total, "two": two.ms_total, "rays": st.rays, "paths": st.paths}))
''' % ROOT
scene = sys.argv[1:6]
libs = sys.argv[6:]
best = {}
for rnd in range(3):
    for lib in libs:
        env = dict(os.environ, RT_B200_LIB=os.path.abspath(lib))
        out = subprocess.run([sys.executable, "-c", CHILD, *scene], env=env, capture_output=True, text=True)
        if out.returncode != 0:
            print(lib, "FAILED", out.stderr[-500:])
            continue
        r = json.loads(out.stdout.strip().splitlines()[-1])
        b = best.setdefault(lib, dict(r))
        for k in ("g", "e", "s", "a", "one", "two"):
            b[k] = min(b[k], r[k])
for lib, b in best.items():
    print(f"{os.path.basename(lib):36s} gen {b['g']:7.2f} extend {b['e']:8.2f} ({b['rays'] / max(b['e'], 1e-9) / 1e6:6.2f} Grays/s) shade {b['s']:8.2f} ({b['rays'] / b['s'] / 1e6:6.2f}) acc {b['a']:6.2f} | "
          f"one lane {b['one']:8.2f} ms, two lanes {b['two']:8.2f} ms = {b['paths'] / b['two'] / 1e3:7.1f} Mpaths/s", flush=True)
