"""What one GPU of an 8-GPU cornell-box run does: rows y % 8 == 0 of 600x600x1024, timed per lanes / queue capacity.
    python scripts/shard_timing.py [row_stride]"""
import importlib
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
stride = int(sys.argv[1]) if len(sys.argv) > 1 else 8
for lanes in ("1", "2", "3", "4"):
    os.environ["RT_B200_LANES"] = lanes
    rt = importlib.import_module("ray-tracer_b200")
    dev = rt.Device(0)
    sc = rt.Scene.example(dev, "cornell-box", 1)
    cam = sc.flat.camera(1.0)
    for cap in (0, 1 << 21, 1 << 20):
        sc.render(cam, 600, 600, 64, 50, row_begin=0, row_stride=stride, queue_capacity=cap, want_sumsq=False)
        best = None
        for rep in range(3):
            t0 = time.perf_counter()
            _, _, st = sc.render(cam, 600, 600, 1024, 50, seed=rep, row_begin=0, row_stride=stride, queue_capacity=cap, want_sumsq=False)
            wall = (time.perf_counter() - t0) * 1e3
            if best is None or st.ms_total < best[0]:
                best = (st.ms_total, wall, st)
        ms, wall, st = best
        print(f"stride {stride} lanes {lanes} cap {cap or 1 << 22:8d}: dev {ms:7.2f} ms wall {wall:7.2f} ms  Mpaths/s {st.paths / ms / 1e3:7.1f} iters {st.iterations} launches {st.launches}", flush=True)
    sc.close()
    dev.close()
