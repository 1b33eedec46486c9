"""Exploratory GPU check (diagnostics, not a test): trace parity, paired-render parity and timing.
Run on the GPU box:  python scripts/gpu_check.py [quick]"""
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
rt = importlib.import_module("ray-tracer_b200")
import oracle_py as orc  # noqa: E402  (checker only)


def rays_for(scene_o, cam_o, w, h, n_random, seed):
    rng = np.random.default_rng(seed)
    cam_rays = orc.camera_rays(cam_o, w, h).astype(np.float32)
    # bounce rays: start at oracle hit points of the camera rays, random directions
    hits = scene_o.trace(cam_rays.astype(np.float64), rel_perturb=0.0)
    pts = hits["p"][hits["hit"] == 1]
    idx = rng.integers(0, len(pts), n_random)
    d = rng.normal(size=(n_random, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    o = pts[idx] + 1e-3 * d * np.maximum(1.0, np.abs(pts[idx]).max(axis=1, keepdims=True))
    bounce = np.concatenate([o, d], axis=1).astype(np.float32)
    return np.concatenate([cam_rays, bounce], axis=0)


def trace_parity(dev, name, w=64, h=36, n_random=4000):
    so = orc.Scene(name, 1)
    sg = rt.Scene.example(dev, name, 1)
    cam_o = so.camera(w / h)
    rays = rays_for(so, cam_o, w, h, n_random, 7)
    ho = so.trace(rays.astype(np.float64), seed=1)
    hg = sg.trace(rays, seed=1)
    stable = ho["stable"] == 1
    hm = (ho["hit"] == hg["hit"])
    pm = (ho["prim"] == hg["prim"])
    both = (ho["hit"] == 1) & (hg["hit"] == 1) & pm
    relt = np.abs(hg["t"][both] - ho["t"][both]) / np.maximum(np.abs(ho["t"][both]), 1e-6)
    scale = np.maximum(1.0, np.abs(ho["p"][both]).max(axis=1))
    relp = np.abs(hg["p"][both] - ho["p"][both]).max(axis=1) / scale
    reln = np.abs(hg["n"][both] - ho["n"][both]).max(axis=1)
    print(f"[trace {name}] rays={len(rays)} stable={stable.mean():.4f} bvh={sg.bvh_info()} hit-mismatch(all)={(~hm).sum()} "
          f"prim-mismatch(all)={(~pm).sum()} mismatch(stable)={(~(hm & pm) & stable).sum()} "
          f"max rel t={relt.max():.2e} p={relp.max():.2e} n={reln.max():.2e} hits={ho['hit'].mean():.3f}")
    bad = np.nonzero(~(hm & pm) & stable)[0][:5]
    for i in bad:
        print("   bad ray", i, rays[i], "oracle", ho[i]["hit"], ho[i]["prim"], ho[i]["t"], "gpu", hg[i]["hit"], hg[i]["prim"], hg[i]["t"])
    return sg, so


def paired_render(dev, name, w, h, spp, depth):
    so = orc.Scene(name, 1)
    sg = rt.Scene.example(dev, name, 1)
    cam = sg.flat.camera(w / h)
    s_g, q_g, st = sg.render(cam, w, h, spp, depth, seed=1)
    s_o, q_o, sto = so.render(w, h, spp, depth, seed=1, mode=orc.PAIRED)
    d = np.abs(s_g - s_o).max(axis=2)
    tol = 1e-3 * np.maximum(1.0, np.abs(s_o).max(axis=2))
    print(f"[paired {name}] {w}x{h}x{spp} gpu rays/path={st.rays / st.paths:.3f} oracle={sto.segments / sto.paths:.3f} "
          f"pixels equal={np.mean(d <= tol):.4f} mean gpu={s_g.mean() / spp:.5f} oracle={s_o.mean() / spp:.5f} "
          f"iters={st.iterations} ms={st.ms_total:.2f}")


def timing(dev, name, w, h, spp, depth, cap=0):
    sg = rt.Scene.example(dev, name, 1)
    cam = sg.flat.camera(w / h)
    sg.render(cam, w, h, max(1, spp // 8), depth, queue_capacity=cap)
    t0 = time.time()
    s, q, st = sg.render(cam, w, h, spp, depth, queue_capacity=cap)
    wall = time.time() - t0
    print(f"[time {name}] {w}x{h}x{spp} d{depth} cap={cap} paths={st.paths} rays={st.rays} iters={st.iterations} dev_ms={st.ms_total:.2f} wall_ms={wall * 1e3:.1f} "
          f"Mpaths/s={st.paths / st.ms_total / 1e3:.1f} Grays/s={st.rays / st.ms_total / 1e6:.3f} "
          f"stages(ms) g={st.ms_generate:.1f} e={st.ms_extend:.1f} s={st.ms_shade:.1f} a={st.ms_accumulate:.1f} mean={s.mean() / spp:.4f}")


if __name__ == "__main__":
    quick = len(sys.argv) > 1 and sys.argv[1] == "quick"
    dev = rt.Device(0)
    for nm in ["book-one", "cornell-box", "book-two", "stress:20000"]:
        trace_parity(dev, nm)
    paired_render(dev, "book-one", 160, 90, 2, 50)
    paired_render(dev, "cornell-box", 100, 100, 4, 50)
    paired_render(dev, "book-two", 100, 100, 2, 50)
    timing(dev, "book-one", 400, 225, 100, 50)
    timing(dev, "cornell-box", 600, 600, 64, 50)
    if not quick:
        timing(dev, "book-one", 1920, 1080, 64, 50)
        timing(dev, "book-one", 1920, 1080, 64, 50, cap=1 << 19)
        timing(dev, "book-one", 1920, 1080, 64, 50, cap=1 << 22)
        timing(dev, "book-two", 800, 800, 32, 50)
        timing(dev, "stress:1000000", 1920, 1080, 8, 16)
