"""Generates the committed fixtures in tests/golden/ from the f64 oracle (oracle/liboracle.so).

The reference itself cannot produce vectors (Rust, OS-seeded RNG, no toolchain here — SURVEY §8(c)), so
these pin the ORACLE's behaviour at the commit that generated them: tests/test_golden.py re-checks the
oracle against them on every run (an edit that changes oracle semantics shows up as a diff here) and the
GPU tests compare the CUDA path with the same files.  Run from the repo root:
    python tests/golden/make_golden.py [scene ...]      (no argument: every scene; name scenes to leave the other fixtures untouched)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_py as orc  # noqa: E402

SCENES = ["book-one", "cornell-box", "book-two", "stress:5000", "checker-lists"]


def fixed_rays(scene, w=24, h=16, n_bounce=600, seed=7):
    """Camera rays at pixel centres + bounce rays leaving oracle hit points in random directions (f32-exact values)."""
    cam = scene.camera(w / h)
    cam_rays = orc.camera_rays(cam, w, h).astype(np.float32)
    hits = scene.trace(cam_rays.astype(np.float64), rel_perturb=0.0)
    pts = hits["p"][hits["hit"] == 1]
    rng = np.random.default_rng(seed)
    idx = rng.integers(0, len(pts), n_bounce)
    d = rng.normal(size=(n_bounce, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    o = pts[idx] + 1e-3 * d * np.maximum(1.0, np.abs(pts[idx]).max(axis=1, keepdims=True))
    return np.concatenate([cam_rays, np.concatenate([o, d], axis=1).astype(np.float32)], axis=0)


def main():
    for name in (sys.argv[1:] or SCENES):
        s = orc.Scene(name, 1)
        rays = fixed_rays(s)
        h = s.trace(rays.astype(np.float64), seed=1)
        w, hh, spp, depth = 20, 12, 4, 50
        img, sq, st = s.render(w, hh, spp, depth, seed=1, mode=orc.PAIRED)
        out = os.path.join(HERE, name.replace(":", "_") + ".npz")
        np.savez_compressed(out, rays=rays, hit=h["hit"].astype(np.int8), prim=h["prim"], stable=h["stable"].astype(np.int8),
                            t=h["t"], p=h["p"], n=h["n"], uv=np.stack([h["u"], h["v"]], axis=1), t_sens=h["t_sens"],
                            p_sens=h["p_sens"], n_sens=h["n_sens"], paired_sum=img, paired_shape=np.array([w, hh, spp, depth]),
                            paired_segments=np.array([st.segments]))
        print(name, len(rays), "stable", h["stable"].mean(), "hits", h["hit"].mean(), os.path.getsize(out))


if __name__ == "__main__":
    main()
