"""Parity of the CUDA path (through the C ABI) against the f64 oracle on the same seeded inputs.

Gate 1 (BASELINE north_star): on fixed ray sets, hit/miss and closest-primitive id match BIT-EXACTLY on every
ray the oracle's margin policy calls stable (no near-tie with another primitive within 1e-4 relative, no
hit/primitive flip under 1e-6 relative perturbations of the ray — oracle/oracle_capi.cpp orc_trace); t, hit
point and normal are within 1e-4 relative, plus the query's own f64 conditioning where a grazing hit
amplifies one f32 ulp of the inputs beyond that.
Gate 2: a rendered image at >= 1024 spp has per-pixel means within 3 sigma of its Monte-Carlo standard error
of a high-spp oracle render (exceedance fraction bounded, the chance rate is 0.27 %), and PSNR >= 40 dB on
the resolved (sqrt-gamma) image.
"""
import math
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REL_TOL = 1e-4  # f32 tolerance stated by BASELINE.json north_star


def _golden(name):
    return np.load(os.path.join(GOLDEN, name.replace(":", "_") + ".npz"))


# Floors of (stable fraction, fraction of stable hits inside the BARE 1e-4 bound) per scene, set 0.004 below what the B200 measures
# (r2: book-one 0.9994 / 0.9960, cornell-box 0.9844 / 1.0000, book-two 0.9646 / 0.9979, stress:100000 0.9992 / 0.9853,
# checker-lists 0.9994 / 0.9998): the conditioning slack in the tolerances below must stay the exception, at its measured size.
TRACE_FLOORS = {"book-one": (0.995, 0.992), "cornell-box": (0.970, 0.996), "book-two": (0.960, 0.993), "stress": (0.995, 0.981),
                "checker-lists": (0.995, 0.995)}


def _floors(label):
    for key, v in TRACE_FLOORS.items():
        if key in label:
            return v
    return (0.95, 0.98)


def _check_trace(hg, ho, label):
    """hg: GPU RtHitOut records, ho: oracle records (dict-like with hit/prim/stable/t/p/n/*_sens)."""
    stable = np.asarray(ho["stable"]) == 1
    floor_stable, floor_bare = _floors(label)
    assert stable.mean() >= floor_stable, f"{label}: only {stable.mean():.4f} of the rays are well-conditioned (floor {floor_stable})"
    hit_o, prim_o = np.asarray(ho["hit"]).astype(np.int32), np.asarray(ho["prim"]).astype(np.int32)
    bad = ((hg["hit"] != hit_o) | (hg["prim"] != prim_o)) & stable
    assert not bad.any(), f"{label}: {bad.sum()} stable rays differ in hit/primitive id, first {np.nonzero(bad)[0][:5]}"
    m = stable & (hit_o == 1)
    t_o, p_o, n_o = np.asarray(ho["t"])[m], np.asarray(ho["p"])[m], np.asarray(ho["n"])[m]
    t_tol = REL_TOL * np.maximum(np.abs(t_o), 1e-3) + 4.0 * np.asarray(ho["t_sens"])[m]
    p_tol = REL_TOL * np.maximum(1.0, np.abs(p_o).max(axis=1)) + 4.0 * np.asarray(ho["p_sens"])[m]
    n_tol = REL_TOL * np.maximum(1.0, np.linalg.norm(n_o, axis=1)) + 4.0 * np.asarray(ho["n_sens"])[m]
    dt = np.abs(hg["t"][m] - t_o)
    dp = np.abs(hg["p"][m] - p_o).max(axis=1)
    dn = np.abs(hg["n"][m] - n_o).max(axis=1)
    assert np.all(dt <= t_tol), f"{label}: t off by up to {np.max(dt / t_tol):.2f} x tolerance"
    assert np.all(dp <= p_tol), f"{label}: p off by up to {np.max(dp / p_tol):.2f} x tolerance"
    assert np.all(dn <= n_tol), f"{label}: n off by up to {np.max(dn / n_tol):.2f} x tolerance"
    # the conditioning slack must stay the exception: nearly all rays pass the bare 1e-4 bound (what is left are
    # grazing exits with t ~ 1e-3 where the f32 ROUNDING OF THE SCENE TABLES themselves — sphere centres at |c| ~ 50 carry
    # 2e-6 — moves t by more than 1e-4 relative; point and normal stay within the bare bound)
    bt, bp, bn = dt <= REL_TOL * np.maximum(np.abs(t_o), 1e-3), dp <= REL_TOL * np.maximum(1.0, np.abs(p_o).max(axis=1)), dn <= REL_TOL * 1.0
    bare = bt & bp & bn
    if bare.mean() < floor_bare:
        worst = np.argsort(-(dt / np.maximum(np.abs(t_o), 1e-3)))[:5]
        print(f"{label}: bare pass t {bt.mean():.4f} p {bp.mean():.4f} n {bn.mean():.4f}; worst t: gpu {hg['t'][m][worst]} oracle {t_o[worst]} "
              f"prim {np.asarray(ho['prim'])[m][worst]} rays {np.nonzero(m)[0][worst]}")
    assert bare.mean() >= floor_bare and bp.mean() > 0.999 and bn.mean() > 0.999, f"{label}: only {bare.mean():.4f} of stable hits are within the bare 1e-4 bound (floor {floor_bare})"
    return stable.mean(), bare.mean()


@pytest.mark.parametrize("name", ["book-one", "cornell-box", "book-two", "stress:5000", "checker-lists"])
def test_trace_matches_golden_fixture(rt, device, name):
    g = _golden(name)
    scene = rt.Scene.example(device, name, 1)
    hg = scene.trace(g["rays"], seed=1)
    _check_trace(hg, g, f"golden {name}")
    m = (g["stable"] == 1) & (g["hit"] == 1)
    assert np.all(np.abs(hg["uv"][m] - g["uv"][m]).max(axis=1) <= 2e-3 + 64 * g["n_sens"][m])  # uv (atan2/acos of the normal)


@pytest.mark.parametrize("name,w,h,n_bounce", [("book-one", 96, 54, 20000), ("cornell-box", 64, 64, 20000), ("book-two", 64, 64, 20000),
                                              ("stress:100000", 64, 36, 20000), ("checker-lists", 96, 54, 20000)])
def test_trace_matches_oracle_on_large_ray_sets(rt, orc, device, name, w, h, n_bounce):
    so = orc.Scene(name, 1)
    scene = rt.Scene.example(device, name, 1)
    cam_rays = orc.camera_rays(so.camera(w / h), w, h).astype(np.float32)
    first = so.trace(cam_rays.astype(np.float64), rel_perturb=0.0)
    pts = first["p"][first["hit"] == 1]
    rng = np.random.default_rng(123)
    idx = rng.integers(0, len(pts), n_bounce)
    d = rng.normal(size=(n_bounce, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    o = pts[idx] + 1e-3 * d * np.maximum(1.0, np.abs(pts[idx]).max(axis=1, keepdims=True))
    rays = np.concatenate([cam_rays, np.concatenate([o, d], axis=1).astype(np.float32)])
    ho = so.trace(rays.astype(np.float64), seed=9)
    hg = scene.trace(rays, seed=9)
    stable, bare = _check_trace(hg, ho, name)
    print(f"{name}: {len(rays)} rays, stable {stable:.4f}, within bare 1e-4 {bare:.4f}")


def test_trace_edge_cases(rt, orc, device):
    scene = rt.Scene.example(device, "cornell-box", 1)
    so = orc.Scene("cornell-box", 1)
    assert len(scene.trace(np.zeros((0, 6), dtype=np.float32))) == 0  # empty batch
    rays = np.array([
        [277.5, 277.5, -800, 0, 0, 1],     # axis-aligned (two zero direction components) down the box
        [277.5, 277.5, -800, 0, 0, -1],    # away from everything: miss
        [277.5, 277.5, 277.5, 0, 1, 0],    # straight up into the light
        [277.5, 277.5, 277.5, 0, -1, 0],   # straight down onto the floor
        [277.5, 277.5, 277.5, 1, 0, 0],    # onto the green wall
        [100.0, 400.0, 277.5, 0, 0, 2.5],  # unnormalised direction: t in units of |d| (clear of both cubes)
        [-100, 277.5, 277.5, 1, 0, 0],     # from outside through the back of the red wall (no culling)
    ], dtype=np.float32)
    ho, hg = so.trace(rays.astype(np.float64)), scene.trace(rays)
    assert list(hg["hit"]) == list(ho["hit"]) == [1, 0, 1, 1, 1, 1, 1]
    assert list(hg["prim"]) == list(ho["prim"])
    assert np.allclose(hg["t"], ho["t"], rtol=1e-5)
    assert hg["t"][5] == pytest.approx((555 - 277.5) / 2.5, rel=1e-6)
    assert np.allclose(hg["n"], ho["n"], atol=1e-6)


def _tables(rt, spheres, materials):
    """Flat tables for translation-only spheres: [(centre, radius, material index)]."""
    prims = np.zeros(len(spheres), dtype=rt.PRIM_DTYPE)
    ident = np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], dtype=np.float32)
    for i, (c, r, m) in enumerate(spheres):
        prims[i]["kind"], prims[i]["material"], prims[i]["order"] = rt.PRIM_SPHERE, m, i
        prims[i]["param"][0] = r
        wfl, lfw = ident.copy(), ident.copy()
        wfl[[3, 7, 11]] = c
        lfw[[3, 7, 11]] = [-v for v in c]
        prims[i]["world_from_local"], prims[i]["local_from_world"] = wfl, lfw
    mats = np.zeros(len(materials), dtype=rt.MATERIAL_DTYPE)
    texs = np.zeros(len(materials), dtype=rt.TEXTURE_DTYPE)
    for i, (kind, rgb, fuzz, ior) in enumerate(materials):
        mats[i] = (kind, i, fuzz, ior)
        texs[i]["kind"], texs[i]["a"], texs[i]["b"] = rt.TEX_SOLID, -1, -1
        texs[i]["rgb"][:3] = rgb
    return prims, mats, texs


def test_custom_scene_single_and_tied_primitives(rt, device):
    # one sphere (BVH root is a leaf), then two coincident spheres: the tie goes to the smaller DFS order (optimize.rs:475,484)
    mats = [(rt.MAT_LAMBERTIAN, (0.5, 0.5, 0.5), 0, 0), (rt.MAT_DIFFUSE_LIGHT, (1, 2, 3), 0, 0)]
    prims, m, t = _tables(rt, [((0, 0, 5), 1.0, 0)], mats)
    one = rt.Scene.from_tables(device, prims, m, t)
    h = one.trace(np.array([[0, 0, 0, 0, 0, 1], [0, 3, 0, 0, 0, 1]], dtype=np.float32))
    assert list(h["hit"]) == [1, 0] and h["t"][0] == 4.0 and list(h["n"][0]) == [0, 0, -1] and h["material"][0] == 0
    prims, m, t = _tables(rt, [((0, 0, 5), 1.0, 0), ((0, 0, 5), 1.0, 1), ((0, 0, 9), 1.0, 1)], mats)
    prims["order"] = [7, 3, 9]
    tie = rt.Scene.from_tables(device, prims, m, t)
    h = tie.trace(np.array([[0, 0, 0, 0, 0, 1]], dtype=np.float32))
    assert h["prim"][0] == 3 and h["material"][0] == 1
    # a light seen directly: every sample returns the emission exactly; depth 0 renders black
    cam = rt.make_camera(rt.camera_args([0, 0, 0], [0, 0, 1], [0, 1, 0], math.radians(10), 1.0, 1.0, 0.0))
    s, q, st = tie.render(cam, 8, 8, 16, 5)
    assert np.allclose(s[4, 4] / 16, [1, 2, 3]) and np.allclose(q[4, 4] / 16, [1, 4, 9]) and st.paths == 8 * 8 * 16
    s0, _, _ = tie.render(cam, 8, 8, 4, 0)
    assert np.all(s0 == 0.0)


def test_error_paths_on_device(rt, device):
    mats = [(rt.MAT_LAMBERTIAN, (0.5, 0.5, 0.5), 0, 0)]
    prims, m, t = _tables(rt, [((0, 0, 5), 1.0, 0)], mats)
    with pytest.raises(rt.BackendError) as e:
        rt.Scene.from_tables(device, prims[:0], m, t)  # empty scene (BVH::new returns None, optimize.rs:367-370)
    assert e.value.code == rt.RT_ERR_INVALID
    bad = prims.copy()
    bad["material"] = 5
    with pytest.raises(rt.BackendError):
        rt.Scene.from_tables(device, bad, m, t)
    bad = prims.copy()
    bad["kind"] = 77
    with pytest.raises(rt.BackendError) as e:
        rt.Scene.from_tables(device, bad, m, t)
    assert e.value.code == rt.RT_ERR_UNSUPPORTED
    scene = rt.Scene.from_tables(device, prims, m, t)
    cam = rt.make_camera(rt.camera_args([0, 0, 0], [0, 0, 1], [0, 1, 0], 0.5, 1.0, 1.0, 0.0))
    for kwargs in ({"width": 0}, {"spp": 0}, {"row_stride": 0}, {"row_begin": 99}, {"max_depth": 300}):
        args = dict(width=8, height=8, spp=1, max_depth=5)
        args.update(kwargs)
        with pytest.raises(rt.BackendError):
            scene.render(cam, args["width"], args["height"], args["spp"], args["max_depth"], row_begin=args.get("row_begin", 0),
                         row_stride=args.get("row_stride", 1))


@pytest.mark.parametrize("name,w,h,spp", [("book-one", 160, 90, 4), ("cornell-box", 96, 96, 8), ("book-two", 96, 96, 4), ("stress:20000", 96, 54, 4),
                                          ("checker-lists", 128, 72, 8)])
def test_paired_paths_match_oracle(rt, orc, device, name, w, h, spp):
    """Same Philox keys on both sides: per-pixel sums agree path for path, except where an f32 rounding flips a
    discrete decision (Schlick coin, near-tie) somewhere along a path."""
    scene = rt.Scene.example(device, name, 1)
    s_g, q_g, st = scene.render(scene.flat.camera(w / h), w, h, spp, 50, seed=1)
    s_o, q_o, sto = orc.Scene(name, 1).render(w, h, spp, 50, seed=1, mode=orc.PAIRED)
    diff = np.abs(s_g - s_o).max(axis=2)
    tol = 2e-3 * np.maximum(1.0, np.abs(s_o).max(axis=2))
    frac = np.mean(diff <= tol)
    print(f"{name}: paired pixel agreement {frac:.4f}, rays/path gpu {st.rays / st.paths:.4f} oracle {sto.segments / sto.paths:.4f}")
    assert frac >= 0.985
    assert st.paths == sto.paths == w * h * spp
    assert abs(st.rays - sto.segments) <= 0.004 * sto.segments
    assert abs(s_g.mean() - s_o.mean()) <= 0.01 * s_o.mean()


def test_golden_paired_render(rt, device):
    for name in ["book-one", "cornell-box", "book-two", "stress:5000", "checker-lists"]:
        g = _golden(name)
        w, h, spp, depth = [int(v) for v in g["paired_shape"]]
        scene = rt.Scene.example(device, name, 1)
        s, _, st = scene.render(scene.flat.camera(w / h), w, h, spp, depth, seed=1)
        ok = np.abs(s - g["paired_sum"]).max(axis=2) <= 2e-3 * np.maximum(1.0, np.abs(g["paired_sum"]).max(axis=2))
        assert ok.mean() >= 0.97, f"{name}: {ok.mean():.3f}"
        assert abs(int(st.rays) - int(g["paired_segments"][0])) <= max(4, 0.01 * int(g["paired_segments"][0])), (name, int(st.rays), int(g["paired_segments"][0]))


def _psnr_resolved(mean_a, mean_b):
    ra, rb = np.sqrt(np.clip(mean_a, 0, 1)), np.sqrt(np.clip(mean_b, 0, 1))  # the example drivers' display transform
    mse = np.mean((ra - rb) ** 2)
    return 10 * math.log10(1.0 / max(mse, 1e-30))


@pytest.mark.parametrize("name,w,h,spp_gpu,spp_cpu,depth", [("book-one", 64, 36, 16384, 4096, 50), ("cornell-box", 32, 32, 131072, 32768, 50),
                                                          ("book-two", 24, 24, 262144, 32768, 50), ("stress:20000", 48, 27, 16384, 2048, 16),
                                                          ("checker-lists", 48, 27, 32768, 8192, 50)])
def test_image_statistics_against_high_spp_oracle(rt, orc, device, name, w, h, spp_gpu, spp_cpu, depth):
    """Gate 2.  The oracle renders with its FAITHFUL sampler (rejection sampling, sequential stream — the reference's
    algorithm) and a different seed: the two estimates are independent."""
    scene = rt.Scene.example(device, name, 1)
    s_g, q_g, st = scene.render(scene.flat.camera(w / h), w, h, spp_gpu, depth, seed=11)
    s_o, q_o, sto = orc.Scene(name, 1).render(w, h, spp_cpu, depth, seed=22, mode=orc.FAITHFUL)
    assert np.all(np.isfinite(s_g)) and np.all(s_g >= 0)
    m_g, m_o = s_g.astype(np.float64) / spp_gpu, s_o / spp_cpu
    var_g = np.maximum(q_g.astype(np.float64) / spp_gpu - m_g ** 2, 0) / spp_gpu
    var_o = np.maximum(q_o / spp_cpu - m_o ** 2, 0) / spp_cpu
    # f32 accumulation floor: a pixel that sees only the sky has zero true variance; its f32 sum / sum-of-squares
    # carry ~1e-6 relative rounding, which must not be read as a Monte-Carlo standard error
    floor = 2e-5 * np.maximum(np.maximum(m_g, m_o), 1e-3)
    se = np.sqrt(var_g + var_o + floor ** 2)
    ok = (var_g + var_o) > 0  # reporting only: every channel enters the z statistics (zero-variance ones through the floor)
    z = (m_g - m_o) / se
    exceed = np.mean(np.abs(z) > 3.0)
    psnr = _psnr_resolved(m_g, m_o)
    print(f"{name}: {ok.mean():.3f} of channels have variance; |z|>3 fraction {exceed:.4f} (chance 0.0027), max |z| {np.abs(z).max():.2f}, "
          f"mean z {z.mean():+.3f}, PSNR {psnr:.1f} dB, rays/path gpu {st.rays / st.paths:.3f} oracle {sto.segments / sto.paths:.3f}")
    assert exceed <= 0.01 and np.abs(z).max() < 6.0
    assert abs(z.mean()) < 0.2  # no systematic brightness offset (a biased f32 self-intersection rule shows up here)
    assert psnr >= 40.0
    assert abs(st.rays / st.paths - sto.segments / sto.paths) < 0.02 * sto.segments / sto.paths


def test_row_sharding_matches_full_image(rt, device):
    """Multi-GPU partition (rows y = rank + k*G): every shard sees the Philox keys of the full image."""
    scene = rt.Scene.example(device, "book-one", 1)
    cam = scene.flat.camera(80 / 45)
    full, fq, st = scene.render(cam, 80, 45, 16, 50, seed=3)
    rays = 0
    for rank in range(4):
        part, pq, ps = scene.render(cam, 80, 45, 16, 50, seed=3, row_begin=rank, row_stride=4)
        assert part.shape == full[rank::4].shape
        assert np.allclose(part, full[rank::4], rtol=1e-5, atol=1e-5)  # f32 atomics: summation order only
        assert np.allclose(pq, fq[rank::4], rtol=1e-5, atol=1e-5)
        rays += ps.rays
    assert rays == st.rays


def test_queue_capacity_does_not_change_the_image(rt, device):
    scene = rt.Scene.example(device, "cornell-box", 1)
    cam = scene.flat.camera(1.0)
    a, _, sa = scene.render(cam, 64, 64, 32, 50, seed=5, queue_capacity=1024)
    b, _, sb = scene.render(cam, 64, 64, 32, 50, seed=5, queue_capacity=1 << 20)
    assert sa.rays == sb.rays and sa.iterations > sb.iterations
    assert np.allclose(a, b, rtol=1e-5, atol=1e-5)
    c, _, _ = scene.render(cam, 64, 64, 32, 50, seed=6)
    assert not np.allclose(a, c, rtol=1e-3)


def test_resolve_u8_matches_oracle(rt, orc, device):
    scene = rt.Scene.example(device, "book-one", 1)
    s, _, _ = scene.render(scene.flat.camera(64 / 36), 64, 36, 32, 50)
    got = device.resolve_u8(s, 32)
    want = orc.resolve_u8(s.astype(np.float64), 32)
    assert np.array_equal(got, want)  # bytes are integer work: bit-exact (the kernel resolves in f64 like book-one.rs:95-97)
    rng = np.random.default_rng(5)
    # every byte boundary (mean = (k / 255)^2) and its f32 neighbours, spp values that do not divide evenly
    k = np.arange(0, 300, dtype=np.float64)
    for spp in (1, 3, 100, 1024, 1000003):
        edge = np.float32(((k / 255.0) ** 2) * spp)
        vals = np.concatenate([edge, np.nextafter(edge, np.float32(np.inf)), np.nextafter(edge, np.float32(-np.inf)),
                               rng.uniform(0, 1.2 * spp, 3 * 5000).astype(np.float32)])
        vals = np.resize(vals, (vals.size // 3 // 7) * 7 * 3).reshape(-1, 7, 3)
        assert np.array_equal(device.resolve_u8(vals, spp), orc.resolve_u8(vals.astype(np.float64), spp)), spp
    special = np.array([[[np.nan, 4 * 4.0, 0.0], [4 * 0.25, -4.0, 4 * 1e-9]]], dtype=np.float32)
    assert np.array_equal(device.resolve_u8(special, 4), orc.resolve_u8(special.astype(np.float64), 4))


def test_full_size_config1_properties(rt, orc, device):
    """BASELINE configs[0] at full size (book-one 400x225, 100 spp, depth 50): size-independent properties."""
    w, h, spp = 400, 225, 100
    scene = rt.Scene.example(device, "book-one", 1)
    s, q, st = scene.render(scene.flat.camera(w / h), w, h, spp, 50, seed=1)
    assert st.paths == w * h * spp and np.all(np.isfinite(s)) and np.all(s >= 0) and np.all(q >= 0)
    assert 2.7 < st.rays / st.paths < 2.9  # 2.79 segments per path
    top = s[-1] / spp
    assert np.allclose(top, [0.5, 0.7, 1.0], atol=1e-4)  # top row sees only the sky light
    assert np.all(s.max(axis=2) / spp <= 1.0 + 1e-4)  # radiance can never exceed the only emitter (0.5, 0.7, 1.0)
    assert np.all(q <= s * 1.0 + 1e-3)  # per-path radiance <= 1 -> sum of squares <= sum
    sub, _, _ = orc.Scene("book-one", 1).render(w, h, 4, 50, seed=5, mode=orc.FAITHFUL, row_begin=0, row_stride=9)
    assert abs(s[0::9].mean() / spp - sub.mean() / 4) < 0.01


def test_deinterleave_kernel_matches_index_arithmetic(rt, device):
    """rt_deinterleave_rows (the kernel behind the NCCL gather) against the pure index arithmetic the gloo tests use."""
    import torch

    for height, width, shards in [(12, 9, 2), (10, 5, 3), (2160, 64, 8), (7, 3, 2)]:
        rows = -(-height // shards)
        gathered = torch.rand((shards * rows, width, 3), device="cuda")
        full = torch.empty((height, width, 3), device="cuda")
        device.deinterleave_rows(gathered.data_ptr(), full.data_ptr(), width, height, shards, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        want = gathered.view(shards, rows, width, 3).permute(1, 0, 2, 3).reshape(rows * shards, width, 3)[:height]
        assert torch.equal(full, want)


def test_render_sharded_single_rank_matches_render(rt, device):
    import importlib

    import torch

    rtdist = importlib.import_module("ray-tracer_b200.dist")
    scene = rt.Scene.example(device, "book-one", 1)
    cam = scene.flat.camera(64 / 36)
    full, st = rtdist.render_sharded(scene, cam, 64, 36, 8, 50, 3, 0, 1)
    ref, _, _ = scene.render(cam, 64, 36, 8, 50, seed=3, want_sumsq=False)
    torch.cuda.synchronize()
    assert np.allclose(full.cpu().numpy(), ref, rtol=1e-6, atol=1e-6)


# ---- the specialised extend kernels against the general one (which the trace tests pin to the oracle) -------------

def _render_with(rt, device, name, w, h, spp, depth=50, seed=3, tables=None, cam=None, queue_capacity=0, scene_opts=None, flags=0, ctx_opts=None):
    """One render with explicit options (RtSceneOptions / RtContextOptions keyword overrides, RtRenderParams.flags)."""
    dev = rt.Device(0, options=ctx_opts) if ctx_opts is not None else device
    scene = rt.Scene.from_tables(dev, *tables, options=scene_opts) if tables else rt.Scene.example(dev, name, 1, options=scene_opts)
    cam = cam if cam is not None else scene.flat.camera(w / h)
    s, q, st = scene.render(cam, w, h, spp, depth, seed=seed, queue_capacity=queue_capacity, flags=flags)
    scene.close()
    if ctx_opts is not None:
        dev.close()
    return s, q, st


@pytest.mark.gpu
@pytest.mark.parametrize("name,w,h,spp", [("book-one", 128, 72, 8), ("book-two", 64, 64, 4), ("stress:20000", 96, 54, 4)])
def test_persistent_extend_matches_one_ray_per_thread(rt, device, name, w, h, spp):
    """k_extend_bvh (lane-refilling persistent warps, cp.async ray feed) and k_extend walk the same tree with the same
    arithmetic: identical closest hits, so identical paths — only the order of the f64 accumulations differs."""
    a, qa, sa = _render_with(rt, device, name, w, h, spp, flags=rt.RENDER_EXTEND_PER_THREAD)
    b, qb, sb = _render_with(rt, device, name, w, h, spp, flags=rt.RENDER_EXTEND_PERSISTENT)
    assert sa.rays == sb.rays and sa.paths == sb.paths
    assert np.allclose(a, b, rtol=1e-5, atol=1e-5) and np.allclose(qa, qb, rtol=1e-5, atol=1e-5)
    # a queue smaller than one feed range and one that is not a multiple of 32
    c, _, sc = _render_with(rt, device, name, w, h, 1, queue_capacity=1056, flags=rt.RENDER_EXTEND_PERSISTENT)
    d, _, sd = _render_with(rt, device, name, w, h, 1, queue_capacity=1056, flags=rt.RENDER_EXTEND_PER_THREAD)
    assert sc.rays == sd.rays and np.allclose(c, d, rtol=1e-5, atol=1e-5)


def _rect_tables(rt, rects, materials):
    """Flat tables for rigidly placed rectangles: [(R 3x3 world<-local rotation, centre, width, height, material)]."""
    prims = np.zeros(len(rects), dtype=rt.PRIM_DTYPE)
    for i, (R, c, wd, ht, m) in enumerate(rects):
        R, c = np.asarray(R, dtype=np.float64), np.asarray(c, dtype=np.float64)
        prims[i]["kind"], prims[i]["material"], prims[i]["order"] = rt.PRIM_RECT, m, i
        prims[i]["param"][0], prims[i]["param"][1] = wd, ht
        prims[i]["world_from_local"] = np.concatenate([R, c[:, None]], axis=1).reshape(12).astype(np.float32)
        prims[i]["local_from_world"] = np.concatenate([R.T, (-R.T @ c)[:, None]], axis=1).reshape(12).astype(np.float32)
    mats = np.zeros(len(materials), dtype=rt.MATERIAL_DTYPE)
    texs = np.zeros(len(materials), dtype=rt.TEXTURE_DTYPE)
    for i, (kind, rgb, fuzz, ior) in enumerate(materials):
        mats[i] = (kind, i, fuzz, ior)
        texs[i]["kind"], texs[i]["a"], texs[i]["b"] = rt.TEX_SOLID, -1, -1
        texs[i]["rgb"][:3] = rgb
    return prims, mats, texs


def _rot(axis, deg):
    a = np.asarray(axis, dtype=np.float64)
    a /= np.linalg.norm(a)
    th = math.radians(deg)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    return np.eye(3) + math.sin(th) * K + (1 - math.cos(th)) * (K @ K)


@pytest.mark.gpu
def test_frame_rect_extend_matches_general_records(rt, device):
    """k_extend_rects (rectangles grouped by the frame they are axis-aligned in) against the general 64-byte rectangle
    records: cornell-box, and a random closed room with boxes in three tilted frames, face normals permuted and
    flipped, a coplanar pair (exact tie -> the DFS-earlier rectangle) and one frame used by a single rectangle."""
    a, qa, sa = _render_with(rt, device, "cornell-box", 96, 96, 16, scene_opts={"flags": rt.SCENE_NO_FRAME_RECTS})
    b, qb, sb = _render_with(rt, device, "cornell-box", 96, 96, 16)
    c, qc, sc = _render_with(rt, device, "cornell-box", 96, 96, 16, scene_opts={"flags": rt.SCENE_NO_FRAME_BOXES})
    assert sa.rays == sc.rays and abs(int(sb.rays) - int(sc.rays)) <= 1e-4 * sc.rays
    assert np.allclose(a, c, rtol=2e-4, atol=2e-4) and np.allclose(qa, qc, rtol=5e-4, atol=5e-4)
    # boxed and loose faces compute the same t = (plane - o_k) / d_k; a ray through the very edge of a cube (its two faces tie, or
    # the slab's t_enter <= t_exit and the rectangles' |p - c| <= h round differently) may take the other face: a few paths in a million
    same = np.abs(c - b).max(axis=2) <= 1e-5 * np.maximum(1.0, np.abs(c).max(axis=2))
    assert same.mean() >= 0.998, same.mean()
    # cornell-box: the two Cube::new boxes, and the room's left / right / floor / ceiling (the 555 x 556 back wall overhangs its face
    # and the light floats below the ceiling: both stay loose)
    sc_box = rt.Scene.example(device, "cornell-box", 1)
    assert sc_box.frame_rect_info() == {"frames": 3, "boxes": 3, "loose": 2}
    assert rt.Scene.example(device, "cornell-box", 1, options={"flags": rt.SCENE_NO_FRAME_BOXES}).frame_rect_info() == {"frames": 3, "boxes": 0, "loose": 18}
    # closest hits ray by ray: rays from inside and outside the room, through the open front, along the axes
    rng = np.random.default_rng(2)
    dd = rng.normal(size=(40000, 3))
    dd /= np.linalg.norm(dd, axis=1, keepdims=True)
    oo = rng.uniform(-300, 900, size=(40000, 3))
    axis = np.concatenate([np.eye(3), -np.eye(3)])[rng.integers(0, 6, 2000)]
    rays = np.concatenate([np.concatenate([oo, dd], axis=1), np.concatenate([rng.uniform(0, 555, size=(2000, 3)), axis], axis=1)]).astype(np.float32)
    h_box = sc_box.trace(rays)  # rt_trace runs the scene's own extend form: frame boxes here, the general records there
    h_gen = rt.Scene.example(device, "cornell-box", 1, options={"flags": rt.SCENE_NO_FRAME_RECTS}).trace(rays)
    differ = (h_box["prim"] != h_gen["prim"]) | (h_box["hit"] != h_gen["hit"])
    assert differ.mean() <= 2e-4, differ.mean()  # edge grazers only
    ok = ~differ & (h_gen["hit"] == 1)
    assert np.allclose(h_box["t"][ok], h_gen["t"][ok], rtol=2e-5, atol=1e-4)

    rng = np.random.default_rng(11)
    mats = [(rt.MAT_LAMBERTIAN, (0.7, 0.6, 0.5), 0, 0), (rt.MAT_DIFFUSE_LIGHT, (4, 4, 4), 0, 0), (rt.MAT_METAL, (0.8, 0.8, 0.9), 0.1, 0),
            (rt.MAT_LAMBERTIAN, (0.2, 0.7, 0.3), 0, 0)]
    I = np.eye(3)
    face = {"+x": _rot([0, 1, 0], 90), "-x": _rot([0, 1, 0], -90), "+y": _rot([1, 0, 0], -90), "-y": _rot([1, 0, 0], 90), "+z": I, "-z": _rot([0, 1, 0], 180)}
    rects = []
    L = 10.0
    for name, sgn, ax in [("+x", 1, 0), ("-x", -1, 0), ("+y", 1, 1), ("-y", -1, 1), ("+z", 1, 2)]:  # room, open towards -z (camera side)
        c = np.zeros(3)
        c[ax] = sgn * L / 2
        inward = face[("-" if sgn > 0 else "+") + name[1]]  # normals face the room, so Lambertian walls scatter back into it
        rects.append((inward, c, L, L, 1 if name == "+y" else (3 if name == "-x" else 0)))
    for k in range(3):  # boxes, each in its own tilted frame, faces through all six permutations / signs
        F = _rot(rng.normal(size=3), rng.uniform(10, 80))
        centre = rng.uniform(-2.5, 2.5, size=3)
        half = rng.uniform(0.6, 1.4, size=3)
        for name, sgn, ax in [("+x", 1, 0), ("-x", -1, 0), ("+y", 1, 1), ("-y", -1, 1), ("+z", 1, 2), ("-z", -1, 2)]:
            n_local = np.zeros(3)
            n_local[ax] = sgn * half[ax]
            Rf = face[name]
            ext = np.abs(Rf.T @ (2 * half))  # extents of the face along its own x / y
            rects.append((F @ Rf, centre + F @ n_local, ext[0], ext[1], 2 if k == 0 else 0))
    rects.append((_rot([1, 2, 3], 33), [0.5, -3.0, 1.0], 1.5, 2.5, 3))  # a frame of its own
    rects.append((I, [0, 0, L / 2], L / 2, L / 2, 3))                  # coplanar with the back wall: ties go to the earlier one
    tables = _rect_tables(rt, rects, mats)
    assert len(rects) <= 32
    cam = rt.make_camera(rt.camera_args([0, 0, -22], [0, 0, 0], [0, 1, 0], math.radians(30), 1.0, 1.0, 0.0))
    a, qa, sa = _render_with(rt, device, None, 96, 96, 16, depth=12, tables=tables, cam=cam, scene_opts={"flags": rt.SCENE_NO_FRAME_RECTS})
    b, qb, sb = _render_with(rt, device, None, 96, 96, 16, depth=12, tables=tables, cam=cam)
    c, qc, sc = _render_with(rt, device, None, 96, 96, 16, depth=12, tables=tables, cam=cam, scene_opts={"flags": rt.SCENE_NO_FRAME_BOXES})
    assert abs(int(sb.rays) - int(sc.rays)) <= 1e-4 * sc.rays
    same = np.abs(c - b).max(axis=2) <= 1e-5 * np.maximum(1.0, np.abs(c).max(axis=2))
    assert same.mean() >= 0.995, same.mean()
    info = rt.Scene.from_tables(device, *tables).frame_rect_info()
    assert info["boxes"] >= 4 and info["boxes"] * 3 + info["loose"] <= len(rects)  # the room and the three tilted boxes
    assert sa.rays > 96 * 96 * 16 * 2  # paths do bounce
    # f32 rounding differs between the two formulations: a path may flip at a grazing edge, the image may not
    assert abs(int(sa.rays) - int(sb.rays)) <= 2e-3 * sa.rays
    close = np.abs(a - b).max(axis=2) <= 2e-3 * np.maximum(1.0, np.abs(a).max(axis=2))
    assert close.mean() >= 0.985, close.mean()
    assert abs(a.mean() - b.mean()) <= 5e-3 * a.mean()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["book-one", "cornell-box", "book-two", "checker-lists"])
def test_color_rays_matches_oracle_color(rt, orc, device, name):
    """rt_color_rays — the reference's per-ray entry point render::color (src/render.rs:5) batched through the wavefront
     — against the oracle's color() on the same rays with the same Philox keys (ray, sample, segment)."""
    so = orc.Scene(name, 1)
    sg = rt.Scene.example(device, name, 1)
    cam_rays = orc.camera_rays(so.camera(16 / 9), 48, 27)
    hits = so.trace(cam_rays, rel_perturb=0.0)
    pts = hits["p"][hits["hit"] == 1]
    rng = np.random.default_rng(3)
    d = rng.normal(size=(1500, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    o = pts[rng.integers(0, len(pts), len(d))]
    o = o + 1e-3 * d * np.maximum(1.0, np.abs(o).max(axis=1, keepdims=True))
    bounce = np.concatenate([o, d * rng.uniform(0.5, 3.0, size=(len(d), 1))], axis=1)  # color() takes un-normalised directions
    rays = np.concatenate([cam_rays, bounce]).astype(np.float32)
    spp = 4
    s_g, q_g, st = sg.color_rays(rays, spp, 50, seed=9, want_sumsq=True)
    s_o, seg = so.color_rays(rays.astype(np.float64), spp, 50, seed=9)
    assert st.paths == len(rays) * spp
    assert abs(int(st.rays) - int(seg.sum())) <= 0.02 * seg.sum()  # ~36 k segments: one flipped decision on a fog path moves dozens
    ok = np.abs(s_g - s_o).max(axis=1) <= 2e-3 * np.maximum(1.0, np.abs(s_o).max(axis=1))
    print(f"{name}: color_rays agreement {ok.mean():.4f} over {len(rays)} rays x {spp}")
    assert ok.mean() >= 0.985
    assert np.all(q_g >= 0.0)
    zero, _, _ = sg.color_rays(rays[:64], 2, 0)
    assert np.all(zero == 0.0)
    with pytest.raises(rt.BackendError):
        sg.color_rays(np.zeros((0, 6), dtype=np.float32), 1, 5)


@pytest.mark.gpu
def test_parallel_bvh_build_is_thread_count_independent(rt, device):
    """The host SAH builder hands subtrees to worker threads (csrc/bvh_build.cpp); the emitted tree must not depend on how many."""
    import time
    out = {}
    for threads in ["1", "5", ""]:
        t0 = time.time()
        scene = rt.Scene.example(device, "stress:200000", 1, options={"bvh_builder": rt.BVH_HOST_SAH, "bvh_threads": int(threads or 0)})  # this size defaults to the device builder
        dt = time.time() - t0
        s, _, st = scene.render(scene.flat.camera(16 / 9), 96, 54, 2, 16, seed=2)
        out[threads] = (scene.bvh_info(), int(st.rays), s)
        print(f"stress:200000 scene build+upload with bvh_threads={threads or 'default'}: {dt:.3f} s, bvh {scene.bvh_info()}")
        scene.close()
    for threads in ["5", ""]:
        assert out[threads][0] == out["1"][0] and out[threads][1] == out["1"][1]
        assert np.allclose(out[threads][2], out["1"][2], rtol=1e-5, atol=1e-5)


@pytest.mark.gpu
def test_full_size_config2_properties(rt, orc, device):
    """BASELINE configs[1] at full size (cornell-box 600x600, 1024 spp, depth 50 — the bench workload): size-independent
    properties.  Radiance is bounded by emission x throughput <= 15; the image mean is resolution independent, so it must agree
    with a coarse FAITHFUL oracle render of the same frame; two interleaved row shards reassemble the full image exactly."""
    w = h = 600
    spp = 1024
    scene = rt.Scene.example(device, "cornell-box", 1)
    cam = scene.flat.camera(1.0)
    s, _, st = scene.render(cam, w, h, spp, 50, seed=1, want_sumsq=False)
    assert st.paths == w * h * spp and np.all(np.isfinite(s)) and np.all(s >= 0)
    assert 6.2 < st.rays / st.paths < 6.6
    assert s.max() / spp <= 15.0 + 1e-3
    assert s.max() / spp > 14.9  # pixels that lie on the light (emission 15, seen from below) return it on every sample
    coarse, _, _ = orc.Scene("cornell-box", 1).render(40, 40, 256, 50, seed=7, mode=orc.FAITHFUL)
    assert abs(s.mean() / spp - coarse.mean() / 256) < 0.02 * coarse.mean() / 256
    parts = [scene.render(cam, w, h, 8, 50, seed=4, row_begin=r, row_stride=2, want_sumsq=False)[0] for r in range(2)]
    full8, _, _ = scene.render(cam, w, h, 8, 50, seed=4, want_sumsq=False)
    assert np.allclose(full8[0::2], parts[0], rtol=1e-5, atol=1e-5) and np.allclose(full8[1::2], parts[1], rtol=1e-5, atol=1e-5)


@pytest.mark.gpu
def test_ingested_earth_texture_renders_like_the_oracle(rt, orc, device):
    """a decoded image installed as book-two's earth texture (examples/main.rs:266-281) on both sides — the GPU
    bakes it 1:1 into its texel table, the oracle evaluates the closure — gives the same paths; and it is not the synthetic one."""
    rng = np.random.default_rng(21)
    earth = np.repeat(np.repeat(rng.integers(0, 256, size=(8, 16, 3), dtype=np.uint8), 6, axis=0), 5, axis=1)  # 48 x 80 blocks
    png = rt.encode_image(earth, rt.IMAGE_PNG_RGBA8)
    decoded = rt.decode_image(png)
    try:  # the format the reference actually opens (examples/main.rs:266 earthmap.jpg): the same picture through the JPEG decoder
        import io
        from PIL import Image
        buf = io.BytesIO()
        Image.fromarray(decoded).save(buf, format="JPEG", quality=90, subsampling=2)
        decoded = rt.decode_image(buf.getvalue())
        assert np.array_equal(decoded, np.asarray(Image.open(io.BytesIO(buf.getvalue())).convert("RGB")))
    except ImportError:
        pass
    w = h = 72
    base, _, _ = rt.Scene.example(device, "book-two", 1).render(rt.FlatScene("book-two", 1).camera(1.0), w, h, 4, 50, seed=1)
    rt.set_earth_image(decoded)
    orc.set_earth_image(decoded)
    try:
        scene = rt.Scene.example(device, "book-two", 1)
        s_g, _, st = scene.render(scene.flat.camera(1.0), w, h, 4, 50, seed=1)
        s_o, _, sto = orc.Scene("book-two", 1).render(w, h, 4, 50, seed=1, mode=orc.PAIRED)
    finally:
        rt.set_earth_image(None)
        orc.set_earth_image(None)
    ok = np.abs(s_g - s_o).max(axis=2) <= 2e-3 * np.maximum(1.0, np.abs(s_o).max(axis=2))
    assert ok.mean() >= 0.985, ok.mean()
    assert abs(int(st.rays) - int(sto.segments)) <= 0.01 * sto.segments
    assert not np.allclose(s_g, base, rtol=1e-3, atol=1e-3)  # the earth sphere is in view: the texture matters


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["stress:20000", "book-two", "stress:300000"])
def test_device_lbvh_build_gives_the_same_hits_as_the_host_sah_tree(rt, orc, device, name):
    """the BVH built on the GPU (csrc/lbvh.cu: Morton sort, Karras radix tree, bottom-up fit) against the host
    SAH tree.  Closest hits do not depend on topology and both trees run the same primitive arithmetic, so rt_trace must return
    identical records, and renders with the same seed identical paths."""
    import time
    so = orc.Scene(name if name != "stress:300000" else "stress:20000", 1)
    rng = np.random.default_rng(4)
    cam_rays = orc.camera_rays(so.camera(16 / 9), 96, 54)
    d = rng.normal(size=(20000, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    o = rng.uniform(-40, 40, size=(20000, 3)) if name.startswith("stress") else rng.uniform(0, 555, size=(20000, 3))
    rays = np.concatenate([cam_rays, np.concatenate([o, d], axis=1)]).astype(np.float32)
    res = {}
    for builder in ["host", "device"]:
        t0 = time.time()
        scene = rt.Scene.example(device, name, 1, options={"bvh_builder": rt.BVH_HOST_SAH if builder == "host" else rt.BVH_DEVICE_LBVH})
        dt = time.time() - t0
        hits = scene.trace(rays, seed=1)
        s, _, st = scene.render(scene.flat.camera(16 / 9), 96, 54, 2, 16, seed=2)
        res[builder] = (hits, s, int(st.rays), scene.bvh_info())
        print(f"{name}: {builder} build+upload {dt:.3f} s, bvh {scene.bvh_info()}")
        scene.close()
    hh, hd = res["host"][0], res["device"][0]
    assert res["host"][3] != res["device"][3]  # really two different trees
    for field in ["hit", "prim", "material"]:
        assert np.array_equal(hh[field], hd[field]), field
    assert np.array_equal(hh["t"], hd["t"]) and np.array_equal(hh["p"], hd["p"]) and np.array_equal(hh["n"], hd["n"])
    assert res["host"][2] == res["device"][2]
    assert np.allclose(res["host"][1], res["device"][1], rtol=1e-5, atol=1e-5)


@pytest.mark.gpu
def test_wavefront_lanes_do_not_change_the_image(rt):
    """rt_render_device runs several interleaved wavefronts (lanes) per render, each on its own streams and host thread
    (capi.cu render_lanes).  Keys are functions of the global pixel: 1, 2 and 3 lanes give the same image, ray count and sums of
    squares; ragged row counts and a caller-side row shard on top of the lane split included."""
    out = {}
    for lanes in ["1", "2", "3"]:
        dev = rt.Device(0, options={"lanes": int(lanes)})
        scene = rt.Scene.example(dev, "cornell-box", 1)
        cam = scene.flat.camera(301 / 203)
        s, q, st = scene.render(cam, 301, 203, 40, 50, seed=3)                                   # 2.4 M paths: enough for 2 lanes
        t, _, st2 = scene.render(cam, 301, 203, 64, 50, seed=4, row_begin=1, row_stride=2, want_sumsq=False)  # shard of a 2-GPU split
        out[lanes] = (s, q, int(st.rays), t, int(st2.rays), int(st.launches))
        scene.close()
        dev.close()
    assert out["2"][5] != out["1"][5]  # the lanes were really used
    for lanes in ["2", "3"]:
        assert out[lanes][2] == out["1"][2] and out[lanes][4] == out["1"][4]
        assert np.allclose(out[lanes][0], out["1"][0], rtol=1e-5, atol=1e-5) and np.allclose(out[lanes][1], out["1"][1], rtol=1e-5, atol=1e-5)
        assert np.allclose(out[lanes][3], out["1"][3], rtol=1e-5, atol=1e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("name,w,h,spp,cap", [("cornell-box", 97, 61, 40, 0), ("book-one", 128, 72, 7, 5000), ("book-two", 64, 64, 33, 0),
                                              ("cornell-box", 3, 2, 70001, 0), ("book-one", 2, 1, 200003, 30000)])
def test_path_order_does_not_change_the_image(rt, device, name, w, h, spp, cap):
    """k_generate issues paths pixel-major (the samples of a pixel next to each other; k_accumulate then folds runs of one pixel in
    the warp before touching the framebuffer) or sample-major (RT_RENDER_SAMPLE_MAJOR).  Philox keys are (global pixel, sample) either
    way: same paths, same ray count, same sums and sums of squares — spp not a multiple of 32, a queue far smaller than a
    pixel's worth of samples × pixels, and more samples per pixel than the generate grid has threads included."""
    a, qa, sa = _render_with(rt, device, name, w, h, spp, queue_capacity=cap, flags=rt.RENDER_SAMPLE_MAJOR, ctx_opts={"lanes": 1})
    b, qb, sb = _render_with(rt, device, name, w, h, spp, queue_capacity=cap, ctx_opts={"lanes": 1})
    assert sa.rays == sb.rays and sa.paths == sb.paths == w * h * spp
    assert np.allclose(a, b, rtol=1e-5, atol=1e-5) and np.allclose(qa, qb, rtol=1e-5, atol=1e-5)


# ---- round 2: surface-origin rays, concurrent contexts, several GPUs behind the C ABI ---------------------------------------------

def _n_gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.gpu
@pytest.mark.parametrize("name,w,h", [("book-one", 96, 54), ("cornell-box", 64, 64), ("book-two", 64, 64), ("stress:20000", 64, 36), ("checker-lists", 96, 54)])
def test_trace_from_surface_origins_matches_oracle(rt, orc, device, name, w, h):
    """The f32 self-intersection rule, ray by ray.  The oracle traces scattered rays whose f64 origin IS the
    previous hit point and rejects the hit at the origin with `t > 1e-6` (src/geometry.rs:57-66,159); the GPU gets the same rays
    rounded to f32 — an origin up to 6e-5 off the r = 1000 ground — plus the primitive each ray starts on (rt_trace_from), which
    is exactly the state a scattered ray carries through the wavefront.  Directions cover both hemispheres: outward rays must not
    re-hit their own surface, inward rays on a sphere must come out at its far side (the same primitive)."""
    so = orc.Scene(name, 1)
    scene = rt.Scene.example(device, name, 1)
    cam_rays = orc.camera_rays(so.camera(w / h), w, h)
    first = so.trace(cam_rays, rel_perturb=0.0)
    prims = so.prims()
    m = (first["hit"] == 1) & (prims["kind"][np.maximum(first["prim"], 0)] <= 1)  # starts on spheres / rectangles (media scatter inside a volume)
    p0, n0, id0 = first["p"][m], first["n"][m], first["prim"][m]
    rng = np.random.default_rng(77)
    reps = max(1, 12000 // len(p0))
    p0, n0, id0 = np.repeat(p0, reps, axis=0), np.repeat(n0, reps, axis=0), np.repeat(id0, reps)
    d = rng.normal(size=p0.shape)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays64 = np.concatenate([p0, d], axis=1)
    ho = so.trace(rays64, seed=5, rel_perturb=0.0)  # (perturbing an origin that lies ON a surface would move it through the surface)
    hg = scene.trace(rays64.astype(np.float32), seed=5, self_order=id0.astype(np.uint32))
    cosn = np.abs(np.einsum("ij,ij->i", d, n0 / np.linalg.norm(n0, axis=1, keepdims=True)))
    # well-conditioned set: no near-tie among the candidates, not grazing the surface the ray starts on, and not grazing the target
    # (an f32 origin error e moves t by e / cos; the bit-exact claim is about rays whose answer does not hinge on that)
    n_hit = ho["n"] / np.maximum(np.linalg.norm(ho["n"], axis=1, keepdims=True), 1e-30)
    cos_hit = np.where(ho["hit"] == 1, np.abs(np.einsum("ij,ij->i", d, n_hit)), 1.0)
    ok = (ho["stable"] == 1) & (cosn > 0.05) & (cos_hit > 0.02)
    assert ok.mean() > 0.8, ok.mean()
    bad = ((hg["hit"] != ho["hit"]) | (hg["prim"] != ho["prim"])) & ok
    inward = np.einsum("ij,ij->i", d, n0) < 0
    is_sphere = prims["kind"][id0] == 0
    print(f"{name}: {len(rays64)} surface rays, well-conditioned {ok.mean():.4f}, mismatches {bad.sum()}, self re-hits expected "
          f"{np.mean(ho['prim'][ok] == id0[ok]):.3f} got {np.mean(hg['prim'][ok] == id0[ok]):.3f}")
    assert not bad.any(), f"{name}: {bad.sum()} well-conditioned surface rays differ, first {np.nonzero(bad)[0][:5]}"
    # inward rays on a sphere leave through the same sphere (far root); nothing ever re-hits a rectangle it starts on
    sel = ok & inward & is_sphere & (ho["prim"] == id0)
    assert sel.sum() > 50 or not is_sphere.any()
    assert np.all(hg["prim"][ok & ~is_sphere] != id0[ok & ~is_sphere])
    hit = ok & (ho["hit"] == 1)
    eo = np.abs(rays64[:, :3] - rays64[:, :3].astype(np.float32)).max(axis=1)  # what rounding the origin to f32 moved ...
    eo = np.maximum(eo, 2.0 ** -23 * np.abs(rays64[:, :3]).max(axis=1))           # ... and one ulp of the scene tables' own f32 coordinates
    t_tol = REL_TOL * np.maximum(np.abs(ho["t"]), 1e-3) + 4.0 * eo / np.maximum(cos_hit, 0.02)
    p_tol = REL_TOL * np.maximum(1.0, np.abs(ho["p"]).max(axis=1))
    dt, dp, dn = np.abs(hg["t"] - ho["t"]), np.abs(hg["p"] - ho["p"]).max(axis=1), np.abs(hg["n"] - ho["n"]).max(axis=1)
    assert np.all(dt[hit] <= t_tol[hit]), f"t off by up to {np.max(dt[hit] / t_tol[hit]):.2f} x tolerance"
    assert np.mean(dp[hit] <= p_tol[hit]) > 0.999 and np.mean(dn[hit] <= 2e-4) > 0.995, (np.mean(dp[hit] <= p_tol[hit]), np.mean(dn[hit] <= 2e-4))


@pytest.mark.gpu
def test_two_contexts_render_different_scenes_concurrently(rt):
    """Launch configuration is per device and per scene (csrc/kernels.cu DeviceState / SceneLaunch), not per process: two contexts —
    on two GPUs when the box has them, else both on GPU 0 — render DIFFERENT scenes (different shared-memory footprints: book-one
    stages 98 KB, cornell-box < 2 KB) at the same time from two host threads and get the images they get alone."""
    import threading
    ids = [0, 1] if _n_gpus() >= 2 else [0, 0]
    devs = [rt.Device(i) for i in ids]
    names = ["book-one", "cornell-box"]
    scenes = [rt.Scene.example(dv, nm, 1) for dv, nm in zip(devs, names)]
    cams = [sc.flat.camera(160 / 90) for sc in scenes]
    alone = [sc.render(cm, 160, 90, 64, 50, seed=9) for sc, cm in zip(scenes, cams)]
    together = [None, None]

    def work(i):
        for _ in range(3):
            together[i] = scenes[i].render(cams[i], 160, 90, 64, 50, seed=9)

    th = [threading.Thread(target=work, args=(i,)) for i in range(2)]
    [t.start() for t in th]
    [t.join() for t in th]
    for i in range(2):
        assert together[i] is not None and together[i][2].rays == alone[i][2].rays
        assert np.allclose(together[i][0], alone[i][0], rtol=1e-5, atol=1e-5)
    # and a scene created on one context cannot be rendered through a context of another device
    for sc in scenes:
        sc.close()
    for dv in devs:
        dv.close()


@pytest.mark.gpu
def test_render_multi_matches_single_device_render(rt, device):
    """rt_init_multi / rt_render_multi (one process, one host thread per device, rows y % n == device, NCCL gather to device 0):
    the image equals the single-context render for every device count the box offers (Philox keys are functions of the global pixel)."""
    flat = rt.FlatScene("book-one", 1)
    w, h, spp = 160, 91, 32  # 91 rows: ragged shards
    cam = flat.camera(w / h)
    ref, refq, rst = rt.Scene(device, flat=flat).render(cam, w, h, spp, 50, seed=4)
    counts = [n for n in (1, 2, 3, 4, 8) if n <= _n_gpus()]
    for n in counts:
        multi = rt.MultiDevice(list(range(n)))
        ms = rt.MultiScene(multi, flat=flat)
        s, q, st = ms.render(cam, w, h, spp, 50, seed=4)
        assert st.paths == rst.paths and st.rays == rst.rays, n
        assert np.allclose(s, ref, rtol=1e-5, atol=1e-5) and np.allclose(q, refq, rtol=1e-5, atol=1e-5), n
        s2, _, _ = ms.render(cam, w, h, spp, 50, seed=4, want_sumsq=False)  # buffers are reused across renders
        assert np.allclose(s2, ref, rtol=1e-5, atol=1e-5)
        tiny, _, _ = ms.render(cam, 5, 1, 4, 50, seed=4, want_sumsq=False)  # fewer rows than devices
        tref, _, _ = rt.Scene(device, flat=flat).render(cam, 5, 1, 4, 50, seed=4, want_sumsq=False)
        assert np.allclose(tiny, tref, rtol=1e-5, atol=1e-5)
        buf = np.zeros((h, w, 3), dtype=np.float32)
        rc = rt.lib.rt_render_multi(ms._h, cam, rt.RtRenderParams(w, h, spp, 50, 4, 1, 2, 0, 0), buf.ctypes.data, None, None)  # caller-side sharding is refused
        assert rc == rt.RT_ERR_INVALID
        ms.close()
        multi.close()
    print(f"rt_render_multi checked on {counts} device(s)")


@pytest.mark.gpu
def test_checker_lists_scene_in_both_device_forms(rt, orc, device):
    """The parity scene for what no example uses: Vec<Arc<dyn Bound>> as the world root, as an inner node
    and as a Sprite's geometry (src/geometry.rs:76-116, src/optimize.rs:256-334), CheckerTexture nested once on a rectangle, a
    sphere, under a Metal and as an emitter (src/material.rs:217-245), and a Cube-bounded ConstantMedium (src/volume.rs:40-43).
    Its 21 primitives run as the warp-uniform DFS list by default; RT_SCENE_NO_SMALL_LIST gives it a tree like any larger scene.
    Both forms must answer like the oracle — ids, t / p / n, uv (the checker reads uv) — and like each other."""
    so = orc.Scene("checker-lists", 1)
    cam_rays = orc.camera_rays(so.camera(16 / 9), 160, 90)
    rng = np.random.default_rng(8)
    d = rng.normal(size=(15000, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    o = np.stack([rng.uniform(-14, 14, len(d)), rng.uniform(0.2, 12, len(d)), rng.uniform(-12, 12, len(d))], axis=1)
    rays = np.concatenate([cam_rays, np.concatenate([o, d], axis=1)]).astype(np.float32)
    ho = so.trace(rays.astype(np.float64), seed=3)
    assert set(np.unique(ho["prim"][ho["hit"] == 1])) >= set(range(0, 21)) - {19}, "every primitive but the sky must be hit by some ray"
    forms = {}
    for label, opts in [("list", None), ("tree", {"flags": rt.SCENE_NO_SMALL_LIST})]:
        scene = rt.Scene.example(device, "checker-lists", 1, options=opts)
        info = scene.bvh_info()
        assert (info["nodes"] == 0) == (label == "list")
        hg = scene.trace(rays, seed=3)
        stable, bare = _check_trace(hg, ho, f"checker-lists {label}")
        m = (ho["stable"] == 1) & (ho["hit"] == 1)
        uv_o = np.stack([ho["u"], ho["v"]], axis=1)
        assert np.all(np.abs(hg["uv"][m] - uv_o[m]).max(axis=1) <= 2e-3 + 64 * ho["n_sens"][m])
        s, q, st = scene.render(scene.flat.camera(16 / 9), 128, 72, 16, 50, seed=2)
        forms[label] = (hg, s, int(st.rays))
        print(f"checker-lists as {label}: bvh {info}, stable {stable:.4f}, bare {bare:.4f}, rays/path {st.rays / st.paths:.3f}")
    a, b = forms["list"], forms["tree"]
    for f in ["hit", "prim", "material"]:
        assert np.array_equal(a[0][f], b[0][f]), f
    assert np.allclose(a[0]["t"], b[0]["t"], rtol=2e-6, atol=0)  # (the tree form tests the two huge spheres through the always-list's own formula)
    assert a[2] == b[2] and np.allclose(a[1], b[1], rtol=1e-5, atol=1e-5)
    # the checker really is in the image: the floor alternates between its three colours
    s_o, _, _ = so.render(128, 72, 16, 50, seed=2, mode=orc.PAIRED)
    ok = np.abs(a[1] - s_o).max(axis=2) <= 2e-3 * np.maximum(1.0, np.abs(s_o).max(axis=2))
    assert ok.mean() >= 0.985, ok.mean()


@pytest.mark.gpu
@pytest.mark.parametrize("builder", ["host", "device"])
def test_trace_config4_one_million_spheres_matches_oracle(rt, orc, device, builder):
    """BASELINE configs[3] at its real size — 1 000 000 spheres under one BVH — against the oracle (the reference's random-axis
    median-split tree, unpruned recursion), ray by ray, for BOTH device trees: the host binned-SAH build and the LBVH built on the GPU."""
    so = orc.Scene("stress:1000000", 1)
    scene = rt.Scene.example(device, "stress:1000000", 1, options={"bvh_builder": rt.BVH_HOST_SAH if builder == "host" else rt.BVH_DEVICE_LBVH})
    info = scene.bvh_info()
    assert info["nodes"] > 1_000_000
    cam_rays = orc.camera_rays(so.camera(16 / 9), 64, 36).astype(np.float32)
    first = so.trace(cam_rays.astype(np.float64), rel_perturb=0.0)
    pts = first["p"][first["hit"] == 1]
    rng = np.random.default_rng(31)
    n_b = 6000
    d = rng.normal(size=(n_b, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    o = pts[rng.integers(0, len(pts), n_b)] + 1e-3 * d * 50.0
    inside = np.concatenate([rng.uniform(-50, 50, size=(3000, 3)), rng.normal(size=(3000, 3))], axis=1)  # rays from anywhere in the cloud
    inside[:, 3:] /= np.linalg.norm(inside[:, 3:], axis=1, keepdims=True)
    rays = np.concatenate([cam_rays, np.concatenate([o, d], axis=1).astype(np.float32), inside.astype(np.float32)])
    ho = so.trace(rays.astype(np.float64), seed=9)
    hg = scene.trace(rays, seed=9)
    stable, bare = _check_trace(hg, ho, f"stress:1000000 ({builder} tree)")
    print(f"stress:1000000 {builder}: bvh {info}, {len(rays)} rays, stable {stable:.4f}, within bare 1e-4 {bare:.4f}, hit fraction {hg['hit'].mean():.3f}")


def _gate2(rt, orc, device, name, w, h, spp_gpu, spp_cpu, depth, block):
    scene = rt.Scene.example(device, name, 1)
    s_g, q_g, st = scene.render(scene.flat.camera(w / h), w, h, spp_gpu, depth, seed=11)
    s_o, q_o, sto = orc.Scene(name, 1).render(w, h, spp_cpu, depth, seed=22, mode=orc.FAITHFUL)
    assert np.all(np.isfinite(s_g)) and np.all(s_g >= 0)
    m_g, m_o = s_g.astype(np.float64) / spp_gpu, s_o / spp_cpu
    var_g = np.maximum(q_g.astype(np.float64) / spp_gpu - m_g ** 2, 0) / spp_gpu
    var_o = np.maximum(q_o / spp_cpu - m_o ** 2, 0) / spp_cpu
    floor = 2e-5 * np.maximum(np.maximum(m_g, m_o), 1e-3)
    z = (m_g - m_o) / np.sqrt(var_g + var_o + floor ** 2)
    exceed = float(np.mean(np.abs(z) > 3.0))
    # block means: the same statistics on (block x block)-pixel averages — 1/block of the per-pixel standard error, so a bias of a
    # fraction of a sigma per pixel, invisible above, shows here
    hb, wb = (h // block) * block, (w // block) * block
    def blk(a):
        return a[:hb, :wb].reshape(hb // block, block, wb // block, block, 3).mean(axis=(1, 3))
    bm_g, bm_o = blk(m_g), blk(m_o)
    bvar = (blk(var_g) + blk(var_o)) / (block * block)
    zb = (bm_g - bm_o) / np.sqrt(bvar + blk(floor) ** 2)
    exceed_b = float(np.mean(np.abs(zb) > 3.0))
    psnr_px = _psnr_resolved(m_g, m_o)
    psnr_blk = _psnr_resolved(bm_g, bm_o)
    # what Monte-Carlo noise alone does to the per-pixel PSNR of two independent renders (delta method on the sqrt display transform)
    mid = np.clip(0.5 * (m_g + m_o), 1e-6, 1.0)
    noise_mse = np.mean(np.where((m_g < 1.0) | (m_o < 1.0), (var_g + var_o) / (4.0 * mid), 0.0))
    psnr_noise = 10 * math.log10(1.0 / max(noise_mse, 1e-30))
    print(f"{name} {w}x{h}: gpu {spp_gpu} spp vs FAITHFUL oracle {spp_cpu} spp ({sto.seconds:.1f} s on {sto.threads} threads): |z|>3 {exceed:.4f} per pixel, "
          f"{exceed_b:.4f} per {block}x{block} block (chance 0.0027), max |z| {np.abs(z).max():.2f}, mean z {z.mean():+.3f}; PSNR {psnr_px:.1f} dB per pixel "
          f"(noise alone: {psnr_noise:.1f}), {psnr_blk:.1f} dB on block means; rays/path gpu {st.rays / st.paths:.4f} oracle {sto.segments / sto.paths:.4f}")
    assert exceed <= 0.005 and np.abs(z).max() < 6.0
    assert exceed_b <= 0.01 and abs(z.mean()) < 0.05
    assert psnr_px >= psnr_noise - 1.0   # the per-pixel difference is what the two renders' own standard errors predict
    assert psnr_blk >= 40.0
    assert abs(st.rays / st.paths - sto.segments / sto.paths) < 0.005 * sto.segments / sto.paths


@pytest.mark.gpu
def test_gate2_config1_full_size(rt, orc, device):
    """Gate 2 at BASELINE configs[0]'s own size: book-one 400x225, depth 50, the GPU at 1024 spp (>= 1024 as the north star asks)
    against an independent FAITHFUL oracle render (rejection sampling, sequential stream: the reference's algorithm) at 1024 spp."""
    _gate2(rt, orc, device, "book-one", 400, 225, 1024, 1024, 50, 5)


@pytest.mark.gpu
def test_gate2_config2_full_size(rt, orc, device):
    """Gate 2 at BASELINE configs[1]'s own size: cornell-box 600x600, 1024 spp, depth 50 on the GPU against the FAITHFUL oracle at
    the SAME 1024 spp (368 M CPU paths, ~50 s on the box's cores).  Equal sample counts matter: per-path radiance is heavy-tailed
    (a 130x105 light of emission 15 found through 6.4 bounces), so a 256-spp estimate is skewed low with an under-estimated
    variance — oracle against oracle at 1024 / 256 spp already shows |z| > 3 in 2.0 % of the channels and a mean z of +0.24, at
    1024 / 1024 it shows 0.15 % and 0.00.  Two 1024-spp renders differ by ~27 dB of Monte-Carlo noise per pixel, so PSNR >= 40 dB
    is asserted on 10x10 block means and the per-pixel PSNR against its noise prediction."""
    _gate2(rt, orc, device, "cornell-box", 600, 600, 1024, 1024, 50, 10)


@pytest.mark.gpu
def test_cpp_example_driver_writes_the_reference_ppm(rt, device, tmp_path):
    """ray-tracer_b200/examples/book_one.cpp — examples/book-one.rs through the C++ host mirror: scene built with the mirrored
    API, ONE gpu::render_image call (host/gpu.hpp), gpu::resolve_u8, PPM P3 as book-one.rs:28-30,90-100 prints it.  The file must
    equal, byte for byte, what the ctypes path produces for the same scene seed and render seed."""
    import subprocess
    pkg = os.path.join(os.path.dirname(GOLDEN), "..", "ray-tracer_b200")
    exe = os.path.join(pkg, "examples", "book_one")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-s", "-C", pkg, "examples/book_one"])
    w, h, spp, depth, seed = 160, 90, 8, 50, 3
    out = tmp_path / "book_one.ppm"
    res = subprocess.run([exe, str(w), str(h), str(spp), str(depth), str(seed), str(out)], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    assert "paths" in res.stderr
    scene = rt.Scene.example(device, "book-one", seed)
    s, _, st = scene.render(scene.flat.camera(w / h), w, h, spp, depth, seed=seed, want_sumsq=False)
    means = (s.astype(np.float64) / spp).astype(np.float32)  # render_image divides in f64 (pixel /= spp, book-one.rs:76); resolve_u8 takes f32
    want = rt.encode_image(device.resolve_u8(means, 1), rt.IMAGE_PPM_P3)
    got = out.read_bytes()
    assert got.startswith(b"P3\n160 90\n255\n")
    assert got == want


@pytest.mark.gpu
def test_every_kernel_form_under_the_checked_library(rt):
    """compute-sanitizer is closed on this GPU pool, so the kernels carry their own bounds / invariant asserts (RT_ASSERT in csrc/,
    compiled in by `make librt_b200_check.so`): every queue slot, BVH node, primitive, general record, rectangle, material, texture,
    traversal-stack and framebuffer index is checked where it is formed, and a violation traps the kernel.  scripts/sanitize_target.py
    drives every kernel form — all extend variants, both shade variants, lanes, tiny queues, surface-origin traces, the LBVH build —
    through that library in a subprocess; it must finish cleanly."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib = os.path.join(root, "ray-tracer_b200", "librt_b200_check.so")
    if not os.path.exists(lib):
        subprocess.check_call(["make", "-s", "-C", os.path.dirname(lib), "librt_b200_check.so"])
    env = dict(os.environ, RT_B200_LIB=lib)
    res = subprocess.run([sys.executable, os.path.join(root, "scripts", "sanitize_target.py")], env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert "ok (36, 64, 3)" in res.stdout or "ok" in res.stdout.splitlines()[-1]
    assert "Assertion" not in res.stderr


@pytest.mark.gpu
def test_box_primitives_match_six_rectangle_leaves(rt, orc, device):
    """Box primitives (csrc/box_prims.h: the six faces of an axis-aligned Cube answered by ONE slab test, intersect.cuh box_hit) against
    the same scene with the faces as six rectangle leaves, ray by ray and image by image; book-two (400 ground blocks) and a table
    scene whose cubes touch, share planes with a loose rectangle and enclose ray origins."""
    box = rt.Scene.example(device, "book-two", 1)
    rec = rt.Scene.example(device, "book-two", 1, options={"flags": rt.SCENE_NO_BOX_PRIMS})
    info = box.extend_info()
    assert info["box_prims"] == 400 and info["leaf_items"] == 3408 - 6 * 400 + 400 and info["whole_tree_threads"] == 1024
    assert rec.extend_info() == {"box_prims": 0, "leaf_items": 3408, "whole_tree_threads": 0}
    assert box.bvh_info()["nodes"] < rec.bvh_info()["nodes"] // 2
    rng = np.random.default_rng(5)
    n = 60000
    dd = rng.normal(size=(n, 3))
    dd /= np.linalg.norm(dd, axis=1, keepdims=True)
    oo = np.stack([rng.uniform(-1100, 1100, n), rng.uniform(-20, 400, n), rng.uniform(-1100, 1100, n)], axis=1)  # above, between and INSIDE the blocks
    axis = np.concatenate([np.eye(3), -np.eye(3)])[rng.integers(0, 6, 4000)]
    ao = np.stack([rng.uniform(-1000, 1000, 4000), rng.uniform(0, 120, 4000), rng.uniform(-1000, 1000, 4000)], axis=1)
    rays = np.concatenate([np.concatenate([oo, dd], axis=1), np.concatenate([ao, axis], axis=1)]).astype(np.float32)
    hb, hr = box.trace(rays), rec.trace(rays)
    differ = (hb["prim"] != hr["prim"]) | (hb["hit"] != hr["hit"])
    assert differ.mean() <= 2e-4, differ.mean()  # rays through the very edge of a block: its two faces tie
    ok = ~differ & (hr["hit"] == 1)
    assert ok.sum() > 0.5 * len(rays)
    assert np.allclose(hb["t"][ok], hr["t"][ok], rtol=2e-5, atol=1e-4)
    assert np.allclose(hb["n"][ok], hr["n"][ok], atol=1e-5) and np.allclose(hb["p"][ok], hr["p"][ok], rtol=2e-5, atol=2e-3)
    # bounce rays that START on faces (self_prim = the face): oracle hits on the blocks, scattered on
    o_sc = orc.Scene("book-two", 1)
    first = o_sc.trace(orc.camera_rays(o_sc.camera(1.0), 96, 96), rel_perturb=0.0)
    kinds = o_sc.prims()["kind"]
    sel = (first["hit"] == 1) & (kinds[np.maximum(first["prim"], 0)] == 1)  # starts on rectangles: the blocks' faces (and the light)
    p, nrm = first["p"][sel], first["n"][sel]
    assert len(p) > 2000
    d2 = rng.normal(size=p.shape)
    d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
    r2 = np.concatenate([p, d2], axis=1).astype(np.float32)  # both hemispheres: into the block (comes out of another face) and away from it
    so = first["prim"][sel].astype(np.uint32)
    h2b, h2r = box.trace(r2, self_order=so), rec.trace(r2, self_order=so)
    d2f = (h2b["prim"] != h2r["prim"]) | (h2b["hit"] != h2r["hit"])
    assert d2f.mean() <= 1e-3, d2f.mean()
    assert np.all(h2b["prim"][h2b["hit"] == 1] != so[h2b["hit"] == 1])  # a face never re-hits itself
    ok2 = ~d2f & (h2r["hit"] == 1)
    assert np.allclose(h2b["t"][ok2], h2r["t"][ok2], rtol=2e-5, atol=1e-4)
    # images: same Philox keys, same hits -> same paths up to the edge grazers
    a, qa, sa = _render_with(rt, device, "book-two", 96, 96, 8, scene_opts={"flags": rt.SCENE_NO_BOX_PRIMS})
    b, qb, sb = _render_with(rt, device, "book-two", 96, 96, 8)
    c, qc, sc = _render_with(rt, device, "book-two", 96, 96, 8, scene_opts={"flags": rt.SCENE_NO_WHOLE_TREE})  # boxes read from global memory
    assert abs(int(sa.rays) - int(sb.rays)) <= 2e-3 * sa.rays and abs(int(sc.rays) - int(sb.rays)) <= 1e-4 * sb.rays
    same = np.abs(c - b).max(axis=2) <= 1e-5 * np.maximum(1.0, np.abs(c).max(axis=2))
    assert same.mean() >= 0.998, same.mean()
    close = np.abs(a - b).max(axis=2) <= 2e-3 * np.maximum(1.0, np.abs(a).max(axis=2))
    assert close.mean() >= 0.97, close.mean()
    assert abs(a.mean() - b.mean()) <= 5e-3 * a.mean()

    # a table scene: touching cubes (shared planes), a cube with a loose rectangle on one of its planes, glass cubes (rays start inside)
    I = np.eye(3)
    face = {"+x": _rot([0, 1, 0], 90), "-x": _rot([0, 1, 0], -90), "+y": _rot([1, 0, 0], -90), "-y": _rot([1, 0, 0], 90), "+z": I, "-z": _rot([0, 1, 0], 180)}
    mats = [(rt.MAT_LAMBERTIAN, (0.7, 0.6, 0.5), 0, 0), (rt.MAT_DIFFUSE_LIGHT, (3, 3, 3), 0, 0), (rt.MAT_METAL, (0.8, 0.8, 0.9), 0.05, 0),
            (rt.MAT_DIELECTRIC, (1, 1, 1), 0, 1.5)]
    rects = []

    def cube(centre, half, m):
        for name, sgn, ax in [("+x", 1, 0), ("-x", -1, 0), ("+y", 1, 1), ("-y", -1, 1), ("+z", 1, 2), ("-z", -1, 2)]:
            off = np.zeros(3)
            off[ax] = sgn * half[ax]
            ext = np.abs(face[name].T @ (2 * np.asarray(half, dtype=np.float64)))
            rects.append((face[name], np.asarray(centre, dtype=np.float64) + off, ext[0], ext[1], m))

    for i in range(5):
        for j in range(5):  # a 5 x 5 floor of touching blocks of random height
            cube([2.0 * i - 4.0, 0.5 * (0.5 + 0.3 * ((i * 7 + j * 3) % 5)) - 2.0, 2.0 * j - 4.0], [1.0, 0.5 * (0.5 + 0.3 * ((i * 7 + j * 3) % 5)), 1.0], (i + j) % 2 * 2)
    cube([0.0, 1.0, 0.0], [0.8, 0.8, 0.8], 3)    # glass: rays refract into it and start on its faces going inward
    cube([2.5, 0.5, -1.0], [0.5, 0.9, 0.5], 2)
    rects.append((face["-y"], [0, 6.0, 0], 6.0, 6.0, 1))          # light
    rects.append((face["+y"], [2.5, 1.4, -1.0], 2.0, 2.0, 0))     # a loose rectangle ON the top plane of the metal cube, overhanging it
    tilted = _rot([1, 1, 0], 30)                                  # a cube that is NOT axis-aligned: stays six rectangles
    for name, sgn, ax in [("+x", 1, 0), ("-x", -1, 0), ("+y", 1, 1), ("-y", -1, 1), ("+z", 1, 2), ("-z", -1, 2)]:
        off = np.zeros(3)
        off[ax] = sgn * 0.6
        rects.append((tilted @ face[name], np.array([-2.5, 1.5, 1.0]) + tilted @ off, 1.2, 1.2, 0))
    tables = _rect_tables(rt, rects, mats)
    sb2 = rt.Scene.from_tables(device, *tables)
    sr2 = rt.Scene.from_tables(device, *tables, options={"flags": rt.SCENE_NO_BOX_PRIMS})
    assert sb2.extend_info()["box_prims"] == 27 and sr2.extend_info()["box_prims"] == 0
    n = 50000
    dd = rng.normal(size=(n, 3))
    dd /= np.linalg.norm(dd, axis=1, keepdims=True)
    oo = rng.uniform(-6, 6, size=(n, 3))
    rays = np.concatenate([oo, dd], axis=1).astype(np.float32)
    hb, hr = sb2.trace(rays), sr2.trace(rays)
    differ = (hb["prim"] != hr["prim"]) | (hb["hit"] != hr["hit"])
    assert differ.mean() <= 5e-4, differ.mean()
    ok = ~differ & (hr["hit"] == 1)
    assert np.allclose(hb["t"][ok], hr["t"][ok], rtol=2e-5, atol=1e-5)
    cam2 = rt.make_camera(rt.camera_args([0, 5, -14], [0, 0, 0], [0, 1, 0], math.radians(40), 1.0, 1.0, 0.0))
    a, qa, sa = _render_with(rt, device, None, 96, 96, 16, depth=20, tables=tables, cam=cam2, scene_opts={"flags": rt.SCENE_NO_BOX_PRIMS})
    b, qb, sb = _render_with(rt, device, None, 96, 96, 16, depth=20, tables=tables, cam=cam2)
    assert sa.rays > 96 * 96 * 16 * 2
    assert abs(int(sa.rays) - int(sb.rays)) <= 3e-3 * sa.rays
    close = np.abs(a - b).max(axis=2) <= 2e-3 * np.maximum(1.0, np.abs(a).max(axis=2))
    assert close.mean() >= 0.97, close.mean()
    assert abs(a.mean() - b.mean()) <= 1e-2 * a.mean()


@pytest.mark.gpu
def test_fused_extend_shade_matches_separate_kernels(rt, device):
    """Small all-rectangle scenes intersect inside k_shade (k_shade<., ., true>: frame_rects_closest, the body of k_extend_rects, then
    the shading of the same thread) instead of through the hit queue.  Same arithmetic on the same rays: identical hits, paths and ray
    counts; the sums differ by the order of the f64 accumulations only.  Queues smaller than a block round and tail iterations
    (the fused kernel rolls the ring counters itself) included."""
    for w, h, spp, cap in [(96, 96, 32, 0), (97, 61, 40, 5000), (64, 64, 3, 1056)]:
        a, qa, sa = _render_with(rt, device, "cornell-box", w, h, spp, queue_capacity=cap, ctx_opts={"flags": rt.CTX_NO_FUSED_RECTS})
        b, qb, sb = _render_with(rt, device, "cornell-box", w, h, spp, queue_capacity=cap, ctx_opts={"flags": 0})
        assert sa.rays == sb.rays and sa.paths == sb.paths == w * h * spp
        assert sb.launches < sa.launches  # one dependent launch fewer per wavefront iteration
        assert np.allclose(a, b, rtol=1e-5, atol=1e-5) and np.allclose(qa, qb, rtol=1e-5, atol=1e-5)
    # a full-size render takes the lanes path (three lanes on frame-rect scenes) in both forms
    a, _, sa = _render_with(rt, device, "cornell-box", 600, 600, 12, ctx_opts={"flags": rt.CTX_NO_FUSED_RECTS})
    b, _, sb = _render_with(rt, device, "cornell-box", 600, 600, 12, ctx_opts={"flags": 0})
    assert sa.rays == sb.rays and np.allclose(a, b, rtol=1e-5, atol=1e-5)
    # scenes that do not use the frame-rect list are not affected by the option
    c, _, sc = _render_with(rt, device, "checker-lists", 64, 36, 4, ctx_opts={"flags": rt.CTX_NO_FUSED_RECTS})
    d, _, sd = _render_with(rt, device, "checker-lists", 64, 36, 4, ctx_opts={"flags": 0})
    assert sc.rays == sd.rays and sc.launches == sd.launches and np.allclose(c, d, rtol=1e-5, atol=1e-5)
