"""Independent pins of what the parity tests share between the two sides.

The oracle and the product build their scenes with the SAME builder (ray-tracer_b200/examples/scenes.hpp), so a wrong wall,
radius or camera constant there is invisible to GPU-vs-oracle comparisons.  Here:
  * every constant of the three example scenes, transcribed by hand from examples/book-one.rs:35-47,103-205,
    examples/cornell-box.rs:31-150 and examples/main.rs:49-61,156-330, is asserted against the FLATTENED tables
    (centres, radii, sizes, outward normals, albedos, fuzz / ior, densities, camera arguments);
  * a second, numpy-only restatement of Sphere::hit / Rectangle::hit / Sprite::hit / ConstantMedium::hit (written from SURVEY
    Appendix A, including the free-flight draw with a numpy Philox4x32-10 and the generic two-probe form over a Cube boundary) runs over
    the flat tables of ALL scenes — book-two's media and the checker-lists smoke box included — against the oracle's BVH recursion."""
import math

import numpy as np
import pytest


def _rot(axis, deg):
    a = np.asarray(axis, dtype=np.float64)
    th = math.radians(deg)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    return np.eye(3) + math.sin(th) * K + (1 - math.cos(th)) * (K @ K)


def _world(prims):
    m = prims["world_from_local"].reshape(-1, 3, 4).astype(np.float64)
    return m[:, :, :3], m[:, :, 3]


def _albedo(flat, prims):
    mats, texs = flat.materials(), flat.textures()
    out = np.full((len(prims), 3), np.nan)
    for i, mi in enumerate(prims["material"]):
        if mi >= 0 and mats["texture"][mi] >= 0 and texs["kind"][mats["texture"][mi]] == 0:
            out[i] = texs["rgb"][mats["texture"][mi]][:3]
    return out, mats["kind"][np.maximum(prims["material"], 0)], mats


def _find(centres, c, tol=1e-4):
    hit = np.nonzero(np.abs(centres - np.asarray(c)).max(axis=1) <= tol)[0]
    assert len(hit) == 1, f"expected exactly one primitive at {c}, found {len(hit)}"
    return hit[0]


def test_cornell_box_constants(rt):  # examples/cornell-box.rs:31-150
    flat = rt.FlatScene("cornell-box", 1)
    prims = flat.prims()
    R, c = _world(prims)
    albedo, kind, mats = _albedo(flat, prims)
    n = R[:, :, 2]  # M (0, 0, 1)
    assert len(prims) == 18 and np.all(prims["kind"] == rt.PRIM_RECT)
    walls = [  # centre, normal, (width, height), colour, material kind
        ((555.0, 277.5, 277.5), (-1, 0, 0), (555, 555), (0.12, 0.45, 0.15), rt.MAT_LAMBERTIAN),  # green, rotated -90 deg about y
        ((0.0, 277.5, 277.5), (1, 0, 0), (555, 555), (0.65, 0.05, 0.05), rt.MAT_LAMBERTIAN),     # red, +90 deg about y
        ((277.5, 554.0, 277.5), (0, -1, 0), (130, 105), (15, 15, 15), rt.MAT_DIFFUSE_LIGHT),     # light, +90 deg about x
        ((277.5, 0.0, 277.5), (0, 1, 0), (555, 555), (0.73, 0.73, 0.73), rt.MAT_LAMBERTIAN),     # floor, -90 deg about x
        ((277.5, 555.0, 277.5), (0, -1, 0), (555, 555), (0.73, 0.73, 0.73), rt.MAT_LAMBERTIAN),  # ceiling
        ((277.5, 277.5, 555.0), (0, 0, -1), (555, 556), (0.73, 0.73, 0.73), rt.MAT_LAMBERTIAN),  # back wall, 555 x 556 (cornell-box.rs:77), 180 deg about y
    ]
    used = set()
    for centre, normal, size, colour, mk in walls:
        i = _find(c, centre)
        used.add(i)
        assert np.allclose(n[i], normal, atol=1e-6)
        assert tuple(prims["param"][i][:2]) == size
        assert np.allclose(albedo[i], colour, atol=1e-6) and kind[i] == mk
    # the light's 130 runs along world x, its 105 along world z (local y -> +z under a +90 deg rotation about x)
    i = _find(c, (277.5, 554.0, 277.5))
    assert np.allclose(R[i][:, 0], (1, 0, 0), atol=1e-6) and np.allclose(R[i][:, 1], (0, 0, 1), atol=1e-6)
    cubes = [((212.5, 82.5, 147.5), (165.0, 165.0, 165.0), -18.0), ((347.5, 165.0, 377.5), (165.0, 330.0, 165.0), 15.0)]
    for centre, (w, h, d), deg in cubes:
        Ry = _rot([0, 1, 0], deg)
        for axis, half, size in [((1, 0, 0), w / 2, {d, h}), ((0, 1, 0), h / 2, {w, d}), ((0, 0, 1), d / 2, {w, h})]:
            for sgn in (1, -1):
                out = Ry @ (sgn * np.asarray(axis, dtype=np.float64))
                i = _find(c, np.asarray(centre) + half * out, tol=1e-3)
                used.add(i)
                assert np.allclose(n[i], out, atol=1e-6), "cube faces carry OUTWARD normals (geometry.rs:250-287)"
                assert {float(prims["param"][i][0]), float(prims["param"][i][1])} == {float(s) for s in size}
                assert np.allclose(albedo[i], (0.73, 0.73, 0.73), atol=1e-6)
    assert len(used) == 18
    a = flat.camera_args(1.0)
    assert list(a.eye) == [277.5, 277.5, -800.0] and list(a.center) == [277.5, 277.5, 0.0] and list(a.up) == [0, 1, 0]
    assert a.fov == math.radians(40.0) and a.focus_distance == 10.0 and a.lens_radius == 0.0


def test_book_one_constants(rt):  # examples/book-one.rs:35-47,103-205
    flat = rt.FlatScene("book-one", 1)
    prims = flat.prims()
    R, c = _world(prims)
    albedo, kind, mats = _albedo(flat, prims)
    r = prims["param"][:, 0].astype(np.float64)
    assert np.all(prims["kind"] == rt.PRIM_SPHERE) and np.allclose(R, np.eye(3))  # translation-only sprites
    for centre, radius, mk, colour, extra in [((0, -1000, 0), 1000.0, rt.MAT_LAMBERTIAN, (0.5, 0.5, 0.5), None),
                                              ((0, 0, 0), 2000.0, rt.MAT_DIFFUSE_LIGHT, (0.5, 0.7, 1.0), None),
                                              ((-4, 1, 0), 1.0, rt.MAT_LAMBERTIAN, (0.4, 0.2, 0.1), None),
                                              ((4, 1, 0), 1.0, rt.MAT_METAL, (0.7, 0.6, 0.5), ("fuzz", 0.0)),
                                              ((0, 1, 0), 1.0, rt.MAT_DIELECTRIC, None, ("ior", 1.5))]:
        i = _find(c[np.isclose(r, radius)], centre)
        i = np.nonzero(np.isclose(r, radius))[0][i]
        assert kind[i] == mk
        if colour:
            assert np.allclose(albedo[i], colour, atol=1e-6)
        if extra:
            assert mats[extra[0]][prims["material"][i]] == np.float32(extra[1])
    small = np.isclose(r, 0.2)
    assert small.sum() == len(prims) - 5 and 22 * 22 - 8 <= small.sum() <= 22 * 22
    cs = c[small]
    assert np.allclose(cs[:, 1], 0.2)
    cell = np.floor(cs[:, [0, 2]] + 1e-12)
    frac = cs[:, [0, 2]] - cell
    assert np.all((cell >= -11) & (cell <= 10)) and np.all((frac >= 0) & (frac < 0.9 + 1e-9))  # a + 0.9 U, b + 0.9 U
    assert len({tuple(v) for v in cell}) == small.sum()                                         # one sphere per grid cell
    assert np.all(np.linalg.norm(cs - [4, 0.2, 0], axis=1) > 0.9)                               # the (4, 0.2, 0) exclusion
    km = kind[small]
    fuzz = mats["fuzz"][prims["material"][small]]
    assert np.all((fuzz[km == rt.MAT_METAL] >= 0) & (fuzz[km == rt.MAT_METAL] < 0.5))
    assert np.all((albedo[small][km == rt.MAT_METAL] >= 0.5) & (albedo[small][km == rt.MAT_METAL] < 1.0))
    assert np.all((albedo[small][km == rt.MAT_LAMBERTIAN] >= 0) & (albedo[small][km == rt.MAT_LAMBERTIAN] < 1.0))
    assert np.all(mats["ior"][prims["material"][small]][km == rt.MAT_DIELECTRIC] == np.float32(1.5))
    a = flat.camera_args(2.0)
    assert list(a.eye) == [13.0, 2.0, 3.0] and list(a.center) == [0, 0, 0] and list(a.up) == [0, 1, 0]
    assert a.fov == math.radians(20.0) and a.focus_distance == 10.0 and a.lens_radius == 0.05 and a.aspect == 2.0


def test_book_two_constants(rt):  # examples/main.rs:49-61,156-330
    flat = rt.FlatScene("book-two", 1)
    prims = flat.prims()
    R, c = _world(prims)
    albedo, kind, mats = _albedo(flat, prims)
    k, par = prims["kind"], prims["param"].astype(np.float64)
    sph, rect, med = k == rt.PRIM_SPHERE, k == rt.PRIM_RECT, k == rt.PRIM_MEDIUM_SPHERE
    assert (sph.sum(), rect.sum(), med.sum()) == (1005, 2401, 2)
    one = lambda centre, mask: np.nonzero(mask)[0][_find(c[mask], centre)]  # noqa: E731
    i = one((273.0, 554.0, 279.5), rect)  # light 300 x 265, +90 deg about x: normal -y, emission 7
    assert tuple(par[i][:2]) == (300.0, 265.0) and np.allclose(R[i][:, 2], (0, -1, 0), atol=1e-6) and np.allclose(albedo[i], 7.0) and kind[i] == rt.MAT_DIFFUSE_LIGHT
    for centre, radius, mk, colour, extra in [((400, 400, 200), 50.0, rt.MAT_LAMBERTIAN, (0.7, 0.3, 0.1), None),
                                              ((260, 150, 45), 50.0, rt.MAT_DIELECTRIC, None, ("ior", 1.5)),
                                              ((0, 150, 145), 50.0, rt.MAT_METAL, (0.8, 0.8, 0.9), ("fuzz", 1.0)),
                                              ((360, 150, 145), 70.0, rt.MAT_DIELECTRIC, None, ("ior", 1.5)),
                                              ((400, 200, 400), 100.0, rt.MAT_LAMBERTIAN, None, None)]:
        i = one(centre, sph & np.isclose(par[:, 0], radius))
        assert kind[i] == mk
        if colour:
            assert np.allclose(albedo[i], colour, atol=1e-6)
        if extra:
            assert mats[extra[0]][prims["material"][i]] == np.float32(extra[1])
    earth = one((400, 200, 400), sph)
    assert flat.textures()["kind"][mats["texture"][prims["material"][earth]]] == rt.TEX_IMAGE  # main.rs:266-283
    i = one((360, 150, 145), med)  # blue subsurface medium inside the r = 70 glass sphere (main.rs:251)
    assert par[i][0] == np.float32(70.0 - 1e-6) and par[i][1] == np.float32(0.03) and kind[i] == rt.MAT_ISOTROPIC and np.allclose(albedo[i], (0.2, 0.4, 0.9), atol=1e-6)
    i = one((0, 0, 0), med)        # scene fog (main.rs:260)
    assert par[i][0] == 5000.0 and par[i][1] == np.float32(0.0001) and np.allclose(albedo[i], 1.0)
    foam = sph & np.isclose(par[:, 0], 10.0)
    assert foam.sum() == 1000 and np.allclose(albedo[foam], 0.73, atol=1e-6)
    lo, hi = c[foam].min(axis=0), c[foam].max(axis=0)
    assert np.all(lo >= [-100, 270, 395]) and np.all(hi <= [65, 435, 560]) and np.all(hi - lo > 150)  # U(0, 165) + (-100, 270, 395)
    # 20 x 20 ground boxes of 100 x h x 100, h in [1, 101), six outward faces each, lambertian (0.48, 0.83, 0.53)
    ground = rect & (np.arange(len(prims)) != one((273.0, 554.0, 279.5), rect))
    assert ground.sum() == 2400 and np.allclose(albedo[ground], (0.48, 0.83, 0.53), atol=1e-6)
    tops = ground & (np.abs(R[:, 1, 2] - 1.0) < 1e-6)  # faces whose normal is +y
    assert tops.sum() == 400
    tc = c[tops]
    assert {(round(x), round(z)) for x, z in tc[:, [0, 2]]} == {(-950 + 100 * i, -950 + 100 * j) for i in range(20) for j in range(20)}
    assert np.all((tc[:, 1] >= 1.0) & (tc[:, 1] < 101.0)) and np.all(par[tops][:, :2] == 100.0)
    a = flat.camera_args(1.0)
    assert list(a.eye) == [477.5, 275.0, -600.0] and list(a.center) == [277.5, 277.5, 0.0] and a.fov == math.radians(40.0) and a.lens_radius == 0.0


# ---- numpy restatement of the closest-hit semantics over the FLAT tables --------------------------------------------------

def _philox_u24(c0, c1, c2, c3, k0, k1):
    """First word of Philox4x32-10 (Random123) as a 24-bit uniform in [0, 1); vectorised over c0 / c1 / k0 / k1."""
    M0, M1, W0, W1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57), np.uint64(0x9E3779B9), np.uint64(0xBB67AE85)
    mask = np.uint64(0xFFFFFFFF)
    c = [np.asarray(v, dtype=np.uint64) & mask for v in np.broadcast_arrays(c0, c1, c2, c3)]
    k = [np.asarray(v, dtype=np.uint64) & mask for v in np.broadcast_arrays(k0, k1, c[0], c[0])[:2]]
    for _ in range(10):
        p0, p1 = M0 * c[0], M1 * c[2]
        c = [(p1 >> np.uint64(32)) ^ c[1] ^ k[0], p1 & mask, (p0 >> np.uint64(32)) ^ c[3] ^ k[1], p0 & mask]
        k = [(k[0] + W0) & mask, (k[1] + W1) & mask]
    return (c[0] >> np.uint64(8)).astype(np.float64) / 16777216.0


def _surface_hit(p, o, d, t_min=1e-6):
    """Sphere::hit / Rectangle::hit behind Sprite::hit on one flat record: (t, world normal); t = inf on a miss."""
    m = p["local_from_world"].reshape(3, 4).astype(np.float64)
    w = p["world_from_local"].reshape(3, 4).astype(np.float64)
    lo, ld = o @ m[:, :3].T + m[:, 3], d @ m[:, :3].T
    if p["kind"] == 1:
        wd, ht = float(p["param"][0]), float(p["param"][1])
        with np.errstate(divide="ignore", invalid="ignore"):
            t = -lo[:, 2] / ld[:, 2]
            x, y = lo[:, 0] + ld[:, 0] * t, lo[:, 1] + ld[:, 1] * t
            bad = ~np.isfinite(t) | (t < t_min) | (x < -wd / 2) | (x > wd / 2) | (y < -ht / 2) | (y > ht / 2)
        return np.where(bad, np.inf, t), np.broadcast_to(w[:, 2], o.shape)
    r = float(p["param"][0])
    a, b, c = np.sum(ld * ld, axis=1), 2 * np.sum(lo * ld, axis=1), np.sum(lo * lo, axis=1) - r * r
    disc = b * b - 4 * a * c
    ok = disc >= 0
    sq = np.sqrt(np.where(ok, disc, 0))
    t1, t2 = (-b - sq) / (2 * a), (-b + sq) / (2 * a)
    t = np.where(ok, np.where(t1 > t_min, t1, np.where(t2 > t_min, t2, np.inf)), np.inf)
    lp = lo + ld * np.where(np.isfinite(t), t, 0.0)[:, None]
    nl = lp / r
    nl = nl / np.maximum(np.linalg.norm(nl, axis=1, keepdims=True), 1e-300)
    return t, nl @ w[:, :3].T


def _boundary_hit(bnd, o, d):
    t = np.full(len(o), np.inf)
    n = np.zeros_like(o)
    for p in bnd:
        tp, np_ = _surface_hit(p, o, d)
        win = tp < t
        t, n = np.where(win, tp, t), np.where(win[:, None], np_, n)
    return t, n


def _medium_hit(p, bnd, o, d, seed):
    """ConstantMedium::hit (volume.rs:46-100) on one flat medium record whose boundary surfaces are `bnd`."""
    n_rays = len(o)
    t1, n1 = _boundary_hit(bnd, o, d)
    hit1 = np.isfinite(t1)
    entering = hit1 & (np.einsum("ij,ij->i", n1, d) < 0.0)
    p1 = o + d * np.where(hit1, t1, 0.0)[:, None]
    t2, _ = _boundary_hit(bnd, p1 + 1e-6 * d, d)  # the re-shot probe from p1 + 1e-6 d (volume.rs:51-54)
    inside = np.where(entering, t2, t1)
    idx = np.arange(n_rays, dtype=np.uint64)
    u = _philox_u24(1, 0x80000000 | int(p["order"]), seed & 0xFFFFFFFF, seed >> 32, idx & np.uint64(0xFFFFFFFF), idx >> np.uint64(32))
    with np.errstate(divide="ignore"):
        dist = (-1.0 / float(p["param"][1])) * np.log(u)
    ok = hit1 & np.isfinite(inside) & ~(dist > inside)
    return np.where(ok, np.where(entering, t1 + dist, dist), np.inf)


def _closest(flat, rays, seed):
    prims, bnd_all = flat.prims(), flat.boundary_prims()
    o, d = rays[:, :3], rays[:, 3:]
    best_t, best_order = np.full(len(rays), np.inf), np.full(len(rays), -1, dtype=np.int64)
    for p in sorted(prims, key=lambda q: q["order"]):  # DFS order + strict < = the reference's tie rule
        if p["kind"] in (0, 1):
            t, _ = _surface_hit(p, o, d)
        elif p["kind"] == 2:  # ConstantMedium<Sphere>: its boundary is the record's own sphere
            t = _medium_hit(p, [p], o, d, seed)
        else:
            first, count = int(p["aux"]), int(p["param"][2])
            t = _medium_hit(p, bnd_all[first:first + count], o, d, seed)
        win = t < best_t
        best_t, best_order = np.where(win, t, best_t), np.where(win, p["order"], best_order)
    return best_t, best_order


@pytest.mark.parametrize("name,w,h", [("book-one", 30, 20), ("cornell-box", 24, 24), ("book-two", 22, 22), ("checker-lists", 32, 18)])
def test_numpy_restatement_agrees_with_the_oracle_on_every_scene(rt, orc, name, w, h):
    flat = rt.FlatScene(name, 1)
    s = orc.Scene(name, 1)
    cam_rays = orc.camera_rays(s.camera(w / h), w, h)
    first = s.trace(cam_rays, rel_perturb=0.0)
    pts = first["p"][first["hit"] == 1]
    rng = np.random.default_rng(2)
    dirs = rng.normal(size=(len(pts), 3))
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    bounce = np.concatenate([pts + 1e-3 * dirs * np.maximum(1.0, np.abs(pts).max(axis=1, keepdims=True)), dirs], axis=1)
    rays = np.concatenate([cam_rays, bounce])
    seed = 5
    hits = s.trace(rays, seed=seed)
    t, order = _closest(flat, rays, seed)
    st = hits["stable"] == 1
    assert st.mean() > 0.9
    assert np.array_equal(np.isfinite(t)[st], hits["hit"][st] == 1)
    assert np.array_equal(order[st], hits["prim"][st])
    both = st & np.isfinite(t)
    tol = 1e-4 * np.maximum(hits["t"][both], 1e-3) + 64 * hits["t_sens"][both]  # f32 matrices / radii in the flat tables
    assert np.all(np.abs(t[both] - hits["t"][both]) <= tol)
    kinds = flat.prims()["kind"]
    if (kinds >= 2).any():  # the media really decide some of these rays
        med_orders = set(flat.prims()["order"][kinds >= 2].tolist())
        assert np.isin(hits["prim"][st], list(med_orders)).sum() > 20
