"""Known-answer tests that pin the f64 CPU oracle (oracle/oracle.hpp).

The reference (aiifabbf/ray-tracer) holds no asserting tests or golden vectors (its only test prints,
src/mat4.rs:398-408) and cannot be run here (Rust), so the oracle is pinned by closed-form answers,
numpy cross-checks and the published Random123 vectors for Philox4x32-10.  Each test cites the
reference lines whose behaviour it checks.
"""
import math

import numpy as np
import pytest


# ---- Philox4x32-10: Random123 kat_vectors ------------------------------------------------------------
@pytest.mark.parametrize("ctr,key,expect", [
    ([0, 0, 0, 0], [0, 0], [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]),
    ([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2, [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]),
    ([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0], [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]),
])
def test_philox_known_answers(orc, ctr, key, expect):
    assert [int(v) for v in orc.philox(ctr, key)] == expect


# ---- Mat4 / Vec4 (src/mat4.rs, src/vec4.rs) ----------------------------------------------------------
def _cm(a):  # column-major [16] -> numpy 4x4
    return np.asarray(a).reshape(4, 4).T


def _rodrigues(theta, axis):
    x, y, z = axis
    K = np.array([[0, -z, y], [z, 0, -x], [-y, x, 0]])
    R = np.eye(3) + math.sin(theta) * K + (1 - math.cos(theta)) * (K @ K)
    M = np.eye(4)
    M[:3, :3] = R
    return M


def test_mat4_translation_layout(orc):  # mat4.rs:36-47: translation in a[12..14]
    m = orc.mat4_translation([1.5, -2.0, 3.25])
    assert list(m[12:15]) == [1.5, -2.0, 3.25] and m[15] == 1.0
    assert np.array_equal(_cm(m)[:3, :3], np.eye(3))


@pytest.mark.parametrize("axis", [(1, 0, 0), (0, 1, 0), (0, 0, 1), (0.6, 0.0, 0.8)])
@pytest.mark.parametrize("deg", [-90.0, -18.0, 15.0, 90.0, 180.0])
def test_mat4_rotation_is_rodrigues(orc, axis, deg):  # mat4.rs:52-80
    m = _cm(orc.mat4_rotation(math.radians(deg), axis))
    assert np.allclose(m, _rodrigues(math.radians(deg), axis), atol=1e-15)


def test_mat4_multiplied_is_self_times_other(orc):  # mat4.rs:82-143: T.multiplied(R) = T*R (rotate, then translate)
    rng = np.random.default_rng(3)
    a, b = rng.normal(size=16), rng.normal(size=16)
    assert np.allclose(_cm(orc.mat4_multiplied(a, b)), _cm(a) @ _cm(b), rtol=1e-14, atol=1e-14)
    t = orc.mat4_translation([5, 0, 0])
    r = orc.mat4_rotation(math.radians(90), [0, 1, 0])
    p = orc.vec4_transformed([0, 0, 1, 1], orc.mat4_multiplied(t, r))  # +z rotates to +x, then moves by 5
    assert np.allclose(p, [6, 0, 0, 1], atol=1e-15)


def test_mat4_inverse_and_singular(orc):  # mat4.rs:146-243, Mat4Cached mat4.rs:413-451
    rng = np.random.default_rng(5)
    for _ in range(20):
        a = rng.normal(size=16)
        inv, det = orc.mat4_inversed(a)
        assert np.isclose(det, np.linalg.det(_cm(a)), rtol=1e-10)
        assert np.allclose(_cm(inv), np.linalg.inv(_cm(a)), rtol=1e-8, atol=1e-10)
    sing = np.zeros(16)
    sing[0] = sing[5] = 1.0  # rank 2
    inv, det = orc.mat4_inversed(sing)
    assert inv is None and det == 0.0


def test_vec4_transformed_point_vs_vector(orc):  # vec4.rs:78-91, vec3.rs:66-73
    t = orc.mat4_translation([1, 2, 3])
    assert np.array_equal(orc.vec4_transformed([1, 1, 1, 1], t), [2, 3, 4, 1])
    assert np.array_equal(orc.vec4_transformed([1, 1, 1, 0], t), [1, 1, 1, 0])


def test_cube_faces_outward(orc):  # geometry.rs:250-287: front, left, back, right, top, bottom
    n, c = orc.cube_faces(2.0, 4.0, 6.0)
    expect_n = np.array([[0, 0, 1], [-1, 0, 0], [0, 0, -1], [1, 0, 0], [0, 1, 0], [0, -1, 0]], dtype=float)
    expect_c = np.array([[0, 0, 3], [-1, 0, 0], [0, 0, -3], [1, 0, 0], [0, 2, 0], [0, -2, 0]], dtype=float)
    assert np.allclose(n, expect_n, atol=1e-15) and np.allclose(c, expect_c, atol=1e-15)


# ---- Sphere::hit (src/geometry.rs:35-74) ------------------------------------------------------------
def test_sphere_outside_inside_behind_tangent(orc):
    h = orc.prim_hit(0, [1.0], [0, 0, -5, 0, 0, 1])
    assert h.hit == 1 and h.t == 4.0 and list(h.p) == [0, 0, -1] and list(h.n) == [0, 0, -1]
    assert h.u == pytest.approx(1.0) and h.v == pytest.approx(0.5)  # unitSphereUv geometry.rs:35-39
    h = orc.prim_hit(0, [1.0], [0, 0, 0, 0, 0, 1])  # origin inside: far root, normal still outward (A.4)
    assert h.hit == 1 and h.t == 1.0 and list(h.n) == [0, 0, 1]
    assert orc.prim_hit(0, [1.0], [0, 0, 5, 0, 0, 1]).hit == 0  # behind
    assert orc.prim_hit(0, [1.0], [1.0 + 1e-9, 0, -5, 0, 0, 1]).hit == 0  # just misses
    h = orc.prim_hit(0, [1.0], [1.0, 0, -5, 0, 0, 1])  # tangent: disc == 0 is a hit (disc < 0 rejects)
    assert h.hit == 1 and h.t == 5.0


def test_sphere_epsilon_and_unnormalised_direction(orc):
    # origin on the surface leaving: roots (-2, 0), neither > 1e-6 -> miss (geometry.rs:57-66)
    assert orc.prim_hit(0, [1.0], [0, 0, 1, 0, 0, 1]).hit == 0
    # origin on the surface going in: roots (0, 2) -> 2
    h = orc.prim_hit(0, [1.0], [0, 0, 1, 0, 0, -1])
    assert h.hit == 1 and h.t == 2.0
    # direction is not normalised; t is in units of |d| (A.5)
    h = orc.prim_hit(0, [1.0], [0, 0, -5, 0, 0, 2])
    assert h.t == 2.0
    # a root of 5e-7 is rejected, 2e-6 accepted
    assert orc.prim_hit(0, [1.0], [0, 0, -1 - 2e-6, 0, 0, 1]).t == pytest.approx(2e-6, rel=1e-6)
    assert orc.prim_hit(0, [1.0], [0, 0, -1 - 5e-7, 0, 0, 1]).t == pytest.approx(2.0, rel=1e-6)


# ---- Rectangle::hit (src/geometry.rs:152-181) --------------------------------------------------------
def test_rectangle_hit_edges_uv(orc):
    h = orc.prim_hit(1, [2.0, 4.0], [0.5, -1.0, 3.0, 0, 0, -1])
    assert h.hit == 1 and h.t == 3.0 and list(h.n) == [0, 0, 1] and (h.u, h.v) == (0.75, 0.25)
    assert orc.prim_hit(1, [2.0, 4.0], [1.0, 2.0, 3.0, 0, 0, -1]).hit == 1  # corner inclusive (geometry.rs:166)
    assert orc.prim_hit(1, [2.0, 4.0], [1.0 + 1e-12, 0, 3.0, 0, 0, -1]).hit == 0
    h = orc.prim_hit(1, [2.0, 4.0], [0, 0, -3.0, 0, 0, 1])  # from behind: still a hit, normal not flipped (A.4)
    assert h.hit == 1 and list(h.n) == [0, 0, 1]
    assert orc.prim_hit(1, [2.0, 4.0], [0, 0, 1.0, 1, 0, 0]).hit == 0  # parallel: t = -inf
    assert orc.prim_hit(1, [2.0, 4.0], [0, 0, 0.0, 1, 0, 0]).hit == 0  # in-plane: t = NaN
    assert orc.prim_hit(1, [2.0, 4.0], [0, 0, 5e-7, 0, 0, -1]).hit == 0  # t < 1e-6
    assert orc.prim_hit(1, [2.0, 4.0], [0, 0, 1e-6, 0, 0, -1]).hit == 1  # t == 1e-6 passes (rejects only t < 1e-6)


# ---- AxisAlignedBoundingBox::hit (src/optimize.rs:60-83) ---------------------------------------------
def test_aabb_slab_semantics(orc):
    box = [-1, -1, -1, 1, 1, 1]
    assert orc.prim_hit(2, box, [0, 0, -5, 0, 0, 1]).hit == 1
    assert orc.prim_hit(2, box, [0, 0, 5, 0, 0, 1]).hit == 0  # behind: tmax <= tmin = 0
    assert orc.prim_hit(2, box, [0, 0, 0, 0, 0, 1]).hit == 1  # origin inside
    assert orc.prim_hit(2, box, [0.5, 0.5, -5, 0, 0, 1]).hit == 1  # zero components, origin inside those slabs
    assert orc.prim_hit(2, box, [2.0, 0, -5, 0, 0, 1]).hit == 0  # zero component, origin outside the slab
    assert orc.prim_hit(2, box, [1.0, 0, -5, 0, 0, 1]).hit == 1  # origin ON the slab, zero component: 0*inf = NaN is ignored
    assert orc.prim_hit(2, [-1, -1, 0, 1, 1, 0], [0, 0, -5, 0, 0, 1]).hit == 0  # zero-thickness box: tmax <= tmin


# ---- PerspectiveCamera (src/camera.rs:25-59,90-107) --------------------------------------------------
def test_camera_frame_book_one(orc):
    cam = orc.camera_args([13, 2, 3], [0, 0, 0], [0, 1, 0], math.radians(20), 400 / 225, 10.0, 0.05)
    ll, hz, vt = orc.camera_frame(cam)
    height = 2 * math.tan(math.radians(20) / 2)
    width = 400 / 225 * height
    # u = up x w is NOT normalised (camera.rs:39): |u| = |v| = sin(angle(up, w)) = 0.98895
    assert np.linalg.norm(hz) / (width * 10) == pytest.approx(0.98895, abs=1e-5)
    assert np.linalg.norm(vt) / (height * 10) == pytest.approx(0.98895, abs=1e-5)
    eye = np.array([13.0, 2, 3])
    w = eye / np.linalg.norm(eye)
    assert np.allclose(ll, eye - hz / 2 - vt / 2 - w * 10, atol=1e-12)
    assert abs(np.dot(hz, vt)) < 1e-12 and abs(np.dot(hz, w)) < 1e-12


def test_camera_pinhole_and_lens_offset(orc):
    cam = orc.camera_args([277.5, 277.5, -800], [277.5, 277.5, 0], [0, 1, 0], math.radians(40), 1.0, 10.0, 0.0)
    r = orc.camera_ray(cam, 0.5, 0.5)
    assert np.allclose(r[:3], [277.5, 277.5, -800]) and np.allclose(r[3:], [0, 0, 1], atol=1e-15)
    # thin lens: the offset is the SCALAR rd.x*u + rd.y*v added to all three components (camera.rs:99-102)
    cam = orc.camera_args([13, 2, 3], [0, 0, 0], [0, 1, 0], math.radians(20), 16 / 9, 10.0, 0.05)
    for key in range(20):
        r = orc.camera_ray(cam, 0.3, 0.8, orc.PAIRED, 1, key, 0)
        off = r[:3] - np.array([13.0, 2, 3])
        assert off[0] == pytest.approx(off[1], abs=1e-12) and off[1] == pytest.approx(off[2], abs=1e-12)
        assert abs(off[0]) <= 0.05 * (0.3 + 0.8) + 1e-12
        assert np.linalg.norm(r[3:]) == pytest.approx(1.0, abs=1e-14)


# ---- materials (src/material.rs) ---------------------------------------------------------------------
def _scatter_many(orc, kind, param, ray, n_vec, mode, count, albedo=(0.5, 0.6, 0.7)):
    out = []
    for k in range(count):
        rc, o6, att = orc.scatter(kind, param, albedo, ray, [0, 0, 0], n_vec, mode, 7, k)
        out.append((rc, o6.copy(), att.copy()))
    return out


def test_lambertian_distribution_both_samplers(orc):  # material.rs:60-70 + util.rs:6-15: normalize(n + U(ball))
    ray = [0, 0, 5, 0, 0, -1]
    rng = np.random.default_rng(11)
    x = rng.normal(size=(200000, 3))
    x *= (rng.random(200000) ** (1 / 3) / np.linalg.norm(x, axis=1))[:, None]
    d = x + np.array([0, 0, 1.0])
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    ref_mean_cos = d[:, 2].mean()
    for mode in (orc.FAITHFUL, orc.PAIRED):
        res = _scatter_many(orc, 0, 0.0, ray, [0, 0, 1], mode, 4000)
        dirs = np.array([r[1][3:] for r in res])
        assert all(r[0] == 1 for r in res)
        assert np.allclose(np.linalg.norm(dirs, axis=1), 1.0, atol=1e-12)
        assert np.all(dirs[:, 2] >= -1e-12)  # n + ball never points below the surface
        assert dirs[:, 2].mean() == pytest.approx(ref_mean_cos, abs=0.012)
        assert abs(dirs[:, 0].mean()) < 0.03 and abs(dirs[:, 1].mean()) < 0.03
        assert np.allclose(res[0][2], [0.5, 0.6, 0.7])  # attenuation = albedo


def test_lambertian_ignores_facing(orc):  # A.4: a hit from behind scatters THROUGH the plane
    rc, o6, _ = orc.scatter(0, 0.0, [1, 1, 1], [0, 0, -5, 0, 0, 1], [0, 0, 0], [0, 0, 1])
    assert rc == 1 and o6[5] > 0.0


def test_metal_mirror_fuzz_and_absorption(orc):  # material.rs:98-119
    d = np.array([1.0, 0.0, -1.0])  # not normalised on purpose: reflected(normalize(d))
    rc, o6, att = orc.scatter(1, 0.0, [0.7, 0.6, 0.5], [0, 0, 5, *d], [0, 0, 0], [0, 0, 1])
    assert rc == 1 and np.allclose(o6[3:], np.array([1, 0, 1]) / math.sqrt(2), atol=1e-15) and np.allclose(att, [0.7, 0.6, 0.5])
    rc, _, _ = orc.scatter(1, 0.0, [1, 1, 1], [0, 0, -5, 0, 0, 1], [0, 0, 0], [0, 0, 1])  # d.n >= 0 -> absorbed
    assert rc == 0
    for mode in (orc.FAITHFUL, orc.PAIRED):
        res = _scatter_many(orc, 1, 0.5, [0, 0, 5, 0, 0, -1], [0, 0, 1], mode, 500)
        dirs = np.array([r[1][3:] for r in res])
        assert np.allclose(np.linalg.norm(dirs, axis=1), 1.0, atol=1e-12)
        # reflect = +z; fuzz 0.5 keeps the direction within asin(0.5) = 30 degrees of it
        assert dirs[:, 2].min() >= math.cos(math.radians(30)) - 1e-9


@pytest.mark.parametrize("cos_in", [1.0, 0.8, 0.5, 0.2])
def test_dielectric_reflect_fraction_matches_schlick(orc, cos_in):  # material.rs:140-193, A.10
    sin_in = math.sqrt(1 - cos_in ** 2)
    d = [sin_in, 0.0, -cos_in]
    ratio = 1 / 1.5
    r0 = ((ratio - 1) / (ratio + 1)) ** 2
    prob = r0 + (1 - r0) * (1 - cos_in) ** 5
    refl = np.array([sin_in, 0, cos_in])
    sin_t = ratio * sin_in
    refr = np.array([sin_t, 0, -math.sqrt(1 - sin_t ** 2)])
    for mode in (orc.FAITHFUL, orc.PAIRED):
        res = _scatter_many(orc, 2, 1.5, [0, 0, 5, *d], [0, 0, 1], mode, 3000)
        dirs = np.array([r[1][3:] for r in res])
        is_refl = dirs[:, 2] > 0
        assert np.allclose(dirs[is_refl], refl, atol=1e-12) and np.allclose(dirs[~is_refl], refr, atol=1e-12)
        sigma = math.sqrt(prob * (1 - prob) / 3000)
        assert abs(is_refl.mean() - prob) < 4 * sigma + 1e-3
        assert np.allclose(res[0][2], [1, 1, 1])


def test_dielectric_total_internal_reflection(orc):  # inside (d.n >= 0), ratio = 1.5, beyond the critical angle
    d = [math.sin(math.radians(60)), 0.0, math.cos(math.radians(60))]
    for k in range(10):
        rc, o6, _ = orc.scatter(2, 1.5, [1, 1, 1], [0, 0, -5, *d], [0, 0, 0], [0, 0, 1], orc.PAIRED, 1, k)
        assert rc == 1 and np.allclose(o6[3:], [d[0], 0, -d[2]], atol=1e-15)


def test_isotropic_uniform_on_sphere(orc):  # material.rs:317-326
    for mode in (orc.FAITHFUL, orc.PAIRED):
        res = _scatter_many(orc, 4, 0.0, [0, 0, 5, 0, 0, -1], [0, 0, 1], mode, 4000)
        dirs = np.array([r[1][3:] for r in res])
        assert np.allclose(np.linalg.norm(dirs, axis=1), 1.0, atol=1e-12)
        assert np.all(np.abs(dirs.mean(axis=0)) < 0.04)
        hist, _ = np.histogram(dirs[:, 2], bins=8, range=(-1, 1))  # z of a uniform direction is uniform
        assert np.all(np.abs(hist - 500) < 5 * math.sqrt(500))


# ---- ConstantMedium::hit (src/volume.rs:46-100, A.9) -------------------------------------------------
def test_medium_free_flight_from_outside(orc):
    r, rho = 2.0, 0.7
    ray = [0, 0, -5, 0, 0, 1]
    chord = 2 * r - 1e-6  # second probe starts 1e-6 past the entry point
    hits = [orc.medium_hit(r, rho, ray, orc.PAIRED, 3, k) for k in range(4000)]
    got = np.array([h.hit for h in hits], dtype=bool)
    p_hit = 1 - math.exp(-rho * chord)
    assert abs(got.mean() - p_hit) < 4 * math.sqrt(p_hit * (1 - p_hit) / 4000)
    t = np.array([h.t for h in hits])[got]
    assert t.min() >= 3.0 and t.max() <= 3.0 + chord + 1e-9  # t = t1 + distance (volume.rs:68)
    dist = t - 3.0
    # conditional CDF (1 - e^{-rho d}) / p_hit at the median
    med = -math.log(1 - 0.5 * p_hit) / rho
    assert abs((dist < med).mean() - 0.5) < 4 * math.sqrt(0.25 / got.sum())
    h = hits[int(np.argmax(got))]
    # quirk A.9 ii: the reported POINT is evaluated on the shadowed ray: o + (t1 + 1e-6) d + t d
    assert h.p[2] == pytest.approx(-5 + (3.0 + 1e-6) + h.t, abs=1e-12)
    assert np.allclose(list(h.n), [0, 0, 0], atol=1e-12)  # (n1 + n2) / 2 of opposite poles
    assert orc.medium_hit(r, rho, [5, 0, -5, 0, 0, 1], orc.PAIRED, 3, 0).hit == 0  # boundary missed


def test_medium_free_flight_from_inside(orc):  # volume.rs:77-96: t = distance, point on the ray, normal n1
    r, rho = 2.0, 0.4
    hits = [orc.medium_hit(r, rho, [0, 0, 0, 0, 0, 1], orc.FAITHFUL, 5, k) for k in range(4000)]
    got = np.array([h.hit for h in hits], dtype=bool)
    p_hit = 1 - math.exp(-rho * r)
    assert abs(got.mean() - p_hit) < 4 * math.sqrt(p_hit * (1 - p_hit) / 4000)
    h = hits[int(np.argmax(got))]
    assert 0 < h.t <= r and h.p[2] == pytest.approx(h.t, abs=1e-15) and list(h.n) == [0, 0, 1]


# ---- render::color + driver loop (src/render.rs:5-29, examples/book-one.rs:62-78) -------------------
def test_depth_semantics_and_black_background(orc):
    s = orc.Scene("cornell-box", 1)
    img0, _, _ = s.render(16, 16, 2, 0)
    assert np.all(img0 == 0.0)  # color(.., 0) = black without tracing (render.rs:6-8)
    img1, _, st = s.render(16, 16, 4, 1)
    assert st.segments == st.paths  # exactly one segment per path
    vals = np.unique(np.round(img1 / 4, 12))
    assert set(vals) <= {0.0, 3.75, 7.5, 11.25, 15.0}  # only direct hits of the emission-15 light count
    sky = orc.Scene("book-one", 1)
    img, _, _ = sky.render(32, 18, 1, 1)
    top = img[-1]  # highest image row looks at the sky sphere: DiffuseLight (0.5, 0.7, 1.0), book-one.rs:112-117
    assert np.allclose(top, [0.5, 0.7, 1.0])


def test_faithful_and_paired_samplers_agree_in_distribution(orc):
    s = orc.Scene("book-one", 1)
    w, h, spp = 24, 14, 256
    a, qa, sta = s.render(w, h, spp, 50, seed=1, mode=orc.FAITHFUL)
    b, qb, stb = s.render(w, h, spp, 50, seed=2, mode=orc.PAIRED)
    ma, mb = a / spp, b / spp
    var = (qa / spp - ma ** 2 + qb / spp - mb ** 2) / spp
    z = (ma - mb) / np.sqrt(np.maximum(var, 1e-12))
    assert np.mean(np.abs(z) > 3.0) < 0.02
    assert abs(z.mean()) < 0.15  # no systematic brightness difference between the two samplers
    assert abs(sta.segments / sta.paths - stb.segments / stb.paths) < 0.05


def test_row_sharding_is_exact(orc):  # rows y = row_begin + k*stride see the same Philox keys as the full image
    s = orc.Scene("cornell-box", 1)
    full, _, _ = s.render(20, 12, 2, 8, seed=4)
    for rank in range(3):
        part, _, _ = s.render(20, 12, 2, 8, seed=4, row_begin=rank, row_stride=3)
        assert np.array_equal(part, full[rank::3])


def test_resolve_u8(orc):  # examples/book-one.rs:90-100, A.12
    s = np.zeros((2, 2, 3))
    s[0, 0] = [4 * 0.25, 4 * 1.0, 4 * 4.0]  # bottom-left
    s[1, 1] = [np.nan, -4.0, 4 * 0.999]
    out = orc.resolve_u8(s, 4)
    assert list(out[1, 0]) == [127, 255, 255]  # rows top -> bottom: y = 0 is the LAST row
    assert out[0, 1, 0] == 255 and out[0, 1, 1] == 255 and out[0, 1, 2] == int(math.sqrt(0.999) * 255)


def test_color_on_caller_rays_known_answers(orc):
    """orc_color_rays = render::color (src/render.rs:5-29) on given rays: a ray that sees the cornell light returns its emission on
    every evaluation (DiffuseLight never scatters, material.rs:290-297), depth 0 is black (render.rs:6-8), a ray leaving through the
    open front is black (render.rs:26), and a diffuse bounce needs more than one segment."""
    s = orc.Scene("cornell-box", 1)
    rays = np.array([[277.5, 277.5, 277.5, 0, 1, 0],      # straight up into the light (y = 554, emission 15)
                     [277.5, 277.5, 277.5, 0, 2.5, 0],    # same, direction not normalised
                     [277.5, 277.5, 277.5, 0, 0, -1],     # out of the open front
                     [277.5, 277.5, 277.5, 0, -1, 0]],    # floor: scatters
                    dtype=np.float64)
    out, seg = s.color_rays(rays, 8, 50, seed=3)
    assert np.allclose(out[0], 15.0 * 8) and np.allclose(out[1], 15.0 * 8) and list(seg[:2]) == [8, 8]
    assert np.all(out[2] == 0.0) and seg[2] == 8
    assert seg[3] > 8 and np.all(out[3] >= 0.0)
    zero, seg0 = s.color_rays(rays, 4, 0, seed=3)
    assert np.all(zero == 0.0) and np.all(seg0 == 0)
    again, _ = s.color_rays(rays, 8, 50, seed=3)
    assert np.array_equal(out, again)  # dimensioned Philox draws: a function of (ray, sample, segment, seed) only
