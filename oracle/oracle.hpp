// ORACLE — TEST INFRASTRUCTURE ONLY.
//
// CPU restatement (C++17, f64) of the render path of aiifabbf/ray-tracer, function by function,
// quirks included.  It is the checker for the CUDA path and the timed host-CPU
// baseline; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// may load it.  The product (ray-tracer_b200/) never includes or links anything from oracle/.
//
// PARITY UNPINNED: the reference has no golden vectors or asserting tests (its only test prints,
// src/mat4.rs:398-408), all of its randomness is OS-seeded, and no Rust toolchain exists here to
// run it.  The oracle is pinned instead by closed-form known answers (tests/test_oracle_*.py) and by
// numpy cross-checks of the matrix code.
//
// Every class/function cites the reference file:line it follows.  Names mirror the reference
// (Sprite, Hit, Bound, BoundingVolumeHierarchyNode, PerspectiveCamera, color ...).
#pragma once
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <functional>
#include <limits>
#include <memory>
#include <optional>
#include <utility>
#include <vector>

#include "philox.hpp"

namespace orc {

constexpr double PI = 3.14159265358979323846;  // std::f64::consts::PI
constexpr double INF = std::numeric_limits<double>::infinity();

// ------------------------------------------------------------------------------------------------
// Vec3 — src/vec3.rs:5-281
// ------------------------------------------------------------------------------------------------
struct Vec3 {
    double x = 0, y = 0, z = 0;
    Vec3() = default;
    Vec3(double x_, double y_, double z_) : x(x_), y(y_), z(z_) {}
    static Vec3 ex() { return {1, 0, 0}; }
    static Vec3 ey() { return {0, 1, 0}; }
    static Vec3 ez() { return {0, 0, 1}; }
    double operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
    double dot(const Vec3& o) const { return x * o.x + y * o.y + z * o.z; }  // vec3.rs:76-78
    Vec3 cross(const Vec3& o) const {                                       // vec3.rs:80-86
        return {y * o.z - z * o.y, -(x * o.z - z * o.x), x * o.y - y * o.x};
    }
    double length() const { return std::sqrt(x * x + y * y + z * z); }  // vec3.rs:88-90
    Vec3 normalized() const { return *this / length(); }               // vec3.rs:96-98
    Vec3 operator-() const { return {-x, -y, -z}; }
    Vec3 operator+(const Vec3& o) const { return {x + o.x, y + o.y, z + o.z}; }
    Vec3 operator-(const Vec3& o) const { return {x - o.x, y - o.y, z - o.z}; }
    Vec3 operator*(const Vec3& o) const { return {x * o.x, y * o.y, z * o.z}; }  // piecewise
    Vec3 operator+(double s) const { return {x + s, y + s, z + s}; }            // vec3.rs:153-163 (broadcast)
    Vec3 operator-(double s) const { return {x - s, y - s, z - s}; }            // vec3.rs:185-195
    Vec3 operator*(double s) const { return {x * s, y * s, z * s}; }
    Vec3 operator/(double s) const { return {x / s, y / s, z / s}; }
    Vec3& operator+=(const Vec3& o) { x += o.x; y += o.y; z += o.z; return *this; }
    // vec3.rs:100-102
    Vec3 reflected(const Vec3& n) const { return *this - n * this->dot(n) * 2.0; }
    // vec3.rs:113-124: dt from the *normalised* self, tangential term from the un-normalised self.
    std::optional<Vec3> refracted(const Vec3& n, double ratio) const {
        const Vec3 uv = normalized();
        const double dt = uv.dot(n);
        const double disc = 1.0 - ratio * ratio * (1.0 - dt * dt);
        if (disc > 0.0) return (*this - n * dt) * ratio - n * std::sqrt(disc);
        return std::nullopt;
    }
};
inline Vec3 operator*(double s, const Vec3& v) { return v * s; }

// ------------------------------------------------------------------------------------------------
// Mat4 / Vec4 / Mat4Cached — src/mat4.rs, src/vec4.rs.  Column-major [f64;16] (mat4.rs:14-17).
// ------------------------------------------------------------------------------------------------
struct Mat4 {
    double a[16] = {0};
    static Mat4 identity() { Mat4 m; m.a[0] = m.a[5] = m.a[10] = m.a[15] = 1.0; return m; }  // mat4.rs:21-28
    static Mat4 zero() { return Mat4{}; }                                                    // mat4.rs:30-32
    static Mat4 translation(const Vec3& o) {                                                 // mat4.rs:36-47
        Mat4 m = identity(); m.a[12] = o.x; m.a[13] = o.y; m.a[14] = o.z; return m;
    }
    static Mat4 rotation(double radians, const Vec3& axis) {  // mat4.rs:52-80 (axis assumed unit)
        const double x = axis.x, y = axis.y, z = axis.z, s = std::sin(radians), c = std::cos(radians), t = 1.0 - c;
        Mat4 m;
        const double v[16] = {x * x * t + c,     y * x * t + z * s, z * x * t - y * s, 0.0,
                              x * y * t - z * s, y * y * t + c,     z * y * t + x * s, 0.0,
                              x * z * t + y * s, y * z * t - x * s, z * z * t + c,     0.0,
                              0.0,               0.0,               0.0,               1.0};
        std::copy(v, v + 16, m.a);
        return m;
    }
    // mat4.rs:85-143: self * other (column-major), i.e. `other` is applied first.
    Mat4 multiplied(const Mat4& o) const {
        Mat4 r;
        for (int col = 0; col < 4; ++col)
            for (int row = 0; row < 4; ++row) {
                // same summation order as the reference: b0*a0r + b1*a1r + b2*a2r + b3*a3r
                r.a[col * 4 + row] = o.a[col * 4 + 0] * a[0 + row] + o.a[col * 4 + 1] * a[4 + row] +
                                     o.a[col * 4 + 2] * a[8 + row] + o.a[col * 4 + 3] * a[12 + row];
            }
        return r;
    }
    struct Cof { double b00, b01, b02, b03, b04, b05, b06, b07, b08, b09, b10, b11; };
    Cof cof() const {  // shared by determinant (mat4.rs:146-181) and inversed (mat4.rs:184-243)
        const double a00 = a[0], a01 = a[1], a02 = a[2], a03 = a[3], a10 = a[4], a11 = a[5], a12 = a[6], a13 = a[7],
                     a20 = a[8], a21 = a[9], a22 = a[10], a23 = a[11], a30 = a[12], a31 = a[13], a32 = a[14], a33 = a[15];
        return {a00 * a11 - a01 * a10, a00 * a12 - a02 * a10, a00 * a13 - a03 * a10, a01 * a12 - a02 * a11,
                a01 * a13 - a03 * a11, a02 * a13 - a03 * a12, a20 * a31 - a21 * a30, a20 * a32 - a22 * a30,
                a20 * a33 - a23 * a30, a21 * a32 - a22 * a31, a21 * a33 - a23 * a31, a22 * a33 - a23 * a32};
    }
    double determinant() const {
        const Cof c = cof();
        return c.b00 * c.b11 - c.b01 * c.b10 + c.b02 * c.b09 + c.b03 * c.b08 - c.b04 * c.b07 + c.b05 * c.b06;
    }
    std::optional<Mat4> inversed() const {
        const double det = determinant();
        if (det == 0.0) return std::nullopt;
        const Cof c = cof();
        const double a00 = a[0], a01 = a[1], a02 = a[2], a03 = a[3], a10 = a[4], a11 = a[5], a12 = a[6], a13 = a[7],
                     a20 = a[8], a21 = a[9], a22 = a[10], a23 = a[11], a30 = a[12], a31 = a[13], a32 = a[14], a33 = a[15];
        Mat4 r;
        const double v[16] = {
            (a11 * c.b11 - a12 * c.b10 + a13 * c.b09) / det, (a02 * c.b10 - a01 * c.b11 - a03 * c.b09) / det,
            (a31 * c.b05 - a32 * c.b04 + a33 * c.b03) / det, (a22 * c.b04 - a21 * c.b05 - a23 * c.b03) / det,
            (a12 * c.b08 - a10 * c.b11 - a13 * c.b07) / det, (a00 * c.b11 - a02 * c.b08 + a03 * c.b07) / det,
            (a32 * c.b02 - a30 * c.b05 - a33 * c.b01) / det, (a20 * c.b05 - a22 * c.b02 + a23 * c.b01) / det,
            (a10 * c.b10 - a11 * c.b08 + a13 * c.b06) / det, (a01 * c.b08 - a00 * c.b10 - a03 * c.b06) / det,
            (a30 * c.b04 - a31 * c.b02 + a33 * c.b00) / det, (a21 * c.b02 - a20 * c.b04 - a23 * c.b00) / det,
            (a11 * c.b07 - a10 * c.b09 - a12 * c.b06) / det, (a00 * c.b09 - a01 * c.b07 + a02 * c.b06) / det,
            (a31 * c.b01 - a30 * c.b03 - a32 * c.b00) / det, (a20 * c.b03 - a21 * c.b01 + a22 * c.b00) / det};
        std::copy(v, v + 16, r.a);
        return r;
    }
};

struct Vec4 {
    double x, y, z, w;
    // vec4.rs:78-91
    Vec4 transformed(const Mat4& t) const {
        const double* m = t.a;
        return {m[0] * x + m[4] * y + m[8] * z + m[12] * w, m[1] * x + m[5] * y + m[9] * z + m[13] * w,
                m[2] * x + m[6] * y + m[10] * z + m[14] * w, m[3] * x + m[7] * y + m[11] * z + m[15] * w};
    }
    Vec3 xyz() const { return {x, y, z}; }
};
inline Vec4 xyz1(const Vec3& v) { return {v.x, v.y, v.z, 1.0}; }  // vec3.rs:66-68 (point)
inline Vec4 xyz0(const Vec3& v) { return {v.x, v.y, v.z, 0.0}; }  // vec3.rs:71-73 (vector)

// mat4.rs:413-451
struct Mat4Cached {
    Mat4 origin, inv;
    double det = 1.0;
    Mat4Cached() : origin(Mat4::identity()), inv(Mat4::identity()), det(1.0) {}
    Mat4Cached(const Mat4& m) : origin(m), det(m.determinant()) {  // NOLINT (Into<Mat4Cached> for Mat4)
        auto i = m.inversed();
        inv = i ? *i : Mat4::zero();
    }
    const Mat4* inversed() const { return det == 0.0 ? nullptr : &inv; }
};

// ------------------------------------------------------------------------------------------------
// Random source.  The reference uses thread_rng()/rand::random (thread-local, OS seeded).  The
// oracle keeps the thread-local shape but draws from Philox (see philox.hpp).  Two modes:
//   FAITHFUL — sequential per-path stream, rejection sampling exactly as util.rs:6-42 (this is the
//              algorithm that is timed as the CPU baseline);
//   PAIRED   — dimensioned draws and direct (non-rejection) sampling, identical in distribution
//               and variate-for-variate identical to the CUDA path, so oracle and GPU can
//              be compared path by path.
// ------------------------------------------------------------------------------------------------
enum SamplerMode : int { FAITHFUL = 0, PAIRED = 1 };

struct Counters {
    uint64_t segments = 0, box_tests = 0, prim_tests = 0, medium_draws = 0;
};

struct PathRng {
    SamplerMode mode = PAIRED;
    uint32_t key0 = 0, key1 = 0;   // (global pixel index, sample index)  — or (ray index, 0) for trace
    uint32_t seed_lo = 0, seed_hi = 0;
    uint32_t bounce = 0;           // 0 = camera; k >= 1 = k-th path segment
    uint32_t seq = 0;              // FAITHFUL: position in the sequential stream
    Counters counters;
    // trace diagnostics: every successful leaf hit (t, leaf id) of the current query, so the caller can
    // find the runner-up primitive (near-tie policy of the fixed-ray-set tests); null = off
    std::vector<std::pair<double, int>>* leaf_log = nullptr;

    static constexpr uint32_t DIM_CAMERA = 0, DIM_SHADE = 1, DIM_SEQ = 2, DIM_MEDIUM = 0x80000000u;

    Philox4 draw(uint32_t b, uint32_t dim) const { return philox4x32_10(b, dim, seed_lo, seed_hi, key0, key1); }
    // FAITHFUL stream: one f64 uniform per call.
    double next() {
        const Philox4 r = draw(seq++, DIM_SEQ);
        return u53(r.v[0], r.v[1]);
    }
    // gen_range(low, high) of rand 0.7 for f64: low + (high-low)*U
    double gen_range(double low, double high) { return low + (high - low) * next(); }
};

inline PathRng*& tls_rng() {
    static thread_local PathRng* p = nullptr;
    return p;
}

// Direct samplers used by PAIRED mode (the CUDA path implements exactly these in f32).
inline Vec3 ball_from_uniforms(double u0, double u1, double u2) {
    const double z = 1.0 - 2.0 * u0, rho = std::sqrt(std::max(0.0, 1.0 - z * z)), phi = 2.0 * PI * u1, r = std::cbrt(u2);
    return {r * rho * std::cos(phi), r * rho * std::sin(phi), r * z};
}
inline Vec3 disk_from_uniforms(double u0, double u1) {
    const double r = std::sqrt(u0), phi = 2.0 * PI * u1;
    return {r * std::cos(phi), r * std::sin(phi), 0.0};
}

// util.rs:6-15
inline Vec3 randomInUnitSphere() {
    PathRng& g = *tls_rng();
    if (g.mode == PAIRED) {
        const Philox4 r = g.draw(g.bounce, PathRng::DIM_SHADE);
        return ball_from_uniforms(u24(r.v[0]), u24(r.v[1]), u24(r.v[2]));
    }
    Vec3 p(1.0, 1.0, 1.0);
    while (p.dot(p) >= 1.0) p = Vec3(g.next(), g.next(), g.next()) * 2.0 - Vec3(1.0, 1.0, 1.0);
    return p;
}
// util.rs:27-42
inline Vec3 randomInUnitDisk() {
    PathRng& g = *tls_rng();
    if (g.mode == PAIRED) {
        const Philox4 r = g.draw(0, PathRng::DIM_CAMERA);
        return disk_from_uniforms(u24(r.v[2]), u24(r.v[3]));
    }
    for (;;) {
        const double px = g.gen_range(-1.0, 1.0), py = g.gen_range(-1.0, 1.0);
        const Vec3 p(px, py, 0.0);
        if (p.length() >= 1.0) continue;
        return p;
    }
}
// thread_rng().gen_range(0.0, 1.0) in material.rs:164 (Dielectric).
inline double randomShadeUniform() {
    PathRng& g = *tls_rng();
    if (g.mode == PAIRED) return u24(g.draw(g.bounce, PathRng::DIM_SHADE).v[0]);
    return g.gen_range(0.0, 1.0);
}
// thread_rng().gen_range(0.0, 1.0) in volume.rs:58-60,80-82 (ConstantMedium); one draw per
// (segment, medium primitive), independent of traversal order in PAIRED mode.
inline double randomMediumUniform(int prim_id) {
    PathRng& g = *tls_rng();
    g.counters.medium_draws++;
    if (g.mode == PAIRED) return u24(g.draw(g.bounce, PathRng::DIM_MEDIUM | (uint32_t)prim_id).v[0]);
    return g.gen_range(0.0, 1.0);
}

// ------------------------------------------------------------------------------------------------
// Ray / HitRecord / Hit — src/ray.rs
// ------------------------------------------------------------------------------------------------
struct Ray {  // ray.rs:6-33
    Vec3 origin, direction;
    Vec3 at(double t) const { return origin + direction * t; }
};

struct Material;

struct HitRecord {  // ray.rs:35-83; `prim` is oracle-only bookkeeping (leaf primitive id, see Scene::enumerate)
    double t = INF;  // ray.rs:89-99: default record has t = +inf so a box "hit" never wins
    Vec3 intersection, normal;
    const Material* material = nullptr;
    double u = 0.0, v = 0.0;
    int prim = -1;
};

struct Hit {  // ray.rs:85-87
    virtual ~Hit() = default;
    virtual std::optional<HitRecord> hit(const Ray& ray) const = 0;
};

// ------------------------------------------------------------------------------------------------
// AABB / Bound — src/optimize.rs:20-95
// ------------------------------------------------------------------------------------------------
struct AxisAlignedBoundingBox {
    Vec3 min, max;
    AxisAlignedBoundingBox merged(const AxisAlignedBoundingBox& o) const {  // optimize.rs:45-57
        return {{std::min(min.x, o.min.x), std::min(min.y, o.min.y), std::min(min.z, o.min.z)},
                {std::max(max.x, o.max.x), std::max(max.y, o.max.y), std::max(max.z, o.max.z)}};
    }
    // optimize.rs:60-83.  Returns only hit / no hit (the reference returns a default record, t = inf).
    bool hit(const Ray& ray) const {
        if (PathRng* g = tls_rng()) g->counters.box_tests++;
        double tmin = 0.0, tmax = INF;
        for (int i = 0; i < 3; ++i) {
            const double inv = 1.0 / ray.direction[i];
            double t0 = (min[i] - ray.origin[i]) * inv;
            double t1 = (max[i] - ray.origin[i]) * inv;
            if (inv < 0.0) std::swap(t0, t1);
            tmin = t0 > tmin ? t0 : tmin;
            tmax = t1 < tmax ? t1 : tmax;
            if (tmax <= tmin) return false;
        }
        return true;
    }
};
using AABB = AxisAlignedBoundingBox;

struct Bound : Hit {  // optimize.rs:88-95 (Bound<AxisAlignedBoundingBox>)
    virtual std::optional<AABB> bound() const = 0;
};
using BoundPtr = std::shared_ptr<const Bound>;

// optimize.rs:128-185 and :188-241 — box of the 8 transformed corners.
inline AABB transformedBound(const AABB& b, const Mat4& m) {
    const double x0 = b.min.x, y0 = b.min.y, z0 = b.min.z, x1 = b.max.x, y1 = b.max.y, z1 = b.max.z;
    const Vec3 corners[8] = {{x0, y0, z0}, {x1, y0, z0}, {x0, y1, z0}, {x0, y0, z1},
                             {x1, y1, z0}, {x1, y0, z1}, {x0, y1, z1}, {x1, y1, z1}};
    double lo[3] = {INF, INF, INF}, hi[3] = {-INF, -INF, -INF};
    for (const Vec3& c : corners) {
        const Vec3 p = xyz1(c).transformed(m).xyz();
        for (int i = 0; i < 3; ++i) {
            if (p[i] < lo[i]) lo[i] = p[i];
            if (p[i] > hi[i]) hi[i] = p[i];
        }
    }
    return {{lo[0], lo[1], lo[2]}, {hi[0], hi[1], hi[2]}};
}

// ------------------------------------------------------------------------------------------------
// Geometry — src/geometry.rs
// ------------------------------------------------------------------------------------------------
struct Sphere : Bound {  // geometry.rs:11-74, optimize.rs:104-113
    double radius;
    mutable int id = -1;
    explicit Sphere(double r) : radius(r) {}
    static std::pair<double, double> unitSphereUv(const Vec3& p) {  // geometry.rs:35-39
        return {0.5 + std::atan2(p.x, p.z) / (2.0 * PI), 1.0 - std::acos(p.y) / PI};
    }
    std::optional<HitRecord> hit(const Ray& ray) const override {  // geometry.rs:42-74
        if (PathRng* g = tls_rng()) g->counters.prim_tests++;
        const Vec3 center(0.0, 0.0, 0.0);
        const Vec3 oc = ray.origin - center;
        const double a = ray.direction.dot(ray.direction);
        const double b = oc.dot(ray.direction) * 2.0;
        const double c = oc.dot(oc) - radius * radius;
        const double disc = b * b - 4.0 * a * c;
        if (disc < 0.0) return std::nullopt;
        double t1 = (-b - std::sqrt(disc)) / (2.0 * a);
        double t2 = (-b + std::sqrt(disc)) / (2.0 * a);
        if (!(t1 < t2)) std::swap(t1, t2);
        const double eps = std::pow(10.0, -6.0);
        double t;
        if (t1 > eps) t = t1;
        else if (t2 > eps) t = t2;
        else return std::nullopt;
        HitRecord rec;
        rec.t = t;
        rec.intersection = ray.at(t);
        rec.normal = ((rec.intersection - center) / radius).normalized();
        const auto uv = unitSphereUv((rec.intersection - center) / radius);
        rec.u = uv.first; rec.v = uv.second;
        rec.prim = id;
        if (id >= 0) if (PathRng* g = tls_rng()) if (g->leaf_log) g->leaf_log->emplace_back(t, id);
        return rec;
    }
    std::optional<AABB> bound() const override {
        return AABB{Vec3(0, 0, 0) - Vec3(radius, radius, radius), Vec3(0, 0, 0) + Vec3(radius, radius, radius)};
    }
};

struct Rectangle : Bound {  // geometry.rs:126-181, optimize.rs:115-126
    double width, height;
    mutable int id = -1;
    Rectangle(double w, double h) : width(w), height(h) {}
    std::optional<HitRecord> hit(const Ray& ray) const override {  // geometry.rs:152-181
        if (PathRng* g = tls_rng()) g->counters.prim_tests++;
        const double z = 0.0;
        const double ax = -width / 2.0, ay = -height / 2.0, bx = width / 2.0, by = height / 2.0;
        const double t = (z - ray.origin.z) / ray.direction.z;
        if (std::isinf(t) || std::isnan(t) || t < 1e-6) return std::nullopt;
        const double x = ray.origin.x + ray.direction.x * t;
        const double y = ray.origin.y + ray.direction.y * t;
        if (x < ax || x > bx || y < ay || y > by) return std::nullopt;
        HitRecord rec;
        rec.t = t;
        rec.intersection = ray.at(t);
        rec.normal = Vec3(0.0, 0.0, 1.0);
        rec.u = (x - ax) / (bx - ax);
        rec.v = (y - ay) / (by - ay);
        rec.prim = id;
        if (id >= 0) if (PathRng* g = tls_rng()) if (g->leaf_log) g->leaf_log->emplace_back(t, id);
        return rec;
    }
    std::optional<AABB> bound() const override {
        return AABB{{-width / 2.0, -height / 2.0, 0.0 - 1e-6}, {width / 2.0, height / 2.0, 0.0 + 1e-6}};
    }
};

// Shared by TransformedGeometry::hit (geometry.rs:210-246) and Sprite::hit (sprite.rs:94-138):
// ray into object space with M^-1 (direction NOT renormalised, so t is shared), inner hit, point and
// normal back with the forward matrix M (no inverse-transpose, no renormalisation).
inline std::optional<HitRecord> transformedHit(const Hit& geometry, const Mat4Cached& transform, const Ray& ray) {
    const Mat4* inversed = transform.inversed();
    if (!inversed) return std::nullopt;  // det == 0: never hit (sprite.rs:131-134)
    const Ray local{xyz1(ray.origin).transformed(*inversed).xyz(), xyz0(ray.direction).transformed(*inversed).xyz()};
    auto rec = geometry.hit(local);
    if (!rec) return std::nullopt;
    HitRecord out = *rec;
    out.intersection = xyz1(rec->intersection).transformed(transform.origin).xyz();
    out.normal = xyz0(rec->normal).transformed(transform.origin).xyz();
    return out;
}

struct TransformedGeometry : Bound {  // geometry.rs:184-246, optimize.rs:188-241
    std::shared_ptr<const Bound> geometry;  // the reference holds T by value
    Mat4Cached transform;
    TransformedGeometry(std::shared_ptr<const Bound> g, const Mat4Cached& t) : geometry(std::move(g)), transform(t) {}
    std::optional<HitRecord> hit(const Ray& ray) const override {
        auto r = transformedHit(*geometry, transform, ray);
        if (r) r->material = nullptr;  // geometry.rs:234
        return r;
    }
    std::optional<AABB> bound() const override {
        auto b = geometry->bound();
        if (!b) return std::nullopt;
        return transformedBound(*b, transform.origin);
    }
};

struct Cube {  // geometry.rs:250-287: six transformed rectangles, normals outward
    static std::vector<std::shared_ptr<const Bound>> make(double width, double height, double depth) {
        auto rad = [](double deg) { return deg * PI / 180.0; };  // f64::to_radians
        auto face = [](double w, double h, const Mat4& m) {
            return std::static_pointer_cast<const Bound>(
                std::make_shared<TransformedGeometry>(std::make_shared<Rectangle>(w, h), Mat4Cached(m)));
        };
        return {
            face(width, height, Mat4::translation({0.0, 0.0, depth / 2.0})),                                                    // front
            face(depth, height, Mat4::translation({-width / 2.0, 0.0, 0.0}).multiplied(Mat4::rotation(rad(-90.0), Vec3::ey()))),  // left
            face(width, height, Mat4::translation({0.0, 0.0, -depth / 2.0}).multiplied(Mat4::rotation(rad(180.0), Vec3::ey()))),  // back
            face(depth, height, Mat4::translation({width / 2.0, 0.0, 0.0}).multiplied(Mat4::rotation(rad(90.0), Vec3::ey()))),    // right
            face(width, depth, Mat4::translation({0.0, height / 2.0, 0.0}).multiplied(Mat4::rotation(rad(-90.0), Vec3::ex()))),   // top
            face(width, depth, Mat4::translation({0.0, -height / 2.0, 0.0}).multiplied(Mat4::rotation(rad(90.0), Vec3::ex()))),   // bottom
        };
    }
};

// geometry.rs:76-116 and optimize.rs:256-334 — linear lists.
struct HitList : Bound {
    std::vector<std::shared_ptr<const Bound>> items;
    std::optional<HitRecord> hit(const Ray& ray) const override {
        std::optional<HitRecord> res;
        for (const auto& v : items)
            if (auto rec = v->hit(ray)) {
                if (!res) res = rec;
                else if (rec->t < res->t) res = rec;
            }
        return res;
    }
    std::optional<AABB> bound() const override {
        std::optional<AABB> res;
        for (const auto& v : items)
            if (auto other = v->bound()) res = res ? res->merged(*other) : *other;
        return res;
    }
};

// ------------------------------------------------------------------------------------------------
// Textures and materials — src/material.rs
// ------------------------------------------------------------------------------------------------
struct Texture {  // material.rs:196-198
    virtual ~Texture() = default;
    virtual Vec3 value(double u, double v, const Vec3& point) const = 0;
};
using TexturePtr = std::shared_ptr<const Texture>;

struct SolidColor : Texture {  // material.rs:200-215
    Vec3 color;
    explicit SolidColor(const Vec3& c) : color(c) {}
    Vec3 value(double, double, const Vec3&) const override { return color; }
};
struct CheckerTexture : Texture {  // material.rs:217-245
    TexturePtr black, white;
    CheckerTexture(TexturePtr b, TexturePtr w) : black(std::move(b)), white(std::move(w)) {}
    Vec3 value(double u, double v, const Vec3& p) const override {
        const double sine = std::sin(2.0 * PI * 10.0 * u) * std::sin(2.0 * PI * 10.0 * v);
        return sine > 0.0 ? black->value(u, v, p) : white->value(u, v, p);
    }
};
struct ImageTexture : Texture {  // material.rs:247-265: arbitrary closure of uv
    std::function<Vec3(double, double)> mapping;
    explicit ImageTexture(std::function<Vec3(double, double)> m) : mapping(std::move(m)) {}
    Vec3 value(double u, double v, const Vec3&) const override { return mapping(u, v); }
};
inline TexturePtr solid(const Vec3& c) { return std::make_shared<SolidColor>(c); }  // Into<Arc<dyn Texture>> for Vec3

struct Material {  // material.rs:13-21
    virtual ~Material() = default;
    virtual std::optional<std::pair<Ray, Vec3>> scatter(const Ray& rayIn, const HitRecord& rec) const = 0;
    virtual Vec3 emitted(double, double, const Vec3&) const { return {0.0, 0.0, 0.0}; }
};
using MaterialPtr = std::shared_ptr<const Material>;

struct Lambertian : Material {  // material.rs:23-70
    TexturePtr albedo;
    explicit Lambertian(TexturePtr a) : albedo(std::move(a)) {}
    explicit Lambertian(const Vec3& c) : albedo(solid(c)) {}
    std::optional<std::pair<Ray, Vec3>> scatter(const Ray&, const HitRecord& rec) const override {
        const Ray scattered{rec.intersection, (rec.normal + randomInUnitSphere()).normalized()};
        return std::make_pair(scattered, albedo->value(rec.u, rec.v, rec.intersection));
    }
};

struct Metal : Material {  // material.rs:72-119
    TexturePtr albedo;
    double fuzziness;
    Metal(TexturePtr a, double f) : albedo(std::move(a)), fuzziness(f) {}
    Metal(const Vec3& c, double f) : albedo(solid(c)), fuzziness(f) {}
    std::optional<std::pair<Ray, Vec3>> scatter(const Ray& rayIn, const HitRecord& rec) const override {
        if (rayIn.direction.dot(rec.normal) < 0.0) {
            const Vec3 reflected = rayIn.direction.normalized().reflected(rec.normal);
            const Ray scattered{rec.intersection,
                                fuzziness == 0.0 ? reflected : (reflected + fuzziness * randomInUnitSphere()).normalized()};
            return std::make_pair(scattered, albedo->value(rec.u, rec.v, rec.intersection));
        }
        return std::nullopt;
    }
};

struct Dielectric : Material {  // material.rs:121-193
    double refractive;
    explicit Dielectric(double r) : refractive(r) {}
    static double schlickReflectionProbability(double theta, double n1, double n2) {  // material.rs:140-143
        const double r0 = std::pow((n1 - n2) / (n1 + n2), 2.0);
        return r0 + (1.0 - r0) * std::pow(1.0 - std::cos(theta), 5.0);
    }
    std::optional<std::pair<Ray, Vec3>> scatter(const Ray& rayIn, const HitRecord& rec) const override {
        const Vec3 attenuation(1.0, 1.0, 1.0);
        double ratio;
        Vec3 normal = rec.normal;
        if (rayIn.direction.dot(rec.normal) < 0.0) ratio = 1.0 / refractive;  // outside
        else { ratio = refractive; normal = -normal; }                        // inside
        if (auto refracted = rayIn.direction.refracted(normal, ratio)) {
            const double theta = std::acos(-rayIn.direction.dot(normal));
            if (randomShadeUniform() < schlickReflectionProbability(theta, ratio, 1.0))
                return std::make_pair(Ray{rec.intersection, rayIn.direction.reflected(rec.normal)}, attenuation);
            return std::make_pair(Ray{rec.intersection, refracted->normalized()}, attenuation);
        }
        return std::make_pair(Ray{rec.intersection, rayIn.direction.reflected(normal)}, attenuation);  // total internal reflection
    }
};

struct DiffuseLight : Material {  // material.rs:273-298
    TexturePtr emission;
    explicit DiffuseLight(TexturePtr e) : emission(std::move(e)) {}
    explicit DiffuseLight(const Vec3& c) : emission(solid(c)) {}
    std::optional<std::pair<Ray, Vec3>> scatter(const Ray&, const HitRecord&) const override { return std::nullopt; }
    Vec3 emitted(double u, double v, const Vec3& p) const override { return emission->value(u, v, p); }
};

struct Isotropic : Material {  // material.rs:300-326
    TexturePtr albedo;
    explicit Isotropic(TexturePtr a) : albedo(std::move(a)) {}
    explicit Isotropic(const Vec3& c) : albedo(solid(c)) {}
    std::optional<std::pair<Ray, Vec3>> scatter(const Ray&, const HitRecord& rec) const override {
        const Ray scattered{rec.intersection, randomInUnitSphere().normalized()};
        return std::make_pair(scattered, albedo->value(rec.u, rec.v, rec.intersection));
    }
};

// ------------------------------------------------------------------------------------------------
// Sprite — src/sprite.rs
// ------------------------------------------------------------------------------------------------
struct Sprite : Bound {
    std::shared_ptr<const Bound> geometry_;  // Option<Arc<T>>
    MaterialPtr material_;                   // Option<Arc<U>>
    Mat4Cached transform_;

    struct Builder {  // sprite.rs:22-64
        std::shared_ptr<Sprite> s = std::make_shared<Sprite>();
        Builder& geometry(std::shared_ptr<const Bound> g) { s->geometry_ = std::move(g); return *this; }
        Builder& material(MaterialPtr m) { s->material_ = std::move(m); return *this; }
        Builder& transform(const Mat4& m) { s->transform_ = Mat4Cached(m); return *this; }
        std::shared_ptr<const Bound> build() { return s; }
    };
    static Builder builder() { return {}; }

    std::optional<HitRecord> hit(const Ray& ray) const override {  // sprite.rs:87-139
        if (!geometry_) return std::nullopt;
        auto r = transformedHit(*geometry_, transform_, ray);
        if (r) r->material = material_.get();  // sprite.rs:124: this sprite's material overrides
        return r;
    }
    std::optional<AABB> bound() const override {  // optimize.rs:128-185
        if (!geometry_) return std::nullopt;
        auto b = geometry_->bound();
        if (!b) return std::nullopt;
        return transformedBound(*b, transform_.origin);
    }
};

// ------------------------------------------------------------------------------------------------
// ConstantMedium — src/volume.rs (free-flight sampling inside hit()), optimize.rs:508-516
// ------------------------------------------------------------------------------------------------
struct ConstantMedium : Bound {
    std::shared_ptr<const Bound> boundary;
    double density;
    mutable int id = -1;
    ConstantMedium(std::shared_ptr<const Bound> b, double d) : boundary(std::move(b)), density(d) {}
    std::optional<HitRecord> hit(const Ray& ray0) const override {  // volume.rs:46-100
        auto record1 = boundary->hit(ray0);
        if (!record1) return std::nullopt;
        if (record1->normal.dot(ray0.direction) < 0.0) {
            // entering: re-shoot from p1 + 1e-6 d (volume.rs:51-54); note `ray` is shadowed
            const Ray ray{record1->intersection + ray0.direction * 1e-6, ray0.direction};
            auto record2 = boundary->hit(ray);
            if (!record2) return std::nullopt;
            const double inside = record2->t;
            const double distance = (-1.0 / density) * std::log(randomMediumUniform(id));
            if (distance > inside) return std::nullopt;
            HitRecord rec;
            rec.t = record1->t + distance;
            rec.intersection = ray.at(record1->t + distance);  // volume.rs:69 — evaluated on the SHADOWED ray (A.9)
            rec.normal = (record1->normal + record2->normal) / 2.0;
            rec.material = nullptr;
            rec.u = record1->u + record2->u;
            rec.v = record1->v + record2->v;
            rec.prim = id;
            if (PathRng* g = tls_rng()) if (g->leaf_log) g->leaf_log->emplace_back(rec.t, id);
            return rec;
        }
        // origin inside the boundary (volume.rs:77-96)
        const double inside = record1->t;
        const double distance = (-1.0 / density) * std::log(randomMediumUniform(id));
        if (distance > inside) return std::nullopt;
        HitRecord rec;
        rec.t = distance;
        rec.intersection = ray0.at(distance);
        rec.normal = record1->normal;
        rec.material = nullptr;
        rec.u = record1->u; rec.v = record1->v;
        rec.prim = id;
        if (PathRng* g = tls_rng()) if (g->leaf_log) g->leaf_log->emplace_back(rec.t, id);
        return rec;
    }
    std::optional<AABB> bound() const override { return boundary->bound(); }
};

// ------------------------------------------------------------------------------------------------
// BoundingVolumeHierarchyNode — src/optimize.rs:336-506
// ------------------------------------------------------------------------------------------------
// thread_rng().gen_range(0, 3) for the split axis (optimize.rs:374-375) comes from this seeded
// sequential stream so that scene builds are reproducible.
struct BuildRng {
    uint64_t seed;
    uint32_t n = 0;
    explicit BuildRng(uint64_t s) : seed(s) {}
    uint32_t next_u32() { return philox4x32_10(n++, 0xB7u, (uint32_t)seed, (uint32_t)(seed >> 32), 0x5CE7Eu, 0).v[0]; }
    int axis() { return (int)(((uint64_t)next_u32() * 3u) >> 32); }
};
inline BuildRng*& tls_build_rng() {
    static thread_local BuildRng* p = nullptr;
    return p;
}

struct BoundingVolumeHierarchyNode : Bound {
    AABB volume;
    std::shared_ptr<const Bound> left, right;

    // optimize.rs:366-440.  Returns nullptr where the reference returns None.
    static std::shared_ptr<const BoundingVolumeHierarchyNode> make(std::vector<std::shared_ptr<const Bound>> objects) {
        if (objects.empty()) return nullptr;
        BuildRng* g = tls_build_rng();
        const int axis = g ? g->axis() : 0;
        auto key = [axis](const std::shared_ptr<const Bound>& v) { return (*v->bound()).min[axis]; };  // .unwrap()
        auto node = std::make_shared<BoundingVolumeHierarchyNode>();
        if (objects.size() == 1) {
            node->left = objects[0];
        } else if (objects.size() == 2) {
            if (key(objects[0]) < key(objects[1])) { node->left = objects[0]; node->right = objects[1]; }
            else { node->left = objects[1]; node->right = objects[0]; }
        } else {
            // The reference's comparator never returns Equal (optimize.rs:396-402); a strict-weak `<`
            // with a stable sort gives the same order on distinct keys.  Keys are computed once per
            // object instead of once per comparison (same values).
            std::vector<std::pair<double, std::shared_ptr<const Bound>>> keyed;
            keyed.reserve(objects.size());
            for (auto& o : objects) keyed.emplace_back(key(o), o);
            std::stable_sort(keyed.begin(), keyed.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
            const size_t middle = keyed.size() / 2;
            std::vector<std::shared_ptr<const Bound>> l, r;
            for (size_t i = 0; i < keyed.size(); ++i) (i < middle ? l : r).push_back(keyed[i].second);
            node->right = make(std::move(r));  // right first, as in the reference (optimize.rs:404-408)
            node->left = make(std::move(l));
        }
        std::optional<AABB> lv, rv;
        if (node->left) lv = node->left->bound();
        if (node->right) rv = node->right->bound();
        if (lv && rv) node->volume = lv->merged(*rv);
        else if (lv) node->volume = *lv;
        else if (rv) node->volume = *rv;
        else return nullptr;
        return node;
    }

    std::optional<HitRecord> hit(const Ray& ray) const override {  // optimize.rs:469-498
        if (!volume.hit(ray)) return std::nullopt;
        HitRecord record;  // t = inf
        if (left)
            if (auto l = left->hit(ray))
                if (l->t < record.t) record = *l;
        if (right)
            if (auto r = right->hit(ray))
                if (r->t < record.t) record = *r;
        if (std::isinf(record.t)) return std::nullopt;
        return record;
    }
    std::optional<AABB> bound() const override { return volume; }  // optimize.rs:502-506
};

// ------------------------------------------------------------------------------------------------
// Camera — src/camera.rs
// ------------------------------------------------------------------------------------------------
struct PerspectiveCamera {
    Vec3 eye, center, up;
    double fov, aspect, focusDistance, lensRadius;
    Vec3 lowerLeft, horizontal, vertical;
    PerspectiveCamera() = default;
    PerspectiveCamera(const Vec3& eye_, const Vec3& center_, const Vec3& up_, double fov_, double aspect_,
                      double focusDistance_, double lensRadius_)  // camera.rs:25-59
        : eye(eye_), center(center_), up(up_.normalized()), fov(fov_), aspect(aspect_), focusDistance(focusDistance_),
          lensRadius(lensRadius_) {
        const double height = std::tan(fov / 2.0) * 2.0;
        const double width = aspect * height;
        const Vec3 w = (eye - center).normalized();
        const Vec3 u = up.cross(w);  // NOT normalised (camera.rs:39)
        const Vec3 v = w.cross(u);
        horizontal = u * width * focusDistance;
        vertical = v * height * focusDistance;
        lowerLeft = eye - horizontal / 2.0 - vertical / 2.0 - w * focusDistance;
    }
    Ray ray(double u, double v) const {  // camera.rs:90-107
        if (lensRadius == 0.0) return {eye, (lowerLeft + horizontal * u + vertical * v - eye).normalized()};
        const Vec3 rd = lensRadius * randomInUnitDisk();
        const double offset = rd.x * u + rd.y * v;  // scalar, broadcast over xyz (camera.rs:99-102)
        return {eye + offset, (lowerLeft + horizontal * u + vertical * v - eye - offset).normalized()};
    }
};

// ------------------------------------------------------------------------------------------------
// render::color — src/render.rs:5-29
// ------------------------------------------------------------------------------------------------
inline Vec3 color(const Ray& ray, const Hit& world, size_t maxDepth) {
    if (maxDepth == 0) return {0.0, 0.0, 0.0};
    PathRng* g = tls_rng();
    if (g) { g->bounce++; g->counters.segments++; }  // segment index for the dimensioned draws
    if (auto record = world.hit(ray)) {
        if (const Material* material = record->material) {
            if (auto s = material->scatter(ray, *record))
                return s->second * color(s->first, world, maxDepth - 1) + material->emitted(record->u, record->v, record->intersection);
            return material->emitted(record->u, record->v, record->intersection);
        }
        return {0.0, 0.0, 0.0};
    }
    return {0.0, 0.0, 0.0};  // background is black (render.rs:26)
}

// One camera sample exactly as the example drivers issue it (examples/book-one.rs:69-75):
// u = (x + U)/W, v = (y + U)/H, camera.ray(u, v), color(ray, world, depth).
inline Vec3 samplePixel(const PerspectiveCamera& camera, const Hit& world, uint32_t x, uint32_t y, uint32_t width,
                        uint32_t height, size_t maxDepth, PathRng& g) {
    tls_rng() = &g;
    g.bounce = 0;
    g.seq = 0;
    double ju, jv;
    if (g.mode == PAIRED) {
        const Philox4 r = g.draw(0, PathRng::DIM_CAMERA);
        ju = u24(r.v[0]); jv = u24(r.v[1]);
    } else {
        ju = g.gen_range(0.0, 1.0); jv = g.gen_range(0.0, 1.0);
    }
    const double u = ((double)x + ju) / (double)width;
    const double v = ((double)y + jv) / (double)height;
    const Ray ray = camera.ray(u, v);
    const Vec3 c = color(ray, world, maxDepth);
    tls_rng() = nullptr;
    return c;
}

}  // namespace orc
