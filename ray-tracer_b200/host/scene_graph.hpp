// Host-side scene-description API — the C++ mirror of the reference's Rust API that stays unchanged
// for users of the GPU backend (BASELINE north_star): Sprite::builder, the Hit / Bound traits,
// BoundingVolumeHierarchyNode, Sphere / Rectangle / Cube / TransformedGeometry / ConstantMedium, the
// material and texture types, PerspectiveCamera.
//
//   trait Hit / Bound            src/ray.rs:85-87, src/optimize.rs:88-95
//   Sphere / Rectangle / Cube    src/geometry.rs:11-33,126-150,250-287
//   TransformedGeometry          src/geometry.rs:184-208
//   Sprite / SpriteBuilder       src/sprite.rs:10-85
//   ConstantMedium               src/volume.rs:17-38
//   BoundingVolumeHierarchyNode  src/optimize.rs:336-461 (random axis, sort by box min, median split)
//   materials / textures         src/material.rs:23-58,72-96,121-138,200-271,273-315
//   PerspectiveCamera            src/camera.rs:9-88
//
// What is different from the reference: nodes do not intersect rays on the CPU (there is no CPU
// path in this backend).  Instead every node implements the flatten hook `flatten()` that
// src/gpu.rs adds to the Rust traits: it walks the graph in the reference's
// left-first DFS order and emits device primitives with pre-composed transforms (flatten.hpp).
// `bound()` is kept bit-compatible with the reference (it decides the BVH topology and therefore
// the DFS order that breaks closest-hit ties).
#pragma once
#include <algorithm>
#include <cstdint>
#include <functional>
#include <memory>
#include <optional>
#include <utility>
#include <vector>

#include "../csrc/philox.h"
#include "math.hpp"

namespace rtb {
namespace host {

struct FlattenCtx;  // flatten.hpp
struct Material;

struct AxisAlignedBoundingBox {  // src/optimize.rs:20-58
    Vec3 min, max;
    AxisAlignedBoundingBox merged(const AxisAlignedBoundingBox& o) const {
        return {{std::min(min.x, o.min.x), std::min(min.y, o.min.y), std::min(min.z, o.min.z)},
                {std::max(max.x, o.max.x), std::max(max.y, o.max.y), std::max(max.z, o.max.z)}};
    }
};
using AABB = AxisAlignedBoundingBox;

// Box of the 8 corners of `b` pushed through `m` (Bound for Sprite / TransformedGeometry, src/optimize.rs:128-241).
inline AABB transformedBound(const AABB& b, const Mat4& m) {
    AABB out{{INFINITY, INFINITY, INFINITY}, {-INFINITY, -INFINITY, -INFINITY}};
    for (int corner = 0; corner < 8; ++corner) {
        const Vec3 c((corner & 1) ? b.max.x : b.min.x, (corner & 2) ? b.max.y : b.min.y, (corner & 4) ? b.max.z : b.min.z);
        const Vec3 p = m.transformPoint(c);
        out.min = {std::min(out.min.x, p.x), std::min(out.min.y, p.y), std::min(out.min.z, p.z)};
        out.max = {std::max(out.max.x, p.x), std::max(out.max.y, p.y), std::max(out.max.z, p.z)};
    }
    return out;
}

// State handed down the graph while flattening: composed transform and the material of the
// OUTERMOST enclosing Sprite (Sprite::hit overwrites the inner record's material, src/sprite.rs:124).
struct FlattenState {
    Mat4 worldFromLocal = Mat4::identity();
    const Material* material = nullptr;
    bool inSprite = false;
};

struct Hit {  // src/ray.rs:85-87 — `hit()` is executed by the GPU (gpu.hpp: World::trace_rays / render_image)
    virtual ~Hit() = default;
    virtual void flatten(FlattenCtx& ctx, const FlattenState& st) const = 0;
};
struct Bound : Hit {  // Bound<AxisAlignedBoundingBox>, src/optimize.rs:88-95
    virtual std::optional<AABB> bound() const = 0;
};
using BoundPtr = std::shared_ptr<const Bound>;

// ---- textures / materials (description only; evaluated on the device) ------------------------------
struct Texture {
    virtual ~Texture() = default;
};
using TexturePtr = std::shared_ptr<const Texture>;
struct SolidColor : Texture {
    Vec3 color;
    explicit SolidColor(const Vec3& c) : color(c) {}
};
struct CheckerTexture : Texture {
    TexturePtr black, white;
    CheckerTexture(TexturePtr b, TexturePtr w) : black(std::move(b)), white(std::move(w)) {}
};
// ImageTexture<F>: an arbitrary closure of (u, v) (src/material.rs:247-265).  A closure cannot run on
// the GPU, so it is baked on a bakeWidth x bakeHeight grid at texel centres when the scene is
// flattened; with the grid equal to the source image this reproduces the example's nearest lookup
// (examples/main.rs:267-280) exactly.
struct ImageTexture : Texture {
    std::function<Vec3(double, double)> mapping;
    uint32_t bakeWidth, bakeHeight;
    explicit ImageTexture(std::function<Vec3(double, double)> m, uint32_t w = 1024, uint32_t h = 512)
        : mapping(std::move(m)), bakeWidth(w), bakeHeight(h) {}
};
inline TexturePtr solid(const Vec3& c) { return std::make_shared<SolidColor>(c); }

struct Material {
    enum Kind { LAMBERTIAN = 0, METAL = 1, DIELECTRIC = 2, DIFFUSE_LIGHT = 3, ISOTROPIC = 4 };
    virtual ~Material() = default;
    virtual Kind kind() const = 0;
    virtual TexturePtr texture() const { return nullptr; }
    virtual double fuzziness() const { return 0.0; }
    virtual double refractive() const { return 1.0; }
};
using MaterialPtr = std::shared_ptr<const Material>;
struct Lambertian : Material {
    TexturePtr albedo;
    explicit Lambertian(TexturePtr a) : albedo(std::move(a)) {}
    explicit Lambertian(const Vec3& c) : albedo(solid(c)) {}
    Kind kind() const override { return LAMBERTIAN; }
    TexturePtr texture() const override { return albedo; }
};
struct Metal : Material {
    TexturePtr albedo;
    double fuzz;
    Metal(TexturePtr a, double f) : albedo(std::move(a)), fuzz(f) {}
    Metal(const Vec3& c, double f) : albedo(solid(c)), fuzz(f) {}
    Kind kind() const override { return METAL; }
    TexturePtr texture() const override { return albedo; }
    double fuzziness() const override { return fuzz; }
};
struct Dielectric : Material {
    double ior;
    explicit Dielectric(double r) : ior(r) {}
    Kind kind() const override { return DIELECTRIC; }
    double refractive() const override { return ior; }
};
struct DiffuseLight : Material {
    TexturePtr emission;
    explicit DiffuseLight(TexturePtr e) : emission(std::move(e)) {}
    explicit DiffuseLight(const Vec3& c) : emission(solid(c)) {}
    Kind kind() const override { return DIFFUSE_LIGHT; }
    TexturePtr texture() const override { return emission; }
};
struct Isotropic : Material {
    TexturePtr albedo;
    explicit Isotropic(TexturePtr a) : albedo(std::move(a)) {}
    explicit Isotropic(const Vec3& c) : albedo(solid(c)) {}
    Kind kind() const override { return ISOTROPIC; }
    TexturePtr texture() const override { return albedo; }
};

// ---- geometry --------------------------------------------------------------------------------------
struct Sphere : Bound {
    double radius;
    explicit Sphere(double r) : radius(r) {}
    void flatten(FlattenCtx& ctx, const FlattenState& st) const override;
    std::optional<AABB> bound() const override { return AABB{Vec3(0, 0, 0) - Vec3(radius, radius, radius), Vec3(0, 0, 0) + Vec3(radius, radius, radius)}; }
};
struct Rectangle : Bound {
    double width, height;
    Rectangle(double w, double h) : width(w), height(h) {}
    void flatten(FlattenCtx& ctx, const FlattenState& st) const override;
    std::optional<AABB> bound() const override { return AABB{{-width / 2.0, -height / 2.0, 0.0 - 1e-6}, {width / 2.0, height / 2.0, 0.0 + 1e-6}}; }
};
struct TransformedGeometry : Bound {
    BoundPtr geometry;
    Mat4Cached transform;
    TransformedGeometry(BoundPtr g, const Mat4Cached& t) : geometry(std::move(g)), transform(t) {}
    void flatten(FlattenCtx& ctx, const FlattenState& st) const override;
    std::optional<AABB> bound() const override {
        auto b = geometry->bound();
        if (!b) return std::nullopt;
        return transformedBound(*b, transform.origin);
    }
};
struct Cube {  // six outward-facing rectangles: front, left, back, right, top, bottom (src/geometry.rs:250-287)
    static std::vector<BoundPtr> make(double width, double height, double depth) {
        const double quarter = 90.0 * 3.14159265358979323846 / 180.0, half = 180.0 * 3.14159265358979323846 / 180.0;
        const Vec3 ex(1, 0, 0), ey(0, 1, 0);
        struct Face { double w, h; Mat4 m; };
        const Face faces[6] = {
            {width, height, Mat4::translation({0.0, 0.0, depth / 2.0})},
            {depth, height, Mat4::translation({-width / 2.0, 0.0, 0.0}).multiplied(Mat4::rotation(-quarter, ey))},
            {width, height, Mat4::translation({0.0, 0.0, -depth / 2.0}).multiplied(Mat4::rotation(half, ey))},
            {depth, height, Mat4::translation({width / 2.0, 0.0, 0.0}).multiplied(Mat4::rotation(quarter, ey))},
            {width, depth, Mat4::translation({0.0, height / 2.0, 0.0}).multiplied(Mat4::rotation(-quarter, ex))},
            {width, depth, Mat4::translation({0.0, -height / 2.0, 0.0}).multiplied(Mat4::rotation(quarter, ex))},
        };
        std::vector<BoundPtr> out;
        for (const Face& f : faces) out.push_back(std::make_shared<TransformedGeometry>(std::make_shared<Rectangle>(f.w, f.h), Mat4Cached(f.m)));
        return out;
    }
};
struct HitList : Bound {  // impl Hit / Bound for Vec<..> (src/geometry.rs:76-116, src/optimize.rs:256-334)
    std::vector<BoundPtr> items;
    void flatten(FlattenCtx& ctx, const FlattenState& st) const override { for (const auto& v : items) v->flatten(ctx, st); }
    std::optional<AABB> bound() const override {
        std::optional<AABB> res;
        for (const auto& v : items)
            if (auto o = v->bound()) res = res ? res->merged(*o) : *o;
        return res;
    }
};
struct ConstantMedium : Bound {
    BoundPtr boundary;
    double density;
    ConstantMedium(BoundPtr b, double d) : boundary(std::move(b)), density(d) {}
    void flatten(FlattenCtx& ctx, const FlattenState& st) const override;
    std::optional<AABB> bound() const override { return boundary->bound(); }
};

// ---- Sprite ----------------------------------------------------------------------------------------
struct Sprite : Bound {
    BoundPtr geometry_;
    MaterialPtr material_;
    Mat4Cached transform_;
    struct Builder {
        std::shared_ptr<Sprite> s = std::make_shared<Sprite>();
        Builder& geometry(BoundPtr g) { s->geometry_ = std::move(g); return *this; }
        Builder& material(MaterialPtr m) { s->material_ = std::move(m); return *this; }
        Builder& transform(const Mat4Cached& m) { s->transform_ = m; return *this; }
        BoundPtr build() { return s; }
    };
    static Builder builder() { return {}; }
    void flatten(FlattenCtx& ctx, const FlattenState& st) const override;
    std::optional<AABB> bound() const override {
        if (!geometry_) return std::nullopt;
        auto b = geometry_->bound();
        if (!b) return std::nullopt;
        return transformedBound(*b, transform_.origin);
    }
};

// ---- BoundingVolumeHierarchyNode ---------------------------------------------------------------------
// The reference draws the split axis from thread_rng (src/optimize.rs:374-375); here it comes from a
// seeded counter-based stream so that scene builds are reproducible.
struct BuildAxisStream {
    uint64_t seed;
    uint32_t n = 0;
    explicit BuildAxisStream(uint64_t s) : seed(s) {}
    int axis() {
        const uint32_t r = philox4x32_10(n++, 0xB7u, (uint32_t)seed, (uint32_t)(seed >> 32), 0x5CE7Eu, 0u).x;
        return (int)(((uint64_t)r * 3u) >> 32);
    }
};
inline BuildAxisStream*& currentAxisStream() {
    static thread_local BuildAxisStream* p = nullptr;
    return p;
}

struct BoundingVolumeHierarchyNode : Bound {
    AABB volume;
    BoundPtr left, right;
    // BoundingVolumeHierarchyNode::new; nullptr where the reference returns None.
    static std::shared_ptr<const BoundingVolumeHierarchyNode> make(std::vector<BoundPtr> objects) {
        if (objects.empty()) return nullptr;
        BuildAxisStream* g = currentAxisStream();
        const int axis = g ? g->axis() : 0;
        auto node = std::make_shared<BoundingVolumeHierarchyNode>();
        const size_t n = objects.size();
        if (n == 1) {
            node->left = objects[0];
        } else {
            std::vector<std::pair<double, BoundPtr>> keyed;
            keyed.reserve(n);
            for (auto& o : objects) keyed.emplace_back((*o->bound()).min[axis], o);
            if (n == 2) {
                const bool firstIsLess = keyed[0].first < keyed[1].first;
                node->left = keyed[firstIsLess ? 0 : 1].second;
                node->right = keyed[firstIsLess ? 1 : 0].second;
            } else {
                std::stable_sort(keyed.begin(), keyed.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
                const size_t middle = n / 2;
                std::vector<BoundPtr> lower, upper;
                for (size_t i = 0; i < n; ++i) (i < middle ? lower : upper).push_back(keyed[i].second);
                node->right = make(std::move(upper));  // right subtree first: it consumes the axis stream first (optimize.rs:404-408)
                node->left = make(std::move(lower));
            }
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
    void flatten(FlattenCtx& ctx, const FlattenState& st) const override {
        if (left) left->flatten(ctx, st);
        if (right) right->flatten(ctx, st);
    }
    std::optional<AABB> bound() const override { return volume; }
};

// ---- camera ----------------------------------------------------------------------------------------
struct PerspectiveCamera {  // src/camera.rs:9-88
    Vec3 eye, center, up;
    double fov, aspect, focusDistance, lensRadius;
    Vec3 lowerLeft, horizontal, vertical;
    PerspectiveCamera(const Vec3& eye_, const Vec3& center_, const Vec3& up_, double fov_, double aspect_, double focusDistance_, double lensRadius_)
        : eye(eye_), center(center_), up(up_.normalized()), fov(fov_), aspect(aspect_), focusDistance(focusDistance_), lensRadius(lensRadius_) {
        const double viewHeight = std::tan(fov / 2.0) * 2.0, viewWidth = aspect * viewHeight;
        const Vec3 w = (eye - center).normalized();
        const Vec3 u = up.cross(w);  // deliberately not normalised: the reference does not (camera.rs:39)
        const Vec3 v = w.cross(u);
        horizontal = u * viewWidth * focusDistance;
        vertical = v * viewHeight * focusDistance;
        lowerLeft = eye - horizontal / 2.0 - vertical / 2.0 - w * focusDistance;
    }
};

}  // namespace host
}  // namespace rtb
