// Scene graph -> flat device description (RtSceneDesc of include/rt_b200.h).
//
// This is the C++ twin of the `flatten` hook that src/gpu.rs adds to the reference's Hit / Bound
// traits: a left-first DFS over BoundingVolumeHierarchyNode / Vec / Sprite /
// TransformedGeometry that pre-composes every transform on the way down
// (Sprite::hit src/sprite.rs:94-138 and TransformedGeometry::hit src/geometry.rs:210-246 become one
// object<-world affine pair per leaf), gives each leaf its DFS rank (`order`, the tie-break of
// src/optimize.rs:475,484) and the material of the outermost enclosing Sprite (src/sprite.rs:124),
// and interns materials / textures into tables.  ImageTexture closures are baked to texels here.
#pragma once
#include <stdexcept>
#include <string>
#include <unordered_map>

#include "../../include/rt_b200.h"
#include "scene_graph.hpp"

namespace rtb {
namespace host {

struct FlattenError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

struct FlatScene {
    std::vector<RtPrim> prims;
    std::vector<RtMaterial> materials;
    std::vector<RtTexture> textures;
    std::vector<float> texels;  // RGBA f32
    std::vector<RtPrim> boundary_prims;  // boundary surfaces of RT_PRIM_MEDIUM primitives (ConstantMedium over anything but a Sphere)
    uint32_t leaves_visited = 0;  // DFS ranks handed out (>= prims.size(): singular transforms are dropped)
    RtSceneDesc desc() const {
        RtSceneDesc d{};
        d.abi_version = RT_ABI_VERSION;
        d.n_prims = (uint32_t)prims.size(); d.n_materials = (uint32_t)materials.size(); d.n_textures = (uint32_t)textures.size();
        d.prims = prims.data(); d.materials = materials.data(); d.textures = textures.data();
        d.texels = texels.data(); d.n_texels = texels.size() / 4;
        d.boundary_prims = boundary_prims.data(); d.n_boundary_prims = (uint32_t)boundary_prims.size();
        return d;
    }
};

struct FlattenCtx {
    FlatScene out;
    std::unordered_map<const Material*, int32_t> materialIds;
    std::unordered_map<const Texture*, int32_t> textureIds;
    bool inBoundary = false;  // flattening the boundary of a ConstantMedium: leaves go to boundary_prims and take no DFS rank

    int32_t internTexture(const Texture* t) {
        if (!t) throw FlattenError("material without texture");
        auto it = textureIds.find(t);
        if (it != textureIds.end()) return it->second;
        RtTexture rec{};
        rec.a = rec.b = -1;
        if (auto s = dynamic_cast<const SolidColor*>(t)) {
            rec.kind = RT_TEX_SOLID;
            rec.rgb[0] = (float)s->color.x; rec.rgb[1] = (float)s->color.y; rec.rgb[2] = (float)s->color.z;
        } else if (auto c = dynamic_cast<const CheckerTexture*>(t)) {
            rec.kind = RT_TEX_CHECKER;
            rec.a = internTexture(c->black.get());
            rec.b = internTexture(c->white.get());
        } else if (auto im = dynamic_cast<const ImageTexture*>(t)) {
            rec.kind = RT_TEX_IMAGE;
            rec.width = im->bakeWidth; rec.height = im->bakeHeight;
            rec.texel_offset = out.texels.size() / 4;
            out.texels.reserve(out.texels.size() + (size_t)rec.width * rec.height * 4);
            for (uint32_t py = 0; py < rec.height; ++py)
                for (uint32_t px = 0; px < rec.width; ++px) {
                    // texel centre; the device looks up px = floor(u W), py = floor((1 - v) H)
                    const double u = ((double)px + 0.5) / (double)rec.width, v = 1.0 - ((double)py + 0.5) / (double)rec.height;
                    const Vec3 c = im->mapping(u, v);
                    out.texels.push_back((float)c.x); out.texels.push_back((float)c.y); out.texels.push_back((float)c.z); out.texels.push_back(1.0f);
                }
        } else {
            throw FlattenError("unsupported Texture implementation (no flatten hook)");
        }
        const int32_t id = (int32_t)out.textures.size();
        out.textures.push_back(rec);
        textureIds.emplace(t, id);
        return id;
    }

    int32_t internMaterial(const Material* m) {
        if (!m) return -1;
        auto it = materialIds.find(m);
        if (it != materialIds.end()) return it->second;
        RtMaterial rec{};
        rec.kind = (uint32_t)m->kind();
        rec.texture = (m->kind() == Material::DIELECTRIC) ? -1 : internTexture(m->texture().get());
        rec.fuzz = (float)m->fuzziness();
        rec.ior = (float)m->refractive();
        const int32_t id = (int32_t)out.materials.size();
        out.materials.push_back(rec);
        materialIds.emplace(m, id);
        return id;
    }

    static void affine(const Mat4& m, float* out12) {
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 4; ++c) out12[4 * r + c] = (float)m.at(r, c);
    }

    void leaf(uint32_t kind, const FlattenState& st, double p0, double p1) {
        const auto inv = st.worldFromLocal.inversed();
        if (inBoundary) {  // a surface of a medium's boundary: never hit on its own, carries no material and no rank
            if (!inv) return;
            RtPrim p{};
            p.kind = kind;
            p.material = -1;
            p.param[0] = (float)p0; p.param[1] = (float)p1;
            affine(*inv, p.local_from_world);
            affine(st.worldFromLocal, p.world_from_local);
            out.boundary_prims.push_back(p);
            return;
        }
        const uint32_t order = out.leaves_visited++;
        if (!inv) return;  // det == 0: the sprite is never hit (src/sprite.rs:131-134); the rank stays reserved
        RtPrim p{};
        p.kind = kind;
        p.material = internMaterial(st.material);
        p.order = order;
        p.param[0] = (float)p0; p.param[1] = (float)p1;
        affine(*inv, p.local_from_world);
        affine(st.worldFromLocal, p.world_from_local);
        out.prims.push_back(p);
    }

    // ConstantMedium<T> over a boundary that is not a plain Sphere (src/volume.rs:40-43 accepts any Hit): the boundary subtree is
    // flattened into `boundary_prims` (world-space transforms, DFS = list order) and ONE primitive stands for the medium.
    void mediumOver(const Bound& boundary, const FlattenState& st, double density) {
        if (inBoundary) throw FlattenError("ConstantMedium inside the boundary of a ConstantMedium has no device implementation");
        const uint32_t order = out.leaves_visited++;
        const uint32_t first = (uint32_t)out.boundary_prims.size();
        inBoundary = true;
        try { boundary.flatten(*this, st); } catch (...) { inBoundary = false; throw; }
        inBoundary = false;
        const uint32_t count = (uint32_t)out.boundary_prims.size() - first;
        if (count == 0) return;  // boundary never hit: neither is the medium
        if (count >= (1u << 24)) throw FlattenError("ConstantMedium boundary with 2^24 or more surfaces");
        RtPrim p{};
        p.kind = RT_PRIM_MEDIUM;
        p.material = internMaterial(st.material);
        p.order = order;
        p.aux = first;
        p.param[1] = (float)density;
        p.param[2] = (float)count;
        affine(Mat4::identity(), p.local_from_world);  // the boundary surfaces carry the transforms
        affine(Mat4::identity(), p.world_from_local);
        out.prims.push_back(p);
    }
};

inline void Sphere::flatten(FlattenCtx& ctx, const FlattenState& st) const { ctx.leaf(RT_PRIM_SPHERE, st, radius, 0.0); }
inline void Rectangle::flatten(FlattenCtx& ctx, const FlattenState& st) const { ctx.leaf(RT_PRIM_RECT, st, width, height); }
inline void TransformedGeometry::flatten(FlattenCtx& ctx, const FlattenState& st) const {
    FlattenState inner = st;
    inner.worldFromLocal = st.worldFromLocal.multiplied(transform.origin);
    geometry->flatten(ctx, inner);
}
inline void Sprite::flatten(FlattenCtx& ctx, const FlattenState& st) const {
    if (!geometry_) return;
    FlattenState inner = st;
    inner.worldFromLocal = st.worldFromLocal.multiplied(transform_.origin);
    if (!st.inSprite) inner.material = material_.get();  // the outermost Sprite's material wins, None included
    inner.inSprite = true;
    geometry_->flatten(ctx, inner);
}
inline void ConstantMedium::flatten(FlattenCtx& ctx, const FlattenState& st) const {
    if (ctx.inBoundary) throw FlattenError("ConstantMedium inside the boundary of a ConstantMedium has no device implementation");
    if (auto s = dynamic_cast<const Sphere*>(boundary.get())) ctx.leaf(RT_PRIM_MEDIUM_SPHERE, st, s->radius, density);  // analytic two-root form
    else ctx.mediumOver(*boundary, st, density);
}

inline FlatScene flattenWorld(const Bound& world) {
    FlattenCtx ctx;
    world.flatten(ctx, FlattenState{});
    return std::move(ctx.out);
}

inline RtCamera flattenCamera(const PerspectiveCamera& c) {
    RtCamera o{};
    const Vec3 v[4] = {c.eye, c.lowerLeft, c.horizontal, c.vertical};
    float* dst[4] = {o.eye, o.lower_left, o.horizontal, o.vertical};
    for (int i = 0; i < 4; ++i) { dst[i][0] = (float)v[i].x; dst[i][1] = (float)v[i].y; dst[i][2] = (float)v[i].z; }
    o.lens_radius = (float)c.lensRadius;
    return o;
}

}  // namespace host
}  // namespace rtb
