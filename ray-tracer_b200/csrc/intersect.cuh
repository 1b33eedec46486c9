// Device intersection routines: ray/AABB slabs, sphere, rectangle, constant medium, and the
// stack-based closest-hit BVH traversal used by the extend and trace kernels.
//
// Semantics follow the reference (quirks in SURVEY.md Appendix A):
//   AxisAlignedBoundingBox::hit   src/optimize.rs:60-83     (tmin = 0, miss iff tmax <= tmin)
//   BoundingVolumeHierarchyNode::hit src/optimize.rs:469-498 (closest hit, strict <, ties -> DFS-earlier)
//   Sphere::hit                   src/geometry.rs:42-74     (first root > 1e-6, outward normal)
//   Rectangle::hit                src/geometry.rs:152-181   (t >= 1e-6, inclusive bounds)
//   Sprite::hit                   src/sprite.rs:94-138      (object-space ray, direction not renormalised)
//   ConstantMedium::hit           src/volume.rs:46-100      (free-flight sampling inside hit())
// The traversal itself is NOT the reference's (which visits both children unconditionally): it is
// ordered, t-pruned and runs over a rebuilt SAH tree; closest-hit results are topology independent.
//
// f32 policy (DESIGN.md "Precision"): the reference relies on f64 and `t > 1e-6` to reject the hit
// at a scattered ray's own origin.  In f32 a hit point on the r = 1000 ground sphere carries ~6e-5 of
// position error, so that test cannot work.  Instead the originating primitive is carried with the
// ray (`self_prim`): a rectangle never re-hits itself; a sphere is skipped when the ray leaves
// outward and answers with its far root (-2 b / a, the c = 0 solution) when the ray goes inward —
// which is exactly what the f64 reference computes for an origin on the surface.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>

#include "device_scene.h"
#include "philox.h"

namespace rtb {

__device__ __forceinline__ float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
__device__ __forceinline__ float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ float3 operator*(float s, float3 a) { return f3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float dot3(float3 a, float3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
__device__ __forceinline__ float3 fma3(float s, float3 a, float3 b) { return f3(fmaf(s, a.x, b.x), fmaf(s, a.y, b.y), fmaf(s, a.z, b.z)); }
__device__ __forceinline__ float3 normalize3(float3 a) {
    const float inv = rsqrtf(dot3(a, a));
    return a * inv;
}
// affine 3x4 row-major
__device__ __forceinline__ float3 xform_point(const float* __restrict__ m, float3 p) {
    return f3(fmaf(m[0], p.x, fmaf(m[1], p.y, fmaf(m[2], p.z, m[3]))), fmaf(m[4], p.x, fmaf(m[5], p.y, fmaf(m[6], p.z, m[7]))),
              fmaf(m[8], p.x, fmaf(m[9], p.y, fmaf(m[10], p.z, m[11]))));
}
__device__ __forceinline__ float3 xform_vector(const float* __restrict__ m, float3 v) {
    return f3(fmaf(m[0], v.x, fmaf(m[1], v.y, m[2] * v.z)), fmaf(m[4], v.x, fmaf(m[5], v.y, m[6] * v.z)),
              fmaf(m[8], v.x, fmaf(m[9], v.y, m[10] * v.z)));
}

constexpr float RT_T_EPS = 1e-6f;  // geometry.rs:57-66,159

// Both roots of |oc + t d|^2 = r^2 (tn <= tf); false when the ray misses.  Discriminant from the
// perpendicular offset l = oc - (b/a) d (no b^2 - a c cancellation), roots from the stable q form.
__device__ __forceinline__ bool sphere_roots(float3 oc, float3 d, float r, float& tn, float& tf) {
    const float a = dot3(d, d);
    const float b = dot3(oc, d);  // half-b
    const float inv_a = 1.0f / a;
    const float3 l = fma3(-b * inv_a, d, oc);
    const float disc = a * fmaf(r, r, -dot3(l, l));  // = b^2 - a c
    if (!(disc >= 0.0f)) return false;
    const float sq = sqrtf(disc);
    const float c = fmaf(-r, r, dot3(oc, oc));
    if (b > 0.0f) {
        const float q = -b - sq;
        tn = q * inv_a;
        tf = c / q;
    } else {
        const float q = sq - b;
        tf = q * inv_a;
        tn = (q != 0.0f) ? c / q : 0.0f;
    }
    if (tn > tf) { const float s = tn; tn = tf; tf = s; }
    return true;
}

// Sphere::hit with the self-origin rule.  Returns t or +inf.
__device__ __forceinline__ float sphere_hit(float3 oc, float3 d, float r, bool is_self) {
    if (is_self) {
        const float b = dot3(oc, d);
        if (b >= 0.0f) return CUDART_INF_F;         // leaving outward: cannot re-hit a convex surface
        const float t = -2.0f * b / dot3(d, d);     // far root with the origin on the surface
        return t > RT_T_EPS ? t : CUDART_INF_F;
    }
    float tn, tf;
    if (!sphere_roots(oc, d, r, tn, tf)) return CUDART_INF_F;
    if (tn > RT_T_EPS) return tn;
    if (tf > RT_T_EPS) return tf;
    return CUDART_INF_F;
}

// Same test in f64 for the primitives of the always-tested list (sky / ground spheres of radius 1e3):
// a ray that starts 1e-3 above the r = 1000 ground and grazes it has c = |oc|^2 - r^2 ~ 2 against an f32
// rounding error of |oc|^2 ~ 0.1, i.e. t would be off by percents.  o - c is exact in f64 for f32 inputs.
__device__ __forceinline__ float sphere_hit_f64(float3 o, float3 ctr, float3 d, float r, bool is_self) {
    const double ox = (double)o.x - (double)ctr.x, oy = (double)o.y - (double)ctr.y, oz = (double)o.z - (double)ctr.z;
    const double dx = d.x, dy = d.y, dz = d.z;
    const double a = dx * dx + dy * dy + dz * dz;
    const double b = ox * dx + oy * dy + oz * dz;  // half-b
    if (is_self) {
        if (b >= 0.0) return CUDART_INF_F;
        const float t = (float)(-2.0 * b / a);
        return t > RT_T_EPS ? t : CUDART_INF_F;
    }
    const double c = ox * ox + oy * oy + oz * oz - (double)r * (double)r;
    const double disc = b * b - a * c;
    if (!(disc >= 0.0)) return CUDART_INF_F;
    const double sq = sqrt(disc);
    const double q = (b > 0.0) ? (-b - sq) : (sq - b);  // stable: no cancellation in q
    double t0 = q / a, t1 = (q != 0.0) ? c / q : 0.0;
    if (t0 > t1) { const double s = t0; t0 = t1; t1 = s; }
    const float tn = (float)t0, tf = (float)t1;
    if (tn > RT_T_EPS) return tn;
    if (tf > RT_T_EPS) return tf;
    return CUDART_INF_F;
}

// Rectangle::hit in object space (plane z = 0).  Returns t or +inf.
__device__ __forceinline__ float rect_hit(float3 o, float3 d, float w, float h) {
    const float t = (0.0f - o.z) / d.z;
    if (!(t >= RT_T_EPS) || isinf(t)) return CUDART_INF_F;  // rejects NaN, inf, t < 1e-6
    const float x = fmaf(d.x, t, o.x), y = fmaf(d.y, t, o.y);
    const float hx = 0.5f * w, hy = 0.5f * h;
    if (x < -hx || x > hx || y < -hy || y > hy) return CUDART_INF_F;
    return t;
}

struct MediumKey {  // Philox key/counter of the free-flight draw: (pixel, sample) or (ray lo, ray hi); segment; seed
    uint32_t k0, k1, segment, seed_lo, seed_hi;
};

// ConstantMedium<Sphere>::hit (volume.rs:46-100).  Returns reported t or +inf; `t1_out`/`entering`
// let shade rebuild the (quirky) reported point.  In the entering branch the reference re-shoots the
// boundary from p1 + 1e-6 d; with a shared direction that second probe's answer is analytically
// t_far - t_near - 1e-6 (and no hit unless that exceeds 1e-6), which is what is computed here.
__device__ __forceinline__ float medium_sphere_hit(float3 o, float3 d, float r, float density, uint32_t order, const MediumKey& key,
                                                   float* t1_out = nullptr, bool* entering_out = nullptr) {
    float tn, tf;
    if (!sphere_roots(o, d, r, tn, tf)) return CUDART_INF_F;
    float t1;
    bool entering;
    if (tn > RT_T_EPS) { t1 = tn; entering = true; }      // outward normal at the near root faces the ray
    else if (tf > RT_T_EPS) { t1 = tf; entering = false; }  // origin inside: normal . d >= 0
    else return CUDART_INF_F;
    // the reference decides the branch by sign(normal . d); for a sphere the near root always has
    // normal . d <= 0 and the far root >= 0, equality only at tangency (handled like the reference: >= 0 -> inside).
    if (entering && !(dot3(fma3(t1, d, o), d) < 0.0f)) entering = false;
    float inside;
    if (entering) {
        inside = (tf - tn) - 1e-6f;
        if (!(inside > RT_T_EPS)) return CUDART_INF_F;
    } else {
        inside = t1;
    }
    const Philox4 rnd = philox4x32_10(key.segment, RT_DIM_MEDIUM | order, key.seed_lo, key.seed_hi, key.k0, key.k1);
    const float distance = (-1.0f / density) * logf(u24f(rnd.x));  // U = 0 -> +inf -> miss (A.9 i)
    if (distance > inside) return CUDART_INF_F;
    if (t1_out) *t1_out = t1;
    if (entering_out) *entering_out = entering;
    return entering ? t1 + distance : distance;
}

// Test device primitive `i`.  Returns t or +inf.  BIG = primitive of the always-tested list (f64 sphere test).
template <bool BIG = false>
__device__ __forceinline__ float prim_hit(const DScene& S, uint32_t i, float3 o, float3 d, uint32_t self_prim, const MediumKey& key) {
    const float4 h = __ldg(&S.prim_head[i]);
    if (h.w > 0.0f) {
        if (BIG) return sphere_hit_f64(o, f3(h.x, h.y, h.z), d, h.w, i == self_prim);
        return sphere_hit(o - f3(h.x, h.y, h.z), d, h.w, i == self_prim);
    }
    const RtPrim* __restrict__ g = &S.gen[__float_as_uint(h.x)];
    const float3 lo = xform_point(g->local_from_world, o);
    const float3 ld = xform_vector(g->local_from_world, d);
    const uint32_t kind = g->kind;
    if (kind == RT_PRIM_RECT) return (i == self_prim) ? CUDART_INF_F : rect_hit(lo, ld, g->param[0], g->param[1]);
    if (kind == RT_PRIM_SPHERE) {
        const float r = g->param[0];
        if (r > 0.0f) return sphere_hit(lo, ld, r, i == self_prim);
        // negative radius (normal flipped by geometry.rs:68): plain roots, no self rule
        float tn, tf;
        if (!sphere_roots(lo, ld, r, tn, tf)) return CUDART_INF_F;
        return tn > RT_T_EPS ? tn : (tf > RT_T_EPS ? tf : CUDART_INF_F);
    }
    return medium_sphere_hit(lo, ld, g->param[0], g->param[1], g->order, key);
}

// ---- BVH traversal -----------------------------------------------------------------------------

struct RayBox {  // per-ray constants of the slab test
    float3 inv, oinv;  // 1/d and o/d
};

// Slab test against [0, tmax]; returns entry distance or +inf.  fminf/fmaxf drop NaNs (0 * inf when the
// origin lies on a slab of a zero direction component), which errs on the side of "hit" like the
// reference's `if t0 > tmin` form.  The far distance is widened by 2 ulp (Ize, robust BVH traversal).
__device__ __forceinline__ float slab(const float4 lo, const float4 hi, const RayBox& rb, float tmax) {
    const float x0 = fmaf(lo.x, rb.inv.x, -rb.oinv.x), x1 = fmaf(hi.x, rb.inv.x, -rb.oinv.x);
    const float y0 = fmaf(lo.y, rb.inv.y, -rb.oinv.y), y1 = fmaf(hi.y, rb.inv.y, -rb.oinv.y);
    const float z0 = fmaf(lo.z, rb.inv.z, -rb.oinv.z), z1 = fmaf(hi.z, rb.inv.z, -rb.oinv.z);
    const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.0f));
    const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), tmax)) * 1.0000004f;
    return tn <= tf ? tn : CUDART_INF_F;
}

struct ClosestHit {
    float t;
    uint32_t prim;  // device primitive index or RT_NO_PRIM
};

// Work counters of the instrumented traversal (RT_RENDER_COUNT_WORK): the algorithmic units of SURVEY §8(d).
struct TraversalWork {
    uint32_t box_tests;   // ray/AABB slab tests (two per visited node pair)
    uint32_t prim_tests;  // primitive intersection tests
};

template <bool BIG = false>
__device__ __forceinline__ void consider(const DScene& S, uint32_t i, float3 o, float3 d, uint32_t self_prim, const MediumKey& key, ClosestHit& best) {
    const float t = prim_hit<BIG>(S, i, o, d, self_prim, key);
    if (t < best.t) { best.t = t; best.prim = i; }
    else if (t == best.t && t < CUDART_INF_F && best.prim != RT_NO_PRIM) {
        // exact tie: the reference keeps the DFS-earlier leaf (optimize.rs:475,484)
        if (__ldg(&S.prim_meta[i]).x < __ldg(&S.prim_meta[best.prim]).x) best.prim = i;
    }
}

// `snodes` = shared-memory copy of nodes[0 .. 2*n_staged) (may be nullptr when n_staged == 0).
__device__ __forceinline__ float4 load_node(const DScene& S, const float4* __restrict__ snodes, uint32_t idx4) {
    return (idx4 < 2u * S.n_staged_nodes) ? snodes[idx4] : __ldg(&S.nodes[idx4]);
}

template <bool COUNT = false>
__device__ __forceinline__ ClosestHit closest_hit(const DScene& S, const float4* __restrict__ snodes, float3 o, float3 d, uint32_t self_prim,
                                                  const MediumKey& key, TraversalWork* work = nullptr) {
    ClosestHit best{CUDART_INF_F, RT_NO_PRIM};
    for (uint32_t i = 0; i < S.n_always; ++i) consider<true>(S, i, o, d, self_prim, key, best);
    if (COUNT) work->prim_tests += S.n_always;
    if (S.n_nodes == 0) return best;

    RayBox rb;
    rb.inv = f3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    rb.oinv = f3(o.x * rb.inv.x, o.y * rb.inv.y, o.z * rb.inv.z);
    // o/d is NaN when both are 0: replace by 0 so that lo*inv - oinv stays +-inf (conservative)
    if (rb.oinv.x != rb.oinv.x) rb.oinv.x = 0.0f;
    if (rb.oinv.y != rb.oinv.y) rb.oinv.y = 0.0f;
    if (rb.oinv.z != rb.oinv.z) rb.oinv.z = 0.0f;

    if (S.root_is_leaf) {
        const float4 lo = load_node(S, snodes, 0), hi = load_node(S, snodes, 1);
        if (COUNT) work->box_tests += 1;
        if (slab(lo, hi, rb, best.t) < CUDART_INF_F) {
            const uint32_t first = __float_as_uint(lo.w), cnt = __float_as_uint(hi.w);
            for (uint32_t i = first; i < first + cnt; ++i) consider(S, i, o, d, self_prim, key, best);
            if (COUNT) work->prim_tests += cnt;
        }
        return best;
    }

    uint32_t stack[RT_STACK_DEPTH];
    int sp = 0;
    uint32_t pair = __float_as_uint(load_node(S, snodes, 0).w);  // left child of the root; right = pair + 1
    for (;;) {
        const float4 llo = load_node(S, snodes, 2u * pair), lhi = load_node(S, snodes, 2u * pair + 1u);
        const float4 rlo = load_node(S, snodes, 2u * pair + 2u), rhi = load_node(S, snodes, 2u * pair + 3u);
        const float tl = slab(llo, lhi, rb, best.t);
        const float tr = slab(rlo, rhi, rb, best.t);
        if (COUNT) work->box_tests += 2;
        // near child first so that the far one can be pruned by what the near one finds
        const bool swap = tr < tl;
        const float t_far = swap ? tl : tr, t_near = swap ? tr : tl;
        const uint32_t near_ref = __float_as_uint(swap ? rlo.w : llo.w), near_cnt = __float_as_uint(swap ? rhi.w : lhi.w);
        const uint32_t far_ref = __float_as_uint(swap ? llo.w : rlo.w), far_cnt = __float_as_uint(swap ? lhi.w : rhi.w);
        uint32_t next_near = RT_NO_PRIM, next_far = RT_NO_PRIM;
        if (t_near < CUDART_INF_F) {
            if (near_cnt) {
                for (uint32_t i = near_ref; i < near_ref + near_cnt; ++i) consider(S, i, o, d, self_prim, key, best);
                if (COUNT) work->prim_tests += near_cnt;
            } else next_near = near_ref;
        }
        if (t_far <= best.t && t_far < CUDART_INF_F) {
            if (far_cnt) {
                for (uint32_t i = far_ref; i < far_ref + far_cnt; ++i) consider(S, i, o, d, self_prim, key, best);
                if (COUNT) work->prim_tests += far_cnt;
            } else next_far = far_ref;
        }
        // interior children: descend into the nearer, push the farther (its children are re-tested against best.t on pop)
        if (next_near != RT_NO_PRIM) {
            if (next_far != RT_NO_PRIM && sp < RT_STACK_DEPTH) stack[sp++] = next_far;
            pair = next_near;
            continue;
        }
        if (next_far != RT_NO_PRIM) { pair = next_far; continue; }
        if (sp == 0) break;
        pair = stack[--sp];
    }
    return best;
}

}  // namespace rtb
