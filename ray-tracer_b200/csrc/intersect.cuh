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
//   Hit for Vec<...>              src/geometry.rs:76-116    (linear list, strict <)
// The traversal itself is NOT the reference's (which visits both children unconditionally): it is
// ordered, t-pruned and runs over a rebuilt SAH tree; closest-hit results are topology independent.
// Scenes of at most RT_SMALL_SCENE primitives skip the tree: every lane of a warp walks the same
// shared-memory primitive list in DFS order (the reference's own linear-list Hit), which costs fewer
// instructions than any traversal and keeps all 32 lanes converged.
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
__device__ __forceinline__ float dot3(float4 a, float3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
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
// 1-ulp-class hardware approximations (MUFU.RCP / MUFU.SQRT + one multiply); the precise sequences cost ~8
// instructions each and sit in the innermost loops.  Errors of a few ulp in t are far inside the 1e-4 budget.
#ifdef RT_PRECISE_RCP
__device__ __forceinline__ float fast_rcp(float x) { return 1.0f / x; }
#else
__device__ __forceinline__ float fast_rcp(float x) { return __fdividef(1.0f, x); }
#endif
__device__ __forceinline__ float fast_sqrt(float x) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

constexpr float RT_T_EPS = 1e-6f;  // geometry.rs:57-66,159

// Per-ray constants shared by every primitive / box test.
struct RayCtx {
    float3 o, d;
    float a, inv_a;    // d.d and its reciprocal
    float3 inv, oinv;  // 1/d and o/d for the slab test
};

// Both roots of |oc + t d|^2 = r^2 (tn <= tf); false when the ray misses.  Discriminant from the
// perpendicular offset l = oc - (b/a) d (no b^2 - a c cancellation), roots from the stable q form.
__device__ __forceinline__ bool sphere_roots(float3 oc, float3 d, float a, float inv_a, float r, float& tn, float& tf) {
    const float b = dot3(oc, d);  // half-b
    const float3 l = fma3(-b * inv_a, d, oc);
    const float disc = a * fmaf(r, r, -dot3(l, l));  // = b^2 - a c
    if (!(disc >= 0.0f)) return false;
    const float sq = fast_sqrt(disc);
    const float c = fmaf(-r, r, dot3(oc, oc));
    const float q = (b > 0.0f) ? (-b - sq) : (sq - b);
    const float t0 = q * inv_a;
    const float t1 = (q != 0.0f) ? __fdividef(c, q) : 0.0f;
    tn = fminf(t0, t1);
    tf = fmaxf(t0, t1);
    return true;
}
__device__ __forceinline__ bool sphere_roots(float3 oc, float3 d, float r, float& tn, float& tf) {
    const float a = dot3(d, d);
    return sphere_roots(oc, d, a, 1.0f / a, r, tn, tf);
}

// Sphere::hit with the self-origin rule.  Returns t or +inf.
__device__ __forceinline__ float sphere_hit(const RayCtx& R, float3 ctr, float r, bool is_self) {
    const float3 oc = R.o - ctr;
    if (is_self) {
        const float b = dot3(oc, R.d);
        if (b >= 0.0f) return CUDART_INF_F;    // leaving outward: cannot re-hit a convex surface
        const float t = -2.0f * b * R.inv_a;   // far root with the origin on the surface
        return t > RT_T_EPS ? t : CUDART_INF_F;
    }
    float tn, tf;
    if (!sphere_roots(oc, R.d, R.a, R.inv_a, r, tn, tf)) return CUDART_INF_F;
    if (tn > RT_T_EPS) return tn;
    if (tf > RT_T_EPS) return tf;
    return CUDART_INF_F;
}

// The same test for the huge spheres of the always-tested list (ground r = 1000, sky r = 2000, fog r = 5000),
// expanded around the WORLD origin: c = |o|^2 - 2 o.ctr + K with K = |ctr|^2 - r^2 rounded once from f64 on the
// host, b = o.d - ctr.d.  For a ray 1e-3 above the r = 1000 ground the centre-relative form has c ~ 2 with an f32
// error of 2 |oc| ulp(|oc|) ~ 0.1 (t off by percents); this form has an error of ~ulp(2 o.ctr) ~ 1e-5.
__device__ __forceinline__ float big_sphere_hit(const RayCtx& R, float3 ctr, float r, float K, bool is_self) {
    const float b = dot3(R.o, R.d) - dot3(ctr, R.d);
    if (is_self) {
        if (b >= 0.0f) return CUDART_INF_F;
        const float t = -2.0f * b * R.inv_a;
        return t > RT_T_EPS ? t : CUDART_INF_F;
    }
    const float c = fmaf(-2.0f, dot3(R.o, ctr), dot3(R.o, R.o)) + K;
    const float disc = fmaf(b, b, -R.a * c);
    if (!(disc >= 0.0f)) return CUDART_INF_F;
    const float sq = fast_sqrt(disc);
    const float q = (b > 0.0f) ? (-b - sq) : (sq - b);
    const float t0 = q * R.inv_a;
    const float t1 = (q != 0.0f) ? __fdividef(c, q) : 0.0f;
    const float tn = fminf(t0, t1), tf = fmaxf(t0, t1);
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

// The same test on the compact 64-byte record the extend kernel reads (device_scene.h "rects"): rows of the
// object<-world matrix, so t = -(row_z . o + tz) / (row_z . d) is the reference's (0 - o'.z) / d'.z, and the
// object-space x, y come from the world-space hit point.
__device__ __forceinline__ float rect_hit(const RayCtx& R, const float4 rz, const float4 rx, const float4 ry, const float4 ext) {
    const float den = dot3(rz, R.d);
    const float num = -(dot3(rz, R.o) + rz.w);
    const float t = __fdividef(num, den);
    if (!(t >= RT_T_EPS) || !(t < CUDART_INF_F)) return CUDART_INF_F;  // rejects NaN, inf, t < 1e-6
    const float3 p = fma3(t, R.d, R.o);
    const float x = dot3(rx, p) + rx.w, y = dot3(ry, p) + ry.w;
    if (fabsf(x) > ext.x || fabsf(y) > ext.y) return CUDART_INF_F;
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

// ---- ConstantMedium<T> over an arbitrary flattened boundary (RT_PRIM_MEDIUM; volume.rs:40-100) -------------------------------
// Both probes of the reference (`boundary.hit(ray)` at volume.rs:47 and, from p1 + 1e-6 d, at volume.rs:54) are closest-hit
// queries over the boundary's surfaces in DFS order with strict `<`.  f32 policy as everywhere: the second probe is evaluated
// on the ORIGINAL ray (no re-rounded origin), measured from t1 + 1e-6; the surface the first probe hit is excluded when it is a
// rectangle (a plane cannot be re-hit) and answers with its far root when it is a sphere — what the f64 reference gets from
// `t >= 1e-6` on the re-shot ray.
struct BoundaryHit {
    float t;
    uint32_t idx;  // index into S.bnd, RT_NO_PRIM = miss
};
// closest boundary surface with t_ray2 = t - t_from >= eps, where ray2 starts at parameter t_from of the ray (0 for the first probe)
__device__ __forceinline__ BoundaryHit boundary_probe(const RtPrim* __restrict__ bnd, uint32_t first, uint32_t count, float3 o, float3 d, float t_from,
                                                      uint32_t skip) {
    BoundaryHit best{CUDART_INF_F, RT_NO_PRIM};
    for (uint32_t i = first; i < first + count; ++i) {
        const RtPrim* __restrict__ g = bnd + i;
        const float3 lo = xform_point(g->local_from_world, o);
        const float3 ld = xform_vector(g->local_from_world, d);
        float t = CUDART_INF_F;
        if (g->kind == RT_PRIM_RECT) {
            if (i != skip) {
                const float tt = rect_hit(lo, ld, g->param[0], g->param[1]);  // rejects t < 1e-6 on the original ray
                if (tt - t_from >= RT_T_EPS) t = tt;
            }
        } else {  // sphere
            float tn, tf;
            if (sphere_roots(lo, ld, g->param[0], tn, tf)) {
                if (i == skip) { if (tf - t_from > RT_T_EPS && tf > tn) t = tf; }  // far root of the surface the ray2 starts on
                else if (tn - t_from > RT_T_EPS) t = tn;
                else if (tf - t_from > RT_T_EPS) t = tf;
            }
        }
        if (t < best.t) { best.t = t; best.idx = i; }
    }
    return best;
}
// outward normal (not renormalised after the forward matrix, sprite.rs:112-115) and uv of boundary surface `idx` at ray parameter t
__device__ __forceinline__ float3 boundary_normal(const RtPrim* __restrict__ bnd, uint32_t idx, float3 o, float3 d, float t, float* u = nullptr, float* v = nullptr) {
    const RtPrim* __restrict__ g = bnd + idx;
    if (g->kind == RT_PRIM_RECT) {
        if (u) {
            const float3 lp = fma3(t, xform_vector(g->local_from_world, d), xform_point(g->local_from_world, o));
            *u = (lp.x + 0.5f * g->param[0]) / g->param[0];
            *v = (lp.y + 0.5f * g->param[1]) / g->param[1];
        }
        return f3(g->world_from_local[2], g->world_from_local[6], g->world_from_local[10]);
    }
    const float3 lp = fma3(t, xform_vector(g->local_from_world, d), xform_point(g->local_from_world, o));
    const float3 q = lp * (1.0f / g->param[0]);
    const float3 unit = normalize3(q);
    if (u) {  // geometry.rs:35-39, acos clamped (A.11)
        *u = 0.5f + atan2f(q.x, q.z) * (1.0f / (2.0f * 3.14159265358979323846f));
        *v = 1.0f - acosf(fminf(fmaxf(q.y, -1.0f), 1.0f)) * (1.0f / 3.14159265358979323846f);
    }
    return xform_vector(g->world_from_local, unit);
}
struct MediumProbe {  // what shade needs to rebuild the record
    float t1, t2;     // first boundary hit; second (entering branch only), both as parameters of the original ray
    uint32_t i1, i2;
    bool entering;
};
__device__ __forceinline__ float medium_generic_hit(const RtPrim* __restrict__ bnd, uint32_t first, uint32_t count, float3 o, float3 d, float density,
                                                    uint32_t order, const MediumKey& key, MediumProbe* probe = nullptr) {
    const BoundaryHit h1 = boundary_probe(bnd, first, count, o, d, 0.0f, RT_NO_PRIM);
    if (h1.idx == RT_NO_PRIM) return CUDART_INF_F;
    const float3 n1 = boundary_normal(bnd, h1.idx, o, d, h1.t);
    const bool entering = dot3(n1, d) < 0.0f;  // volume.rs:49
    float inside, t2 = 0.0f;
    uint32_t i2 = RT_NO_PRIM;
    if (entering) {
        const BoundaryHit h2 = boundary_probe(bnd, first, count, o, d, h1.t + 1e-6f, h1.idx);
        if (h2.idx == RT_NO_PRIM) return CUDART_INF_F;  // volume.rs:74-76
        inside = (h2.t - h1.t) - 1e-6f;
        t2 = h2.t; i2 = h2.idx;
    } else {
        inside = h1.t;
    }
    const Philox4 rnd = philox4x32_10(key.segment, RT_DIM_MEDIUM | order, key.seed_lo, key.seed_hi, key.k0, key.k1);
    const float distance = (-1.0f / density) * logf(u24f(rnd.x));
    if (distance > inside) return CUDART_INF_F;
    if (probe) { probe->t1 = h1.t; probe->t2 = t2; probe->i1 = h1.idx; probe->i2 = i2; probe->entering = entering; }
    return entering ? h1.t + distance : distance;
}

// Primitive of a 128-byte general record (rotated / scaled sphere, negative radius, medium).  Returns t or +inf.
// Kept out of line: it is the rare path and would otherwise triple the size of the traversal loop.
#ifndef RT_GENERAL_INLINE
#define RT_GENERAL_INLINE __noinline__
#endif
// (The boundary surfaces of an RT_PRIM_MEDIUM record live in the same table, `aux` records AFTER the record itself — see rt_scene_create —
// so that this out-of-line function needs no second table pointer: an extra argument here cost k_extend<., true> 24 bytes of spills.)
__device__ RT_GENERAL_INLINE float general_hit(const RtPrim* __restrict__ g, float3 o, float3 d, bool is_self, const MediumKey key) {
    const uint32_t kind = g->kind;
    if (kind == RT_PRIM_MEDIUM) return medium_generic_hit(g + g->aux, 0u, (uint32_t)g->param[2], o, d, g->param[1], g->order, key);
    const float3 lo = xform_point(g->local_from_world, o);
    const float3 ld = xform_vector(g->local_from_world, d);
    if (kind == RT_PRIM_RECT) return is_self ? CUDART_INF_F : rect_hit(lo, ld, g->param[0], g->param[1]);
    if (kind == RT_PRIM_SPHERE) {
        const float r = g->param[0];
        float tn, tf;
        if (r > 0.0f && is_self) {
            const float b = dot3(lo, ld);
            if (b >= 0.0f) return CUDART_INF_F;
            const float t = -2.0f * b / dot3(ld, ld);
            return t > RT_T_EPS ? t : CUDART_INF_F;
        }
        // (a negative radius flips the normal, geometry.rs:68: plain roots, no self rule)
        if (!sphere_roots(lo, ld, r, tn, tf)) return CUDART_INF_F;
        return tn > RT_T_EPS ? tn : (tf > RT_T_EPS ? tf : CUDART_INF_F);
    }
    return medium_sphere_hit(lo, ld, g->param[0], g->param[1], g->order, key);
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

// Shared-memory staging of the scene prefix (see k_extend): `nodes` = copy of S.nodes[0 .. 2*n_staged_nodes),
// `heads` / `rects` = copies of the primitive tables when the whole scene is small enough (S.n_staged_prims).
// All four are 32-bit SHARED-WINDOW byte addresses, read with explicit ld.shared / st.shared: handing the compiler
// generic pointers made it merge the shared and global paths into address-space-agnostic LD (ncu/SASS: 16 generic LD,
// no LDS, no LDG.CONSTANT) and cost 25-40 % on scenes whose nodes live in global memory.
struct Staged {
    uint32_t nodes;
    uint32_t heads;
    uint32_t rects;
    uint32_t boxes;  // (ALL only; otherwise the box records are read from global memory)
    uint32_t stack;  // this thread's column of the shared-memory traversal stack: entry k at stack + k * 4 * blockDim.x
};
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }

// Test device primitive `i`.  Returns t or +inf.  BIG = primitive of the always-tested list.
// ALL = every node and primitive head of the scene is staged in shared memory (k_extend<.., true>): no per-access range check.
// GEN = some primitive of the scene is reached through its general 128-byte record (S.has_general: rotated / scaled spheres, media).  Scenes without any — book-one,
// cornell-box, the sphere cloud — run kernels from which the out-of-line general path is compiled away: its mere presence (call ABI,
// argument marshalling) costs registers in the traversal loop (k_extend<., true>: 24 bytes of spills, -3 %; k_shade: -7 to -11 %).
// Box primitive (csrc/box_prims.h): the closest hit among the six face rectangles of an axis-aligned box, each a Rectangle::hit
// (src/geometry.rs:152-181: its own plane, t >= 1e-6) whose extents are the whole face.  The line crosses the box surface at t_en and
// t_ex; the entry crossing answers unless it lies behind the origin or on the ray's own face, then the exit crossing.  t is
// (plane - o_k) / d_k — the difference first, so a ray that starts near a plane does not lose its t to cancellation.
__device__ __forceinline__ float box_hit(const RayCtx& R, const float4 L, const float4 H, uint32_t self_prim, uint32_t& id) {
    const float x0 = (L.x - R.o.x) * R.inv.x, x1 = (H.x - R.o.x) * R.inv.x;
    const float y0 = (L.y - R.o.y) * R.inv.y, y1 = (H.y - R.o.y) * R.inv.y;
    const float z0 = (L.z - R.o.z) * R.inv.z, z1 = (H.z - R.o.z) * R.inv.z;
    // fminf / fmaxf drop a NaN operand (0 * inf: origin on a plane of a zero direction component): that axis then does not constrain
    const float nx = fminf(x0, x1), ny = fminf(y0, y1), nz = fminf(z0, z1);
    const float fx = fmaxf(x0, x1), fy = fmaxf(y0, y1), fz = fmaxf(z0, z1);
    const float t_en = fmaxf(fmaxf(nx, ny), nz), t_ex = fminf(fminf(fx, fy), fz);
    if (!(t_en <= t_ex)) return CUDART_INF_F;  // (also NaN) the line misses the box: it crosses none of the faces inside their bounds
    const uint32_t ke = (nx == t_en) ? 0u : ((ny == t_en) ? 1u : 2u), kx = (fx == t_ex) ? 0u : ((fy == t_ex) ? 1u : 2u);
    const float de = ke == 0u ? R.d.x : (ke == 1u ? R.d.y : R.d.z), dx = kx == 0u ? R.d.x : (kx == 1u ? R.d.y : R.d.z);
    const uint32_t first = __float_as_uint(L.w);
    const uint32_t id_en = first + 2u * ke + (de > 0.0f ? 0u : 1u), id_ex = first + 2u * kx + (dx > 0.0f ? 1u : 0u);
    if (t_en >= RT_T_EPS && id_en != self_prim) { id = id_en; return t_en; }
    if (t_ex >= RT_T_EPS && t_ex < CUDART_INF_F && id_ex != self_prim) { id = id_ex; return t_ex; }
    return CUDART_INF_F;
}

// `id` = the device primitive the returned t belongs to: i itself, or the crossed face of a box primitive.
template <bool BIG, bool ALL = false, bool GEN = true>
__device__ __forceinline__ float prim_hit(const DScene& S, const Staged& sm, uint32_t i, const RayCtx& R, uint32_t self_prim, const MediumKey& key, uint32_t& id) {
    RT_ASSERT(i < S.n_leaf_prims);
    id = i;
    const bool staged = ALL || i < S.n_staged_prims;
    float4 h;
    if (staged) h = lds128(sm.heads + 16u * i);
    else h = __ldg(&S.prim_head[i]);
    if (h.w > 0.0f) {
        if (BIG) return big_sphere_hit(R, f3(h.x, h.y, h.z), h.w, S.always_k[i], i == self_prim);
        return sphere_hit(R, f3(h.x, h.y, h.z), h.w, i == self_prim);
    }
    if (__float_as_uint(h.y) == (uint32_t)RT_PRIM_RECT) {
        if (i == self_prim) return CUDART_INF_F;  // a plane cannot re-hit itself
        const uint32_t ri = __float_as_uint(h.z);
        RT_ASSERT(ri < S.n_leaf_rects);
        if (staged) {
            const uint32_t q = sm.rects + 64u * ri;
            return rect_hit(R, lds128(q), lds128(q + 16u), lds128(q + 32u), lds128(q + 48u));
        }
        const float4* __restrict__ q = S.rects + 4u * (size_t)ri;
        return rect_hit(R, __ldg(q), __ldg(q + 1), __ldg(q + 2), __ldg(q + 3));
    }
    if (__float_as_uint(h.y) == RT_DPRIM_BOX) {
        const uint32_t bi = __float_as_uint(h.x);
        RT_ASSERT(bi < S.n_boxes);
        if (ALL) return box_hit(R, lds128(sm.boxes + 32u * bi), lds128(sm.boxes + 32u * bi + 16u), self_prim, id);
        const float4* __restrict__ q = S.boxes + 2u * (size_t)bi;
        return box_hit(R, __ldg(q), __ldg(q + 1), self_prim, id);
    }
    RT_ASSERT(GEN && __float_as_uint(h.x) < S.n_gen);
    if (!GEN) return CUDART_INF_F;  // unreachable: a scene rendered by a GEN = false kernel has no such record
    return general_hit(&S.gen[__float_as_uint(h.x)], R.o, R.d, i == self_prim, key);
}
// Closest-hit update with the reference's tie rule (strict <, the DFS-earlier leaf wins an exact tie).
template <bool ORDERED>
__device__ __forceinline__ void take(const DScene& S, uint32_t i, float t, ClosestHit& best) {
    if (t < best.t) { best.t = t; best.prim = i; return; }
    if (ORDERED) return;  // candidates arrive in DFS order: the earlier one already holds the tie
    if (t == best.t && t < CUDART_INF_F) {
        if (__ldg(&S.prim_meta[i]).x < __ldg(&S.prim_meta[best.prim]).x) best.prim = i;
    }
}

// test primitive i and keep the closer hit
template <bool BIG, bool ORDERED, bool ALL = false, bool GEN = true>
__device__ __forceinline__ void test_prim(const DScene& S, const Staged& sm, uint32_t i, const RayCtx& R, uint32_t self_prim, const MediumKey& key, ClosestHit& best) {
    uint32_t id;
    const float t = prim_hit<BIG, ALL, GEN>(S, sm, i, R, self_prim, key, id);
    take<ORDERED>(S, id, t, best);
}

// Slab test against [0, tmax]; returns entry distance or +inf.  fminf/fmaxf drop NaNs (0 * inf when the
// origin lies on a slab of a zero direction component), which errs on the side of "hit" like the
// reference's `if t0 > tmin` form.  1/d comes from MUFU.RCP, so the far distance is widened by 8 ulp
// (Ize, robust BVH traversal): the test can only err towards visiting a node.
__device__ __forceinline__ bool slab_hit(const float4 lo, const float4 hi, const RayCtx& R, float tmax, float& tn) {
    const float x0 = fmaf(lo.x, R.inv.x, -R.oinv.x), x1 = fmaf(hi.x, R.inv.x, -R.oinv.x);
    const float y0 = fmaf(lo.y, R.inv.y, -R.oinv.y), y1 = fmaf(hi.y, R.inv.y, -R.oinv.y);
    const float z0 = fmaf(lo.z, R.inv.z, -R.oinv.z), z1 = fmaf(hi.z, R.inv.z, -R.oinv.z);
    tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.0f));
    const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), tmax)) * 1.000001f;
    return tn <= tf;
}
__device__ __forceinline__ float slab(const float4 lo, const float4 hi, const RayCtx& R, float tmax) {
    float tn;
    return slab_hit(lo, hi, R, tmax, tn) ? tn : CUDART_INF_F;
}

__device__ __forceinline__ float4 load_node(const DScene& S, const Staged& sm, uint32_t idx4) {
    if (idx4 < 2u * S.n_staged_nodes) return lds128(sm.nodes + 16u * idx4);
    return __ldg(&S.nodes[idx4]);
}
// The four float4 of a sibling pair (64 contiguous bytes); `pair` is even, so the pair is staged or not as a whole.
template <bool ALL = false>
__device__ __forceinline__ void load_pair(const DScene& S, const Staged& sm, uint32_t pair, float4& llo, float4& lhi, float4& rlo, float4& rhi) {
    RT_ASSERT((pair & 1u) == 0u && pair + 1u < S.n_nodes + (S.n_nodes & 1u));
    if (ALL || pair < S.n_staged_nodes) {
        const uint32_t a = sm.nodes + 32u * pair;
        llo = lds128(a); lhi = lds128(a + 16u); rlo = lds128(a + 32u); rhi = lds128(a + 48u);
    } else {
        // one 32-byte node per 256-bit load (sm_100: LDG.E.256): two L1 requests per pair instead of four — the r1k profile of
        // the 1 M-sphere scene has the kernel at 70 % of L1 (l1tex) throughput with four divergent LDG.128 per visit
        const float4* __restrict__ g = S.nodes + 2u * (size_t)pair;
        asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=f"(llo.x), "=f"(llo.y), "=f"(llo.z), "=f"(llo.w), "=f"(lhi.x), "=f"(lhi.y), "=f"(lhi.z), "=f"(lhi.w) : "l"(g));
        asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=f"(rlo.x), "=f"(rlo.y), "=f"(rlo.z), "=f"(rlo.w), "=f"(rhi.x), "=f"(rhi.y), "=f"(rhi.z), "=f"(rhi.w) : "l"(g + 2));
    }
}

__device__ __forceinline__ RayCtx make_ray(float3 o, float3 d) {
    RayCtx R;
    R.o = o;
    R.d = d;
    R.a = dot3(d, d);
    R.inv_a = fast_rcp(R.a);
    R.inv = f3(fast_rcp(d.x), fast_rcp(d.y), fast_rcp(d.z));
    R.oinv = f3(o.x * R.inv.x, o.y * R.inv.y, o.z * R.inv.z);
    // o/d is NaN when both are 0: replace by 0 so that lo*inv - oinv stays +-inf (conservative)
    if (R.oinv.x != R.oinv.x) R.oinv.x = 0.0f;
    if (R.oinv.y != R.oinv.y) R.oinv.y = 0.0f;
    if (R.oinv.z != R.oinv.z) R.oinv.z = 0.0f;
    return R;
}

// One NODE step of the ordered, t-pruned traversal: test the two children of `pair`; children that are leaves are not
// intersected here but returned as ONE pending primitive range [leaf_begin, leaf_end) (sibling leaves are adjacent in
// the primitive order); descend into the nearer interior child (pushing the farther) or pop.  Returns the next pair,
// RT_NO_PRIM when the traversal is over.  Primitive tests are deferred so that the lanes of a warp run them together
// ("while-while" traversal, Aila & Laine): in the if-if form the ~140-instruction leaf loop ran on nearly every warp
// iteration for ~2 of 32 lanes (r1k profile of the 1 M-sphere scene: 192 warp-instructions per iteration, of which the
// node step is ~60).
// What a node step found below `pair`: which children are leaves the ray enters, and their primitive ranges.  The range
// [begin, end) is only formed when a leaf was found (`leaf_range`), not on every step: the step is ALU-pipe bound and the
// IADD3 / SEL / ISETP chain of the range was a fifth of its ALU instructions (r2i SASS).
struct LeafPair {
    bool left, right;
    uint32_t lref, lcnt, rref, rcnt;
    __device__ __forceinline__ bool any() const { return left || right; }
};
__device__ __forceinline__ void leaf_range(const LeafPair& lp, uint32_t& begin, uint32_t& end) {
    begin = 0u;
    end = 0u;
    if (lp.left) { begin = lp.lref; end = lp.lref + lp.lcnt; }
    if (lp.right) { if (end == 0u) begin = lp.rref; end = lp.rref + lp.rcnt; }
}

// ALL (k_extend<.., true>): tree and primitive heads entirely in shared memory and no deeper than the shared-memory stack —
// no "staged or global" range checks, no spill path.
template <bool COUNT, bool ALL = false>
__device__ __forceinline__ uint32_t node_step_core(const DScene& S, const Staged& sm, const RayCtx& R, const ClosestHit& best, uint32_t pair, int& sp,
                                                   uint32_t* stack_spill, LeafPair& lp, TraversalWork* work) {
    const uint32_t sstride = 4u * blockDim.x;
    float4 llo, lhi, rlo, rhi;
    load_pair<ALL>(S, sm, pair, llo, lhi, rlo, rhi);
    // hit flags straight from the slab compare (the entry distances only order the two children): the extend kernels are bound by
    // the ALU pipe (FMNMX / FSETP / SEL at half the FMA rate, r2i profile: 67 % vs 15 %), a +inf sentinel cost 2 FSEL + 2 FSETP here
    float tl, tr;
    const bool lhit = slab_hit(llo, lhi, R, best.t, tl);
    const bool rhit = slab_hit(rlo, rhi, R, best.t, tr);
    if (COUNT) work->box_tests += 2;
    lp.lcnt = __float_as_uint(lhi.w); lp.rcnt = __float_as_uint(rhi.w);
    lp.lref = __float_as_uint(llo.w); lp.rref = __float_as_uint(rlo.w);
    lp.left = lhit && lp.lcnt;
    lp.right = rhit && lp.rcnt;
    // interior children: descend into the nearer one, push the farther
    const bool lin = lhit && !lp.lcnt, rin = rhit && !lp.rcnt;
    uint32_t next = RT_NO_PRIM;
    if (lin && rin) {
        const bool swap = tr < tl;
        const uint32_t pushed = swap ? lp.lref : lp.rref;
        if (ALL) {
            RT_ASSERT(sp >= 0 && (uint32_t)sp < S.bvh_depth + 2u);
            sts32(sm.stack + sp * sstride, pushed);
            ++sp;
        } else {
            RT_ASSERT(sp >= 0 && sp < RT_STACK_DEPTH);
            if (sp < (int)RT_SMEM_STACK) sts32(sm.stack + sp * sstride, pushed);
            else stack_spill[sp - RT_SMEM_STACK] = pushed;
            sp += (sp < RT_STACK_DEPTH - 1);
        }
        next = swap ? lp.rref : lp.lref;
    } else if (lin) {
        next = lp.lref;
    } else if (rin) {
        next = lp.rref;
    } else if (sp > 0) {
        --sp;
        if (ALL) next = lds32(sm.stack + sp * sstride);
        else next = (sp < (int)RT_SMEM_STACK) ? lds32(sm.stack + sp * sstride) : stack_spill[sp - RT_SMEM_STACK];
    }
    return next;
}
template <bool COUNT, bool ALL = false>
__device__ __forceinline__ uint32_t node_step(const DScene& S, const Staged& sm, const RayCtx& R, const ClosestHit& best, uint32_t pair, int& sp,
                                              uint32_t* stack_spill, uint32_t& leaf_begin, uint32_t& leaf_end, TraversalWork* work) {
    LeafPair lp;
    const uint32_t next = node_step_core<COUNT, ALL>(S, sm, R, best, pair, sp, stack_spill, lp, work);
    leaf_range(lp, leaf_begin, leaf_end);
    return next;
}

// ---- node step of the all-in-shared-memory variant: "octant-ready" pair records --------------------------------------------------
// k_extend<.., true> re-lays the tree out while staging it: per sibling pair 7 float4 (112 B)
//   X+ = (l.lo.x, l.hi.x, r.lo.x, r.hi.x)   X- = (l.hi.x, l.lo.x, r.hi.x, r.lo.x)   Y+ Y- Z+ Z- likewise   M = (lref, lcnt, rref, rcnt)
// A ray picks X+ or X- once, by the sign of its direction, and then every LDS.128 delivers (near, far) of both children for that
// axis: the 12 FMNMX per pair that sorted the slab planes are gone.  The step is bound by the ALU pipe (FMNMX / FSETP / ISETP /
// SEL at half the FMA rate; r2i: alu 67 %, fma 15 %), so this is a third of its ALU work.  Same plane distances, same decisions.
constexpr uint32_t RT_OCT_PAIR_BYTES = 112u;
struct OctBase {
    uint32_t x, y, z, m;  // shared-window byte addresses of the chosen X, Y, Z records and of M, for pair 0
};
__device__ __forceinline__ OctBase oct_base(const Staged& sm, const RayCtx& R) {
    OctBase ob{sm.nodes + (R.inv.x < 0.0f ? 16u : 0u), sm.nodes + 32u + (R.inv.y < 0.0f ? 16u : 0u), sm.nodes + 64u + (R.inv.z < 0.0f ? 16u : 0u),
               sm.nodes + 96u};
    // keep the three chosen addresses in registers: ptxas otherwise re-derives them from the signs in every step (3 FSETP + 6 integer
    // ALU instructions per step, on the pipe the step is bound by)
    asm volatile("" : "+r"(ob.x), "+r"(ob.y), "+r"(ob.z));
    return ob;
}
template <bool COUNT>
__device__ __forceinline__ uint32_t node_step_oct(const Staged& sm, const OctBase& ob, const RayCtx& R, const ClosestHit& best, uint32_t pair, int& sp,
                                                  LeafPair& lp, TraversalWork* work) {
    const uint32_t sstride = 4u * blockDim.x;
    RT_ASSERT((pair & 1u) == 0u && sp >= 0);
    const uint32_t a = pair * (RT_OCT_PAIR_BYTES / 2u);  // `pair` is the (even) node index of the left child
    const float4 X = lds128(a + ob.x), Y = lds128(a + ob.y), Z = lds128(a + ob.z), M = lds128(a + ob.m);
    const float lnx = fmaf(X.x, R.inv.x, -R.oinv.x), lfx = fmaf(X.y, R.inv.x, -R.oinv.x);
    const float rnx = fmaf(X.z, R.inv.x, -R.oinv.x), rfx = fmaf(X.w, R.inv.x, -R.oinv.x);
    const float lny = fmaf(Y.x, R.inv.y, -R.oinv.y), lfy = fmaf(Y.y, R.inv.y, -R.oinv.y);
    const float rny = fmaf(Y.z, R.inv.y, -R.oinv.y), rfy = fmaf(Y.w, R.inv.y, -R.oinv.y);
    const float lnz = fmaf(Z.x, R.inv.z, -R.oinv.z), lfz = fmaf(Z.y, R.inv.z, -R.oinv.z);
    const float rnz = fmaf(Z.z, R.inv.z, -R.oinv.z), rfz = fmaf(Z.w, R.inv.z, -R.oinv.z);
    const float tl = fmaxf(fmaxf(lnx, lny), fmaxf(lnz, 0.0f)), tr = fmaxf(fmaxf(rnx, rny), fmaxf(rnz, 0.0f));
    const float fl = fminf(fminf(lfx, lfy), fminf(lfz, best.t)) * 1.000001f, fr = fminf(fminf(rfx, rfy), fminf(rfz, best.t)) * 1.000001f;
    const bool lhit = tl <= fl, rhit = tr <= fr;
    if (COUNT) work->box_tests += 2;
    lp.lref = __float_as_uint(M.x); lp.lcnt = __float_as_uint(M.y);
    lp.rref = __float_as_uint(M.z); lp.rcnt = __float_as_uint(M.w);
    lp.left = lhit && lp.lcnt;
    lp.right = rhit && lp.rcnt;
    const bool lin = lhit && !lp.lcnt, rin = rhit && !lp.rcnt;
    uint32_t next = RT_NO_PRIM;
    if (lin && rin) {
        const bool swap = tr < tl;
        sts32(sm.stack + sp * sstride, swap ? lp.lref : lp.rref);
        ++sp;
        next = swap ? lp.rref : lp.lref;
    } else if (lin) {
        next = lp.lref;
    } else if (rin) {
        next = lp.rref;
    } else if (sp > 0) {
        --sp;
        next = lds32(sm.stack + sp * sstride);
    }
    return next;
}

template <bool COUNT, bool ALL = false, bool GEN = true>
__device__ __forceinline__ void leaf_tests(const DScene& S, const Staged& sm, const RayCtx& R, uint32_t self_prim, const MediumKey& key, ClosestHit& best,
                                           uint32_t begin, uint32_t end, TraversalWork* work) {
    RT_ASSERT(begin <= end && end <= S.n_leaf_prims);
    for (uint32_t i = begin; i < end; ++i) test_prim<false, false, ALL, GEN>(S, sm, i, R, self_prim, key, best);
    if (COUNT) work->prim_tests += end - begin;
}

template <bool COUNT = false, bool ALL = false, bool GEN = true>
__device__ __forceinline__ ClosestHit closest_hit(const DScene& S, const Staged& sm, float3 o, float3 d, uint32_t self_prim, const MediumKey& key,
                                                  TraversalWork* work = nullptr) {
    ClosestHit best{CUDART_INF_F, RT_NO_PRIM};
    const RayCtx R = make_ray(o, d);

    if (S.small_all_rects) {
        // small scene made of rectangles only (cornell-box: 18): rectangle i IS primitive i, records in shared memory.
        // Branch-free body (every reject is a predicate on the final select) so that ptxas can unroll and overlap the
        // LDS of the next rectangle with the arithmetic of this one; strict `<` in DFS order = the reference's tie rule.
#pragma unroll 3
        for (uint32_t i = 0; i < S.n_small; ++i) {
            const uint32_t q = sm.rects + 64u * i;
            const float4 rz = lds128(q), rx = lds128(q + 16u), ry = lds128(q + 32u), ext = lds128(q + 48u);
            const float t = __fdividef(-(dot3(rz, R.o) + rz.w), dot3(rz, R.d));
            const float3 p = fma3(t, R.d, R.o);
            const float x = dot3(rx, p) + rx.w, y = dot3(ry, p) + ry.w;
            const bool ok = (t >= RT_T_EPS) & (t < best.t) & (fabsf(x) <= ext.x) & (fabsf(y) <= ext.y) & (i != self_prim);
            best.t = ok ? t : best.t;
            best.prim = ok ? i : best.prim;
        }
        if (COUNT) work->prim_tests += S.n_small;
        return best;
    }
    if (S.n_small) {
        // small scene: the whole primitive list in DFS order, warp-uniform (geometry.rs:76-116)
        for (uint32_t i = 0; i < S.n_small; ++i) test_prim<false, true, false, GEN>(S, sm, i, R, self_prim, key, best);
        if (COUNT) work->prim_tests += S.n_small;
        return best;
    }
    for (uint32_t i = 0; i < S.n_always; ++i) test_prim<true, false, ALL, GEN>(S, sm, i, R, self_prim, key, best);
    if (COUNT) work->prim_tests += S.n_always;
    if (S.n_nodes == 0) return best;

    if (S.root_is_leaf) {
        const float4 lo = load_node(S, sm, 0), hi = load_node(S, sm, 1);
        if (COUNT) work->box_tests += 1;
        if (slab(lo, hi, R, best.t) < CUDART_INF_F) {
            const uint32_t first = __float_as_uint(lo.w), cnt = __float_as_uint(hi.w);
            for (uint32_t i = first; i < first + cnt; ++i) test_prim<false, false, false, GEN>(S, sm, i, R, self_prim, key, best);
            if (COUNT) work->prim_tests += cnt;
        }
        return best;
    }

    // Traversal stack: the first RT_SMEM_STACK entries live in shared memory (conflict-free: one column per thread),
    // so the per-thread local-memory stack no longer competes with BVH nodes for L1; deeper entries (trees deeper
    // than RT_SMEM_STACK, i.e. > 16 M well-balanced primitives) spill to a local array.
    uint32_t stack_spill[RT_STACK_DEPTH - RT_SMEM_STACK];
    int sp = 0;
    uint32_t pair = ALL ? 0u : __float_as_uint(load_node(S, sm, 0).w);  // left child of the root; right = pair + 1
    // while-while: lanes walk nodes until each has found a leaf range (or finished), then the warp intersects the pending
    // ranges together.  One back-edge per loop, so the warp reconverges at both loop heads.
    LeafPair lp;
    if (ALL) {
        const OctBase ob = oct_base(sm, R);
        pair = lds32(ob.m);  // M.lref of pair 0 = left child of the root
        do {
            bool found = false;
            while (pair != RT_NO_PRIM) {
                pair = node_step_oct<COUNT>(sm, ob, R, best, pair, sp, lp, work);
                if (lp.any()) { found = true; break; }
            }
            if (found) {
                uint32_t leaf_begin, leaf_end;
                leaf_range(lp, leaf_begin, leaf_end);
                leaf_tests<COUNT, true, GEN>(S, sm, R, self_prim, key, best, leaf_begin, leaf_end, work);
            }
        } while (pair != RT_NO_PRIM);
        return best;
    }
    // trees that reach into global memory: the range is formed in the step (the loop-carried form of the flags costs this variant
    // more than the range arithmetic: book-two 542 vs 566 ms of extend per image, r2j)
    uint32_t leaf_begin = 0u, leaf_end = 0u;
    do {
        while (pair != RT_NO_PRIM && leaf_end == leaf_begin) pair = node_step<COUNT, false>(S, sm, R, best, pair, sp, stack_spill, leaf_begin, leaf_end, work);
        leaf_tests<COUNT, false, GEN>(S, sm, R, self_prim, key, best, leaf_begin, leaf_end, work);
        leaf_begin = leaf_end = 0u;
    } while (pair != RT_NO_PRIM);
    return best;
}

}  // namespace rtb
