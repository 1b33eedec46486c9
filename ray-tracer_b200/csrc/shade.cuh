// Device shading routines: hit-record reconstruction (point, normal, uv), textures, material
// scatter / emission, camera ray generation and the direct samplers.
//
// Reference semantics:
//   Sprite::hit point/normal back-transform  src/sprite.rs:108-126 (forward matrix, not renormalised)
//   unitSphereUv                             src/geometry.rs:35-39
//   Lambertian / Metal / Dielectric / DiffuseLight / Isotropic   src/material.rs:60-70, 98-119, 140-193, 274-298, 317-326
//   SolidColor / CheckerTexture / ImageTexture src/material.rs:200-265, lookup examples/main.rs:267-280
//   PerspectiveCamera::ray                   src/camera.rs:90-107 (scalar lens offset broadcast over xyz)
//   randomInUnitSphere / randomInUnitDisk    src/util.rs:6-42 — replaced by direct (non-rejection)
//     samplers with the identical distribution, driven by Philox (philox.h).
#pragma once
#include "intersect.cuh"

namespace rtb {

constexpr float RT_PI = 3.14159265358979323846f;

struct HitDetail {
    float3 p, n;
    float t;  // t of the record (refined in f64 for translation-only spheres)
    float u, v;
    bool uv_valid;
};

// Hit point and unit normal of a translation-only sphere for the primitive extend chose.  A far, small sphere (r = 0.2 seen from 15
// units, r = 0.25 from 200) leaves extend's f32 t only ~1e-5 of absolute accuracy along the ray, i.e. up to 1e-4 (1e-3 at 200 units) of
// normal error, so t is refined here — once per path SEGMENT, not per candidate.  Round 1 re-solved the quadratic in f64 (~100
// DFMA-class instructions, 14 % of k_shade, the FP64 pipe at 13 %); this is the same refinement in compensated f32:
//   * oc = o - c as an unevaluated sum hi + lo (Dekker's TwoDiff: lo is the exact rounding error of hi), so the centre-relative
//     origin is exact although |o|, |c| >> r;
//   * q(t) = oc + t d through ONE fma per component: the product is exact inside the fma and the sum is rounded once, to the
//     precision of the RESULT, which is small (|q| = r) — the cancellation costs nothing;
//   * one Newton step on f(t) = |q(t)|^2 - r^2 (f' = 2 q.d) from extend's t: quadratic convergence from a 1e-5-class error leaves
//     r * 1e-7 / cos(theta) of error in t, and the normal q / |q| accurate to a few f32 ulps.
// Checked by gate 1 on every scene, the 1 M-sphere cloud included (tests/test_gpu_parity.py: normals within 1e-4 of the f64 oracle).
__device__ __forceinline__ void two_diff(float a, float b, float& hi, float& lo) {
    hi = a - b;
    const float v = hi - a;
    lo = (a - (hi - v)) - (b + v);
}
__device__ __forceinline__ void sphere_record(float3 o, float3 d, float3 ctr, float r, float t32, float3& p, float3& unit, float& t_out) {
    if (r > 8.0f * t32) {
        // A sphere that is HUGE against the hit distance (the r = 1000 ground seen from a few units) is the opposite case: extend's t is
        // already as good as f32 gets (intersect.cuh big_sphere_hit expands around the world origin), while here |q| = r, so f = |q|^2 - r^2
        // would cancel at r^2 ~ 1e6 and c + r unit would round at ulp(r).  The reference's own formulas are the accurate ones in f32:
        // intersection = ray.at(t) (geometry.rs:67), normal = normalize((p - c) / r) (geometry.rs:68).
        p = fma3(t32, d, o);
        unit = normalize3(p - ctr);
        t_out = t32;
        return;
    }
    float3 hi, lo;
    two_diff(o.x, ctr.x, hi.x, lo.x);
    two_diff(o.y, ctr.y, hi.y, lo.y);
    two_diff(o.z, ctr.z, hi.z, lo.z);
    float t = t32;
#pragma unroll
    for (int it = 0; it < 2; ++it) {  // (the second step costs 15 instructions and removes any dependence on how good extend's t was)
        const float3 q = f3(fmaf(t, d.x, hi.x) + lo.x, fmaf(t, d.y, hi.y) + lo.y, fmaf(t, d.z, hi.z) + lo.z);
        const float f = fmaf(-r, r, dot3(q, q)), g = 2.0f * dot3(q, d);
        const float step = __fdividef(f, g);
        // a grazing hit has f' = 2 q.d -> 0 (the two roots meet): a step that is not tiny against the sphere is no refinement, keep t
        if (fabsf(step) * sqrtf(dot3(d, d)) <= 0.2f * r) t -= step;
    }
    const float3 q = f3(fmaf(t, d.x, hi.x) + lo.x, fmaf(t, d.y, hi.y) + lo.y, fmaf(t, d.z, hi.z) + lo.z);
    unit = q * rsqrtf(dot3(q, q));  // geometry.rs:68 normalize(p / r)
    p = fma3(r, unit, ctr);
    t_out = t;
}

__device__ __forceinline__ void sphere_uv(float3 unit, float& u, float& v) {  // geometry.rs:35-39, acos clamped (A.11)
    u = 0.5f + atan2f(unit.x, unit.z) * (1.0f / (2.0f * RT_PI));
    v = 1.0f - acosf(fminf(fmaxf(unit.y, -1.0f), 1.0f)) * (1.0f / RT_PI);
}

// General 128-byte records (rotated / scaled spheres, media): the rare path.
// inline by default: k_shade<1, true> on book-two 20.6 vs 19.8 Grays/s out of line (profiles/r2_20_shade2_variants.txt); the kernels of scenes
// WITHOUT general records (GEN = false) do not contain this code at all
#ifndef RT_GENERAL_DETAIL_INLINE
#define RT_GENERAL_DETAIL_INLINE 1
#endif
#if RT_GENERAL_DETAIL_INLINE
__device__ __forceinline__ HitDetail general_detail(
#else
__device__ __noinline__ HitDetail general_detail(
#endif
    const RtPrim* __restrict__ g, float3 o, float3 d, float t, bool want_uv, const MediumKey key) {
    HitDetail h;
    h.u = h.v = 0.0f;
    h.t = t;
    h.uv_valid = want_uv;
    const uint32_t kind = g->kind;
    if (kind == RT_PRIM_MEDIUM) {
        // redo both boundary probes (same arithmetic as extend): which surfaces, where, which branch
        MediumProbe pr;
        pr.t1 = pr.t2 = 0.0f; pr.i1 = pr.i2 = 0u; pr.entering = false;
        const RtPrim* __restrict__ bnd = g + g->aux;  // this medium's boundary surfaces (intersect.cuh general_hit)
        medium_generic_hit(bnd, 0u, (uint32_t)g->param[2], o, d, g->param[1], g->order, key, &pr);
        float u1 = 0.0f, v1 = 0.0f, u2 = 0.0f, v2 = 0.0f;
        const float3 n1 = boundary_normal(bnd, pr.i1, o, d, pr.t1, want_uv ? &u1 : nullptr, &v1);
        if (pr.entering) {
            const float3 n2 = boundary_normal(bnd, pr.i2, o, d, pr.t2, want_uv ? &u2 : nullptr, &v2);
            h.p = fma3(pr.t1 + 1e-6f + t, d, o);  // volume.rs:69 on the shadowed ray (A.9 ii)
            h.n = (n1 + n2) * 0.5f;
            h.u = u1 + u2; h.v = v1 + v2;
        } else {
            h.p = fma3(t, d, o);
            h.n = n1;
            h.u = u1; h.v = v1;
        }
        return h;
    }
    const float3 lo = xform_point(g->local_from_world, o);
    const float3 ld = xform_vector(g->local_from_world, d);
    if (kind == RT_PRIM_SPHERE) {
        const float r = g->param[0];
        const float3 q = fma3(t, ld, lo) * (1.0f / r);
        const float3 unit = normalize3(q);
        h.p = xform_point(g->world_from_local, unit * r);
        h.n = xform_vector(g->world_from_local, unit);
        if (want_uv) sphere_uv(q, h.u, h.v);
        return h;
    }
    // constant medium: redo the boundary query (same arithmetic as extend) for t1 and the branch
    float t1 = 0.0f;
    bool entering = false;
    const float r = g->param[0];
    medium_sphere_hit(lo, ld, r, g->param[1], g->order, key, &t1, &entering);
    float tn = 0.0f, tf = 0.0f;
    sphere_roots(lo, ld, r, tn, tf);
    if (entering) {
        // volume.rs:69 evaluates `ray.at(t1 + distance)` on the SHADOWED ray that starts at p1 + 1e-6 d:
        // point = o + (t1 + 1e-6 + t) d with t = t1 + distance (A.9 ii)
        const float3 lp = fma3(t1 + 1e-6f + t, ld, lo);
        const float3 n1 = normalize3(fma3(tn, ld, lo) * (1.0f / r)), n2 = normalize3(fma3(tf, ld, lo) * (1.0f / r));
        h.p = xform_point(g->world_from_local, lp);
        h.n = xform_vector(g->world_from_local, (n1 + n2) * 0.5f);
        if (want_uv) { float u1, v1, u2, v2; sphere_uv(n1, u1, v1); sphere_uv(n2, u2, v2); h.u = u1 + u2; h.v = v1 + v2; }
    } else {
        const float3 lp = fma3(t, ld, lo);
        const float3 n1 = normalize3(fma3(t1, ld, lo) * (1.0f / r));
        h.p = xform_point(g->world_from_local, lp);
        h.n = xform_vector(g->world_from_local, n1);
        if (want_uv) sphere_uv(n1, h.u, h.v);
    }
    return h;
}

// Rebuild the reference's HitRecord for (ray, t, prim).  `want_uv` skips the atan2/acos when the
// material's texture does not read uv.
// S.prim_head / S.rect_nrm are read with plain (generic) loads: k_shade points them at its shared-memory copies when
// the tables are small enough to be staged (ld.global.nc on a shared address would be illegal).
template <bool GEN = true>
__device__ __forceinline__ HitDetail hit_detail(const DScene& S, uint32_t i, float3 o, float3 d, float t, bool want_uv, const MediumKey& key,
                                                const float4 head) {
    HitDetail h;
    h.u = h.v = 0.0f;
    h.t = t;
    h.uv_valid = want_uv;
    if (head.w > 0.0f) {  // translation-only sphere
        sphere_record(o, d, f3(head.x, head.y, head.z), head.w, t, h.p, h.n, h.t);
        if (want_uv) sphere_uv(h.n, h.u, h.v);
        return h;
    }
    if (__float_as_uint(head.y) == (uint32_t)RT_PRIM_RECT) {
        // Sprite::hit returns M (o' + t d') = o + t d and M (0,0,1) (sprite.rs:108-115): the world-space point needs no
        // matrix at all and the normal is one precomputed column — 16 B instead of the 128-byte record
        const uint32_t ri = __float_as_uint(head.z);
        const float4 nrm = S.rect_nrm[ri];
        h.p = fma3(t, d, o);
        h.n = f3(nrm.x, nrm.y, nrm.z);
        if (want_uv) {  // geometry.rs:170-171
            const float4* __restrict__ q = S.rects + 4u * (size_t)ri;
            const float4 rx = __ldg(q + 1), ry = __ldg(q + 2), ext = __ldg(q + 3);
            h.u = (dot3(rx, h.p) + rx.w + ext.x) / (2.0f * ext.x);
            h.v = (dot3(ry, h.p) + ry.w + ext.y) / (2.0f * ext.y);
        }
        return h;
    }
    if (!GEN) { h.p = fma3(t, d, o); h.n = f3(0.0f, 0.0f, 0.0f); return h; }  // unreachable (prim_hit<.., GEN = false>)
    return general_detail(&S.gen[__float_as_uint(head.x)], o, d, t, want_uv, key);
}

// ---- textures ------------------------------------------------------------------------------------

__device__ __forceinline__ float3 texture_value(const DScene& S, int32_t tex, float u, float v) {
    // CheckerTexture nests (material.rs:217-245): walk down at most 8 levels
    for (int depth = 0; depth < 8; ++depth) {
        RT_ASSERT(tex >= 0 && (uint32_t)tex < S.n_texs);
        const RtTexture* __restrict__ t = &S.texs[tex];
        const uint32_t kind = t->kind;
        if (kind == RT_TEX_SOLID) return f3(t->rgb[0], t->rgb[1], t->rgb[2]);
        if (kind == RT_TEX_CHECKER) {
            const float sine = sinpif(20.0f * u) * sinpif(20.0f * v);  // sin(2 pi 10 u) sin(2 pi 10 v), material.rs:238
            tex = sine > 0.0f ? t->a : t->b;
            continue;
        }
        // baked image, nearest texel, clamped (main.rs:271-274; A.11)
        const float fx = u * (float)t->width, fy = (1.0f - v) * (float)t->height;
        uint32_t px = fx > 0.0f ? (uint32_t)fx : 0u, py = fy > 0.0f ? (uint32_t)fy : 0u;
        px = min(px, t->width - 1u);
        py = min(py, t->height - 1u);
        const float4 c = __ldg(&S.texels[t->texel_offset + (uint64_t)py * t->width + px]);
        return f3(c.x, c.y, c.z);
    }
    return f3(0.0f, 0.0f, 0.0f);
}

// ---- samplers (identical in distribution to util.rs:6-42; identical variates to the oracle's PAIRED mode) ----

// Sampling maps on the hardware approximations (MUFU.SIN / COS / LG2 / EX2 / SQRT): absolute errors of a few 1e-7 in a unit vector's
// components and ~1e-6 relative in the ball radius, against ~60 instructions for sincospif + cbrtf + sqrtf — on cornell-box, where every
// bounce is Lambertian, that was 5 % of the fused kernel's instructions.  RT_FAST_SAMPLING=0 keeps the libm forms for A/B.
#ifndef RT_FAST_SAMPLING
#define RT_FAST_SAMPLING 1
#endif
// sin and cos of 2 pi u, u in [0, 1)
__device__ __forceinline__ void sincos_2pi(float u, float& s, float& c) {
#if RT_FAST_SAMPLING
    const float t = 6.283185307179586f * (u - 0.5f);  // [-pi, pi): the range in which __sinf / __cosf are accurate to 2^-21
    s = -__sinf(t);
    c = -__cosf(t);
#else
    sincospif(2.0f * u, &s, &c);
#endif
}
__device__ __forceinline__ float3 ball_from_uniforms(float u0, float u1, float u2) {
#if RT_FAST_SAMPLING
    const float z = 1.0f - 2.0f * u0, rho = fast_sqrt(fmaxf(0.0f, 1.0f - z * z)), r = exp2f(__log2f(u2) * 0.33333334f);  // u2 = 0: exp2(-inf) = 0
#else
    const float z = 1.0f - 2.0f * u0, rho = sqrtf(fmaxf(0.0f, 1.0f - z * z)), r = cbrtf(u2);
#endif
    float s, c;
    sincos_2pi(u1, s, c);
    return f3(r * rho * c, r * rho * s, r * z);
}
__device__ __forceinline__ float2 disk_from_uniforms(float u0, float u1) {
#if RT_FAST_SAMPLING
    const float r = fast_sqrt(u0);
#else
    const float r = sqrtf(u0);
#endif
    float s, c;
    sincos_2pi(u1, s, c);
    return make_float2(r * c, r * s);
}

// ---- camera (camera.rs:90-107 + the driver's jitter, examples/book-one.rs:71-72) --------------------

__device__ __forceinline__ void camera_ray(const RtCamera& cam, uint32_t x, uint32_t y, uint32_t width, uint32_t height, const Philox4& rnd,
                                           float3& o, float3& d) {
#if RT_FAST_SAMPLING
    const float u = ((float)x + u24f(rnd.x)) * __frcp_rn((float)width);  // (the compiler hoists the two reciprocals out of k_generate's loop)
    const float v = ((float)y + u24f(rnd.y)) * __frcp_rn((float)height);
#else
    const float u = ((float)x + u24f(rnd.x)) / (float)width;
    const float v = ((float)y + u24f(rnd.y)) / (float)height;
#endif
    const float3 eye = f3(cam.eye[0], cam.eye[1], cam.eye[2]);
    const float3 ll = f3(cam.lower_left[0], cam.lower_left[1], cam.lower_left[2]);
    const float3 hz = f3(cam.horizontal[0], cam.horizontal[1], cam.horizontal[2]);
    const float3 vt = f3(cam.vertical[0], cam.vertical[1], cam.vertical[2]);
    const float3 target = fma3(v, vt, fma3(u, hz, ll));
    if (cam.lens_radius == 0.0f) {
        o = eye;
        d = normalize3(target - eye);
        return;
    }
    const float2 rd = disk_from_uniforms(u24f(rnd.z), u24f(rnd.w));
    const float offset = cam.lens_radius * (rd.x * u + rd.y * v);  // scalar, broadcast over xyz (A.7)
    o = f3(eye.x + offset, eye.y + offset, eye.z + offset);
    d = normalize3(target - o);
}

// ---- materials -----------------------------------------------------------------------------------

struct ScatterResult {
    bool scattered;
    float3 dir;          // next direction (unit)
    float3 attenuation;  // or emission when !scattered
};

__device__ __forceinline__ float3 reflect3(float3 v, float3 n) { return fma3(-2.0f * dot3(v, n), n, v); }  // vec3.rs:100-102

// material.scatter + material.emitted for one hit.  `rnd` = Philox(segment, RT_DIM_SHADE).
__device__ __forceinline__ ScatterResult shade_material(const DScene& S, const DMaterial& m, float3 d, const HitDetail& h, const Philox4& rnd) {
    ScatterResult r;
    r.scattered = false;
    r.dir = f3(0.0f, 0.0f, 0.0f);
    const float3 tex = (m.texture < 0) ? f3(m.r, m.g, m.b) : texture_value(S, m.texture, h.u, h.v);
    r.attenuation = tex;
    switch (m.kind) {
        case RT_MAT_LAMBERTIAN: {
            r.dir = normalize3(h.n + ball_from_uniforms(u24f(rnd.x), u24f(rnd.y), u24f(rnd.z)));
            r.scattered = true;
        } break;
        case RT_MAT_METAL: {
            if (dot3(d, h.n) < 0.0f) {
                const float3 refl = reflect3(normalize3(d), h.n);
                r.dir = (m.fuzz == 0.0f) ? refl : normalize3(fma3(m.fuzz, ball_from_uniforms(u24f(rnd.x), u24f(rnd.y), u24f(rnd.z)), refl));
                r.scattered = true;
            } else {
                r.attenuation = f3(0.0f, 0.0f, 0.0f);  // absorbed; emitted() is black
            }
        } break;
        case RT_MAT_DIELECTRIC: {
            r.attenuation = f3(1.0f, 1.0f, 1.0f);
            r.scattered = true;
            float ratio;
            float3 n = h.n;
            if (dot3(d, h.n) < 0.0f) ratio = 1.0f / m.ior;
            else { ratio = m.ior; n = -n; }
            const float3 unit = normalize3(d);
            const float dt = dot3(unit, n);
            const float disc = 1.0f - ratio * ratio * (1.0f - dt * dt);
            if (disc > 0.0f) {
                const float r0s = (ratio - 1.0f) / (ratio + 1.0f), r0 = r0s * r0s;
                // material.rs:163-166: theta = acos(-d.n), reflect iff rand < r0 + (1 - r0)(1 - cos theta)^5.  acos of |x| > 1 is NaN and
                // `rand < NaN` is false: under a SCALING transform (normals are not renormalised, sprite.rs:112-115) the reference always
                // refracts — mirrored here.  Within a few f32 ulps of 1 the excess is rounding of a unit vector, not scaling: clamped.
                const float c = -dot3(d, n);
                const float om = 1.0f - fminf(fmaxf(c, -1.0f), 1.0f);
                const float om2 = om * om;
                const float prob = (fabsf(c) > 1.0f + 1e-5f) ? -1.0f : r0 + (1.0f - r0) * om2 * om2 * om;
                if (u24f(rnd.x) < prob) r.dir = reflect3(d, h.n);
                else r.dir = normalize3(fma3(-sqrtf(disc), n, (d - n * dt) * ratio));  // vec3.rs:113-124
            } else {
                r.dir = reflect3(d, n);  // total internal reflection
            }
        } break;
        case RT_MAT_DIFFUSE_LIGHT:
            break;  // never scatters; attenuation carries the emission
        case RT_MAT_ISOTROPIC: {
            r.dir = normalize3(ball_from_uniforms(u24f(rnd.x), u24f(rnd.y), u24f(rnd.z)));
            r.scattered = true;
        } break;
        default:
            r.attenuation = f3(0.0f, 0.0f, 0.0f);
    }
    return r;
}

}  // namespace rtb
