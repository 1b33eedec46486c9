/* rt_b200.h — C ABI of the B200 (sm_100a) wavefront path-tracing backend for aiifabbf/ray-tracer.
 *
 * The reference has no FFI of its own: its render path is reachable only through its Rust API
 * (render::color src/render.rs:5, trait Hit src/ray.rs:85-87, trait Bound src/optimize.rs:88-95,
 * BoundingVolumeHierarchyNode::new src/optimize.rs:366, Sprite::builder src/sprite.rs:26-64,
 * PerspectiveCamera::new src/camera.rs:25-33).  This header is what a new Rust module `src/gpu.rs`
 * binds (INTEGRATION.md shows the `extern "C"` block and build.rs); the C++ host mirror in
 * ray-tracer_b200/host/ and the ctypes binding in ray-tracer_b200/__init__.py bind the same symbols.
 *
 * Conventions: every struct is plain-old-data, little-endian; all entry points return 0 on success
 * and a negative RtStatus on failure (never abort/unwind across the boundary); rt_last_error()
 * returns a thread-local message.  Opaque handles own all device memory.  One RtContext is driven
 * by one host thread at a time, and renders of DIFFERENT scenes must not overlap within one process (the
 * kernels' launch configuration — shared-memory sizes, occupancy — is per process and set by the scene being
 * rendered); rt_render_device itself runs its wavefront lanes on internal host threads.  There is no CPU
 * fallback: without a CUDA device rt_init fails.
 */
#ifndef RT_B200_H
#define RT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_ABI_VERSION 1u

typedef enum RtStatus {
    RT_OK = 0,
    RT_ERR_INVALID = -1,     /* bad argument / malformed descriptor */
    RT_ERR_UNSUPPORTED = -2, /* scene node the flattener / device cannot express */
    RT_ERR_CUDA = -3,        /* CUDA runtime error (message in rt_last_error) */
    RT_ERR_NO_DEVICE = -4,   /* no usable sm_100 device */
    RT_ERR_OOM = -5
} RtStatus;

/* ---- flat scene description (what src/gpu.rs / host/flatten.hpp produce) ------------------------ */

typedef enum RtPrimKind {
    RT_PRIM_SPHERE = 0,        /* Sphere::hit, src/geometry.rs:42-74; param[0] = radius */
    RT_PRIM_RECT = 1,          /* Rectangle::hit, src/geometry.rs:152-181; param[0..1] = width, height */
    RT_PRIM_MEDIUM_SPHERE = 2  /* ConstantMedium<Sphere>::hit, src/volume.rs:46-100; param[0] = boundary radius, param[1] = density */
} RtPrimKind;

/* One leaf primitive with every enclosing Sprite / TransformedGeometry transform pre-composed
 * (Sprite::hit src/sprite.rs:94-138, TransformedGeometry::hit src/geometry.rs:210-246).
 * Matrices are affine 3x4, row-major: x' = m[0]*x + m[1]*y + m[2]*z + m[3], ...            */
typedef struct RtPrim {
    uint32_t kind;      /* RtPrimKind */
    int32_t material;   /* index into materials, -1 = none (hit returns black, src/render.rs:18-20) */
    uint32_t order;     /* rank in the reference's left-first DFS; closest-hit ties go to the smaller order
                           (src/optimize.rs:475,484) and it is the id rt_trace reports */
    uint32_t flags;     /* reserved, 0 */
    float param[4];
    float local_from_world[12];
    float world_from_local[12];
} RtPrim; /* 128 B */

typedef enum RtMaterialKind {
    RT_MAT_LAMBERTIAN = 0,    /* src/material.rs:60-70 */
    RT_MAT_METAL = 1,         /* src/material.rs:98-119; fuzz */
    RT_MAT_DIELECTRIC = 2,    /* src/material.rs:140-193; ior */
    RT_MAT_DIFFUSE_LIGHT = 3, /* src/material.rs:274-298 */
    RT_MAT_ISOTROPIC = 4      /* src/material.rs:317-326 */
} RtMaterialKind;

typedef struct RtMaterial {
    uint32_t kind;   /* RtMaterialKind */
    int32_t texture; /* albedo / emission texture index */
    float fuzz;
    float ior;
} RtMaterial; /* 16 B */

typedef enum RtTextureKind {
    RT_TEX_SOLID = 0,   /* SolidColor, src/material.rs:200-215; rgb */
    RT_TEX_CHECKER = 1, /* CheckerTexture, src/material.rs:217-245; a = "black" child, b = "white" child */
    RT_TEX_IMAGE = 2    /* ImageTexture<F> baked on a width x height grid (row 0 = v near 1), nearest lookup
                           as in examples/main.rs:267-280; texels at texel_offset in RtSceneDesc.texels */
} RtTextureKind;

typedef struct RtTexture {
    uint32_t kind;
    int32_t a, b;
    uint32_t width, height;
    uint32_t reserved;
    uint64_t texel_offset; /* in texels (RGBA f32 quadruples) */
    float rgb[4];
} RtTexture; /* 48 B */

typedef struct RtSceneDesc {
    uint32_t abi_version; /* RT_ABI_VERSION */
    uint32_t n_prims, n_materials, n_textures;
    const RtPrim* prims;
    const RtMaterial* materials;
    const RtTexture* textures;
    const float* texels; /* RGBA f32, n_texels * 4 floats */
    uint64_t n_texels;
} RtSceneDesc;

/* Camera constants exactly as PerspectiveCamera::new derives them in f64 (src/camera.rs:34-45),
 * rounded to f32 by the caller. */
typedef struct RtCamera {
    float eye[3], lower_left[3], horizontal[3], vertical[3];
    float lens_radius;
    float reserved[3];
} RtCamera; /* 64 B */

typedef struct RtRenderParams {
    uint32_t width, height, spp, max_depth;
    uint64_t seed;
    uint32_t row_begin, row_stride; /* this call renders rows y = row_begin + k*row_stride (multi-GPU sharding,
                                       the reference's thread partition examples/book-one.rs:63) */
    uint32_t queue_capacity;        /* live-ray queue size; 0 = default */
    uint32_t flags;                 /* RtRenderFlags */
} RtRenderParams;

typedef enum RtRenderFlags {
    RT_RENDER_STAGE_TIMING = 1u, /* bracket the stage launches of every 8th wavefront iteration with CUDA events (no extra syncs)
                                    and fill RtStats.ms_<stage> with sum over the timed iterations * iterations / timed_iterations */
    RT_RENDER_COUNT_WORK = 2u    /* run the instrumented extend kernel and fill RtStats.box_tests / prim_tests (slower; not for timing) */
} RtRenderFlags;

typedef struct RtStats {
    uint64_t paths;      /* camera samples traced */
    uint64_t rays;       /* path segments = closest-hit queries issued by extend */
    uint64_t iterations; /* wavefront iterations */
    uint64_t launches;   /* kernels launched */
    float ms_total;      /* device time of the whole render (CUDA events on the render stream) */
    float ms_generate, ms_extend, ms_shade, ms_accumulate; /* whole-render stage times estimated from the timed iterations
                                                              (RT_RENDER_STAGE_TIMING), else 0 */
    uint32_t bvh_nodes, bvh_always; /* device BVH size / primitives in the always-tested list */
    uint64_t box_tests, prim_tests; /* ray/AABB and ray/primitive tests of the whole render (RT_RENDER_COUNT_WORK), else 0 */
    uint64_t timed_iterations;      /* iterations whose launches were bracketed by events (RT_RENDER_STAGE_TIMING), else 0 */
} RtStats;

typedef struct RtRayIn {
    float origin[3];
    float direction[3];
} RtRayIn; /* 24 B */

typedef struct RtHitOut {
    int32_t hit;  /* 0 / 1 */
    int32_t prim; /* RtPrim.order of the closest primitive, -1 on miss */
    float t;
    float p[3];
    float n[3];
    float uv[2];
    int32_t material; /* material index of the hit, -1 none */
} RtHitOut; /* 48 B */

typedef struct RtContext RtContext;
typedef struct RtScene RtScene;

/* ---- entry points ------------------------------------------------------------------------------ */

/* Bind a context to one CUDA device (one process per GPU; multi-GPU = one context per rank with
 * row_begin/row_stride sharding and one framebuffer gather by the caller). */
int rt_init(int device_id, RtContext** out);
void rt_shutdown(RtContext* ctx);

/* Copies the description, rebuilds a device BVH (binned SAH) over the flattened primitives and
 * uploads everything.  The caller keeps ownership of the input arrays. */
int rt_scene_create(RtContext* ctx, const RtSceneDesc* desc, RtScene** out);
void rt_scene_destroy(RtScene* scene);

/* Replaces the `for y / for x / for _ in 0..spp { color(camera.ray(u,v), world, depth) }` nest of the
 * example drivers (examples/book-one.rs:62-78).  Output: per-pixel radiance SUMS over spp (divide by
 * spp for the mean, book-one.rs:76), rows of this shard in increasing y, [rows][width][3] f32;
 * out_sumsq (nullable) = per-pixel sum of squared per-path radiance (for standard errors).
 * Host-pointer variant (synchronous, includes the device->host copy). */
int rt_render(RtScene* scene, const RtCamera* cam, const RtRenderParams* params, float* out_sum_rgb,
              float* out_sumsq_rgb, RtStats* stats);
/* Device-pointer variant: outputs stay in HBM (d_sum / d_sumsq are device pointers, zeroed by the
 * call), work is enqueued on `cuda_stream` (a cudaStream_t, 0 = the context's stream) and the call
 * returns after the stream has drained. */
int rt_render_device(RtScene* scene, const RtCamera* cam, const RtRenderParams* params, float* d_sum_rgb,
                     float* d_sumsq_rgb, void* cuda_stream, RtStats* stats);

/* Batched `render::color(&ray, world, max_depth)` (src/render.rs:5-29) on caller-supplied rays — the literal per-ray entry point
 * of the reference, fed through the same wavefront.  Ray i is evaluated `spp` times with independent
 * random streams (Philox key = (i, sample)); out_sum_rgb[3 i ..] receives the SUM of the spp radiances (divide by spp for the
 * mean), out_sumsq_rgb (nullable) the sum of their squares.  The direction is used as given (not normalised), like color(). */
int rt_color_rays(RtScene* scene, const RtRayIn* rays, uint64_t n, uint32_t spp, uint32_t max_depth, uint64_t seed,
                  float* out_sum_rgb, float* out_sumsq_rgb, RtStats* stats);

/* Closest-hit query on caller-supplied rays (batched equivalent of `world.hit(&ray)`,
 * src/optimize.rs:469-498); medium free-flight draws are keyed by (ray index, prim order, seed). */
int rt_trace(RtScene* scene, const RtRayIn* rays, uint64_t n, uint64_t seed, RtHitOut* out);

/* mean -> sqrt -> *255 -> min(255) -> truncation, rows top to bottom (examples/book-one.rs:90-100).
 * sum_rgb: host [height][width][3] sums over spp, y up. */
int rt_resolve_u8(RtContext* ctx, const float* sum_rgb, uint32_t width, uint32_t height, uint32_t spp,
                  uint8_t* out_rgb_top_down);

/* Multi-GPU reassembly of the reference's row partition (thread i owns rows y % n == i, examples/book-one.rs:63):
 * d_gathered holds the n_shards row shards back to back, each padded to ceil(height / n_shards) rows
 * ([n_shards][rows_padded][width][3] f32, e.g. the output of one NCCL all-gather of the rt_render_device
 * buffers); d_full receives [height][width][3] with row y taken from shard y % n_shards, row y / n_shards.
 * Enqueued on `cuda_stream` (0 = the context's stream, which is then drained before returning). */
int rt_deinterleave_rows(RtContext* ctx, const float* d_gathered, float* d_full, uint32_t width, uint32_t height, uint32_t n_shards,
                         void* cuda_stream);

/* Introspection used by tests / the roofline accounting. */
int rt_scene_bvh_info(const RtScene* scene, uint32_t* n_nodes, uint32_t* n_always, uint32_t* max_depth);
/* Bytes of device memory the scene occupies (= host->device bytes rt_scene_create copied). */
uint64_t rt_scene_device_bytes(const RtScene* scene);
const char* rt_last_error(void);
uint32_t rt_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_H */
