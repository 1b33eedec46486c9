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
 * by one host thread at a time; different contexts — on the same or on different devices of one process —
 * may render concurrently from different host threads (launch configuration is kept per device and per
 * scene, nothing per process); rt_render_device itself runs its wavefront lanes on internal host threads.
 * Tuning options are fixed when the context / scene is created (RtContextOptions, RtSceneOptions; the
 * RT_B200_* environment variables are read ONCE, by rt_init / rt_init_multi, as their defaults) or passed per
 * render in RtRenderParams.flags — no entry point reads the environment afterwards.  There is no CPU
 * fallback: without a CUDA device rt_init fails.
 */
#ifndef RT_B200_H
#define RT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_ABI_VERSION 2u
#define RT_NO_PRIM 0xFFFFFFFFu /* "no primitive": rt_trace_from's self_order for a ray that starts on nothing */

typedef enum RtStatus {
    RT_OK = 0,
    RT_ERR_INVALID = -1,     /* bad argument / malformed descriptor */
    RT_ERR_UNSUPPORTED = -2, /* scene node the flattener / device cannot express */
    RT_ERR_CUDA = -3,        /* CUDA runtime error (message in rt_last_error) */
    RT_ERR_NO_DEVICE = -4,   /* no usable sm_100 device */
    RT_ERR_OOM = -5,
    RT_ERR_NCCL = -6         /* NCCL could not be loaded or a collective failed (multi-GPU entry points only) */
} RtStatus;

/* ---- flat scene description (what src/gpu.rs / host/flatten.hpp produce) ------------------------ */

typedef enum RtPrimKind {
    RT_PRIM_SPHERE = 0,        /* Sphere::hit, src/geometry.rs:42-74; param[0] = radius */
    RT_PRIM_RECT = 1,          /* Rectangle::hit, src/geometry.rs:152-181; param[0..1] = width, height */
    RT_PRIM_MEDIUM_SPHERE = 2, /* ConstantMedium<Sphere>::hit, src/volume.rs:46-100; param[0] = boundary radius, param[1] = density */
    RT_PRIM_MEDIUM = 3         /* ConstantMedium<T: Hit>::hit over any flattened boundary (src/volume.rs:40-43 accepts every Hit; the book's
                                  smoke boxes are Cube boundaries): param[1] = density, aux = index of the first boundary primitive in
                                  RtSceneDesc.boundary_prims, param[2] = their count (integral, < 2^24).  Both probes of volume.rs:47,54
                                  are closest-hit queries over those primitives, in their list (= DFS) order with strict <. */
} RtPrimKind;

/* One leaf primitive with every enclosing Sprite / TransformedGeometry transform pre-composed
 * (Sprite::hit src/sprite.rs:94-138, TransformedGeometry::hit src/geometry.rs:210-246).
 * Matrices are affine 3x4, row-major: x' = m[0]*x + m[1]*y + m[2]*z + m[3], ...            */
typedef struct RtPrim {
    uint32_t kind;      /* RtPrimKind */
    int32_t material;   /* index into materials, -1 = none (hit returns black, src/render.rs:18-20) */
    uint32_t order;     /* rank in the reference's left-first DFS; closest-hit ties go to the smaller order
                           (src/optimize.rs:475,484) and it is the id rt_trace reports */
    uint32_t aux;       /* RT_PRIM_MEDIUM: first boundary primitive; else reserved, 0 */
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
    const RtPrim* boundary_prims; /* boundary surfaces of RT_PRIM_MEDIUM primitives (spheres / rectangles, transforms pre-composed to
                                     WORLD space like every RtPrim; material and order unused); never hit by rays on their own */
    uint32_t n_boundary_prims;
    uint32_t reserved;
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
    RT_RENDER_COUNT_WORK = 2u,   /* run the instrumented extend kernel and fill RtStats.box_tests / prim_tests (slower; not for timing) */
    RT_RENDER_SAMPLE_MAJOR = 4u, /* issue paths sample-major (path = sample * pixels + pixel) instead of pixel-major; same image */
    RT_RENDER_EXTEND_PER_THREAD = 8u,   /* extend: one ray per thread whatever the tree size (default: by tree size) */
    RT_RENDER_EXTEND_PERSISTENT = 16u,  /* extend: persistent lane-refilling warps wherever a tree exists */
    RT_RENDER_TIME_EVERY_ITERATION = 32u /* with RT_RENDER_STAGE_TIMING: bracket every iteration instead of every 8th */
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
typedef struct RtMulti RtMulti;           /* several devices of one process (rt_init_multi) */
typedef struct RtMultiScene RtMultiScene; /* a scene replicated on every device of an RtMulti */

/* Options of a context; all-zero = defaults.  rt_init(device, ..) == rt_init_ex(device, NULL, ..): defaults from the RT_B200_*
 * environment variables (RT_B200_LANES, RT_B200_QUEUE_CAPACITY, RT_B200_OVERLAP_ACCUMULATE, RT_B200_STAGE_TIMING,
 * RT_B200_GEN_BLOCKS, RT_B200_ACC_BLOCKS, RT_B200_SHADE_RAYS2, RT_B200_FUSE_RECTS and the RtSceneOptions ones), read once. */
typedef enum RtContextFlags {
    RT_CTX_NO_ACCUMULATE_OVERLAP = 1u, /* run accumulate on the render stream instead of a side stream */
    RT_CTX_STAGE_TIMING = 2u,          /* every render behaves as if RT_RENDER_STAGE_TIMING were set */
    RT_CTX_SHADE_ONE_RAY = 4u,         /* k_shade shades one ray per thread per round on every scene */
    RT_CTX_NO_FUSED_RECTS = 8u         /* small all-rectangle scenes: extend stays its own kernel (k_extend_rects) instead of running inside k_shade */
} RtContextFlags;
typedef struct RtContextOptions {
    uint32_t lanes;          /* concurrent wavefronts per render, 1..4; 0 = default (2) */
    uint32_t queue_capacity; /* live-ray queue size per lane; 0 = default (4 Mi rays) */
    uint32_t flags;          /* RtContextFlags */
    uint32_t gen_blocks, acc_blocks; /* blocks per SM of k_generate / k_accumulate; 0 = default (4 / 8) */
    uint32_t reserved[3];
} RtContextOptions; /* 32 B */

typedef enum RtBvhBuilder { RT_BVH_AUTO = 0, RT_BVH_HOST_SAH = 1, RT_BVH_DEVICE_LBVH = 2 } RtBvhBuilder;
typedef enum RtSceneFlags {
    RT_SCENE_NO_FRAME_RECTS = 1u,  /* small all-rectangle scenes keep the general rectangle records (no frame-rect list) */
    RT_SCENE_NO_STAGED_PRIMS = 2u, /* primitive heads stay in global memory */
    RT_SCENE_NO_SMALL_LIST = 4u,   /* scenes of <= 32 primitives get a tree like any other */
    RT_SCENE_NO_FRAME_BOXES = 8u,  /* frame-rect list without grouping the faces of a box into one slab test */
    RT_SCENE_NO_BOX_PRIMS = 16u,   /* the six faces of an axis-aligned Cube stay six leaves of the tree (no box primitive) */
    RT_SCENE_NO_WHOLE_TREE = 32u   /* extend never runs in its whole-tree-in-shared-memory form */
} RtSceneFlags;
/* Options of a scene build; all-zero = defaults (a context's defaults come from RT_B200_BVH_BUILDER=host|device,
 * RT_B200_BVH_MAX_LEAF, RT_B200_BVH_TREELET, RT_B200_BVH_THREADS, RT_B200_STAGED_NODES, RT_B200_RECT_FRAMES=0, RT_B200_RECT_BOXES=0, RT_B200_BOX_PRIMS=0, RT_B200_WHOLE_TREE=0, RT_B200_STAGE_PRIMS=0,
 * RT_B200_SMALL_SCENE=0 at rt_init). */
typedef struct RtSceneOptions {
    uint32_t bvh_builder;   /* RtBvhBuilder: AUTO = device LBVH from 65 536 primitives, host binned SAH below */
    uint32_t bvh_max_leaf;  /* primitives per leaf, 1..16; 0 = default (4 host, 1 device) */
    uint32_t bvh_treelet_nodes; /* 0 = default (64) */
    uint32_t staged_nodes;  /* BVH nodes staged in shared memory by extend, + 1 (so 0 = default: whole tree up to 2048 nodes, else 256) */
    uint32_t flags;         /* RtSceneFlags */
    uint32_t bvh_threads;   /* host builder threads; 0 = one per core (the tree does not depend on it) */
    uint32_t reserved[2];
} RtSceneOptions; /* 32 B */

/* ---- entry points ------------------------------------------------------------------------------ */

/* Bind a context to one CUDA device.  Multi-GPU: either one context per process / rank with row_begin / row_stride
 * sharding and the gather done by the caller (ray-tracer_b200/dist.py over torch.distributed), or rt_init_multi below. */
int rt_init(int device_id, RtContext** out);
int rt_init_ex(int device_id, const RtContextOptions* options /* NULL: from the environment */, RtContext** out);
void rt_shutdown(RtContext* ctx);
/* The defaults rt_init would use right now (environment read at the time of this call).  Either pointer may be NULL. */
void rt_options_from_env(RtContextOptions* ctx_options, RtSceneOptions* scene_options);

/* Copies the description, rebuilds a device BVH (binned SAH on the host, or LBVH on the device for large scenes) over the
 * flattened primitives and uploads everything.  The caller keeps ownership of the input arrays. */
int rt_scene_create(RtContext* ctx, const RtSceneDesc* desc, RtScene** out);
int rt_scene_create_ex(RtContext* ctx, const RtSceneDesc* desc, const RtSceneOptions* options /* NULL: the context's defaults */, RtScene** out);
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
/* The same query for rays that START ON a primitive — scattered rays (src/render.rs:12-14) whose origin is the previous hit point.
 * The reference relies on f64 and `t > 1e-6` (src/geometry.rs:57-66,159) to reject the hit at the ray's own origin; in f32 the
 * origin carries up to ~6e-5 of position error (r = 1000 ground sphere), so the wavefront carries the originating primitive with
 * the ray instead (DESIGN.md "f32 policy").  self_order[i] = RtPrim.order of the primitive ray i starts on, 0xFFFFFFFF = none
 * (e.g. a ray scattered inside a ConstantMedium); self_order == NULL is rt_trace. */
int rt_trace_from(RtScene* scene, const RtRayIn* rays, const uint32_t* self_order, uint64_t n, uint64_t seed, RtHitOut* out);

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

/* ---- several GPUs of one process (SURVEY §8(b) `rt_init(n_devices, device_ids)`, §8(e)) -----------------------------------------
 * One host thread + stream set per device, scene replicated, device i renders rows y % n == i — the reference's own thread
 * partition (examples/book-one.rs:54-66) — and the row shards travel to device 0 with ONE NCCL collective (grouped ncclSend /
 * ncclRecv over NVLink, communicators from ncclCommInitAll), where a kernel de-interleaves them; the main thread's `buffer[y][x] =
 * pixel` reassembly of book-one.rs:83-88.  libnccl.so.2 is loaded with dlopen on the first rt_init_multi with n > 1 (whichever NCCL
 * the process already holds is used; single-GPU use of the library needs no NCCL).  Philox keys are functions of the global
 * pixel: the image does not depend on n. */
int rt_init_multi(int n_devices, const int* device_ids, const RtContextOptions* options /* NULL: from the environment */, RtMulti** out);
void rt_shutdown_multi(RtMulti* multi);
int rt_multi_device_count(const RtMulti* multi);
RtContext* rt_multi_context(RtMulti* multi, int index); /* context of the index-th device (owned by `multi`) */
int rt_multi_scene_create(RtMulti* multi, const RtSceneDesc* desc, const RtSceneOptions* options, RtMultiScene** out);
void rt_multi_scene_destroy(RtMultiScene* scene);
/* rt_render over all devices: params->row_begin / row_stride must be 0 / 1; out_sum_rgb / out_sumsq_rgb are HOST [height][width][3].
 * stats: paths / rays / launches summed over devices, iterations = max, ms_total = device 0's clock from the start of the render to
 * the end of the de-interleave. */
int rt_render_multi(RtMultiScene* scene, const RtCamera* cam, const RtRenderParams* params, float* out_sum_rgb, float* out_sumsq_rgb,
                    RtStats* stats);

/* rt_render without any allocation.  No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments, image and
 * RtStats as rt_render (host output buffers; the Philox keys do not depend on queue size or lane count, so the sums agree with
 * rt_render's up to the order of the f64 accumulations), but ALL device scratch — ray queues of RT_NOALLOC_QUEUE_RAYS rays, hit and
 * finished-path records, counters, f64 accumulators, device framebuffers — is one static device array of RT_NOALLOC_SCRATCH_BYTES that
 * every GPU which loads the library holds until exit.  The call performs no cudaMalloc / cudaFree / cudaMallocHost and creates no stream
 * or event; scene and context must exist already (rt_init / rt_scene_create allocate).  Limits: width x rendered rows <=
 * RT_NOALLOC_MAX_PIXELS, else RT_ERR_INVALID with a message naming the limit; RT_RENDER_STAGE_TIMING / RT_RENDER_COUNT_WORK are ignored.
 * Every call on a device shares the one array (it keeps the previous call's contents; the call clears what must start at zero): calls
 * that overlap on the same GPU — from any context or thread — race.  One wavefront lane, small queues: slower than rt_render. */
#define RT_NOALLOC_MAX_PIXELS 65536u
#define RT_NOALLOC_QUEUE_RAYS 32768u
#define RT_NOALLOC_SCRATCH_BYTES (9u << 20)
int rt_render_noalloc(RtScene* scene, const RtCamera* cam, const RtRenderParams* params, float* out_sum_rgb, float* out_sumsq_rgb, RtStats* stats);

/* Introspection used by tests / the roofline accounting. */
int rt_scene_bvh_info(const RtScene* scene, uint32_t* n_nodes, uint32_t* n_always, uint32_t* max_depth);
/* Frame-rect list of a small all-rectangle scene (k_extend_rects): frames, frame boxes (faces of one box tested with one slab test)
 * and loose rectangles; all 0 when the scene does not use it. */
int rt_scene_frame_rect_info(const RtScene* scene, uint32_t* n_frames, uint32_t* n_boxes, uint32_t* n_loose);
/* How extend sees the scene: box primitives (six face rectangles of an axis-aligned Cube answered by one slab test), leaf items of the
 * tree (primitives + boxes - 6 x boxes), and the threads per block of the whole-tree-in-shared-memory form (0: not used). */
int rt_scene_extend_info(const RtScene* scene, uint32_t* n_box_prims, uint32_t* n_leaf_items, uint32_t* whole_tree_threads);
/* The same answer WITHOUT a device (host planning only: box detection + host SAH build; no CUDA call): what rt_scene_create_ex would
 * hand to extend.  n_nodes / bvh_depth are 0 for scenes whose tree would be built on the device (RT_BVH_DEVICE_LBVH). */
typedef struct RtScenePlan {
    uint32_t n_box_prims, n_leaf_items, n_nodes, bvh_depth, n_always;
    uint32_t whole_tree_threads;   /* 0: the tables do not fit one block's shared memory */
    uint64_t whole_tree_smem;      /* bytes of dynamic shared memory of that form */
} RtScenePlan;
int rt_scene_plan(const RtSceneDesc* desc, const RtSceneOptions* options, RtScenePlan* out);
/* Bytes of device memory the scene occupies (= host->device bytes rt_scene_create copied). */
uint64_t rt_scene_device_bytes(const RtScene* scene);
const char* rt_last_error(void);
uint32_t rt_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_H */
