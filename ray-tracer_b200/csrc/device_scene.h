// Device-resident scene layout (HBM) and wavefront state shared by the kernels and the C ABI.
//
// Layout decisions (DESIGN.md "Data layout in HBM"):
//  * BVH: 32-byte nodes of two float4 {lo.xyz, leftOrFirst} {hi.xyz, count}; the two children of an
//    interior node are adjacent (one 64-byte aligned pair), nodes are emitted breadth-first so the top
//    levels form a prefix that the extend kernel stages into shared memory.
//  * Primitives are reordered into leaf order.  `prim_head[i]` (float4) is all the extend kernel needs
//    for the common case — a translation-only sphere (cx, cy, cz, r>0).  Anything else (rectangles,
//    rotated / scaled spheres, media) stores w = -1, its kind in y and the index of its 128-byte general
//    record in x; rectangles also carry in z the index of a compact 64-byte record in `rects` (rows z, x, y of
//    the object<-world matrix + half extents) which is what the extend kernel tests.
//  * Scenes of at most RT_SMALL_SCENE primitives have no tree (n_small = n_prims, primitives stay in the
//    reference's DFS order) and are tested as a warp-uniform list out of shared memory.
//  * `prim_meta[i]` = {order (reference DFS rank), material index} is read by shade / on exact t ties.
//  * Materials are 32-byte records with the solid colour inlined (no texture hop for SolidColor).
#pragma once
#include <stdint.h>

#include <cuda_runtime.h>

#include "../../include/rt_b200.h"

// Bounds / invariant checks of the debug library (make librt_b200_check.so: -DRT_CHECK_BOUNDS).  compute-sanitizer is closed on the
// GPU pool this backend is developed on, so the kernels carry their own asserts: every queue slot, node, primitive, general record,
// rectangle, material, stack and framebuffer index is checked where it is formed; a violation traps the kernel (cudaErrorAssert) and
// the render returns RT_ERR_CUDA.  scripts/sanitize_target.py drives every kernel form through the checked library.
#ifdef RT_CHECK_BOUNDS
#include <cassert>
#define RT_ASSERT(cond) assert(cond)
#else
#define RT_ASSERT(cond) ((void)0)
#endif

namespace rtb {

#undef RT_NO_PRIM  // (the public header's macro of the same value; inside the library it is a typed constant)
constexpr uint32_t RT_NO_PRIM = 0xFFFFFFFFu;
constexpr int RT_STACK_DEPTH = 64;       // traversal stack entries (builder caps tree depth below this)
#ifndef RT_SMEM_STACK_N
#define RT_SMEM_STACK_N 24
#endif
constexpr uint32_t RT_SMEM_STACK = RT_SMEM_STACK_N;   // of which this many per thread live in shared memory (96 B/thread)
constexpr uint32_t RT_K_TAIL = 0x80000000u;  // flag on the iteration index handed to extend / shade: tail of a render, no k_generate (kernels.cu)
constexpr uint32_t RT_RING = 4;          // ring length of the per-iteration queue counters
constexpr uint32_t RT_MAX_STAGED_NODES = 2048;  // a BVH of at most this many nodes (64 KB) is staged whole in shared memory by extend
constexpr uint32_t RT_TOP_STAGED_NODES = 256;   // larger trees: only this breadth-first prefix (top 8 levels, 8 KB)
constexpr uint32_t RT_MAX_STAGED_MATS = 512;    // 16 KB of material records staged in shared memory by shade
constexpr uint32_t RT_MAX_STAGED_PRIMS = 1024;  // primitive heads (16 KB) staged by extend when the whole table fits
constexpr uint32_t RT_MAX_STAGED_RECTS = 256;   // compact rectangle records (16 KB) staged together with the heads
constexpr uint32_t RT_SMALL_SCENE = 32;         // at most this many primitives: linear list instead of a BVH
constexpr uint32_t RT_MAX_ALWAYS = 8;           // always-tested list (sky / ground / fog)
constexpr uint32_t RT_DEVICE_BUILD_MIN_PRIMS = 65536;  // from this many primitives the BVH is built on the GPU (lbvh.cu)
constexpr uint32_t RT_DPRIM_BOX = 16;            // device-only primitive kind (prim_head.y): a box of six face rectangles, record index in prim_head.x
constexpr uint32_t RT_MAX_FRAMES = 8;           // frame-rect list: distinct orthonormal frames the rectangles of a small scene are aligned to

struct DMaterial {       // 32 B, read as two float4
    uint32_t kind;       // RtMaterialKind
    int32_t texture;     // -1: solid colour inlined in rgb; else texture table index
    float fuzz, ior;
    float r, g, b, pad;
};

struct DScene {
    const float4* nodes;       // 2 float4 per node
    const float4* prim_head;   // per device primitive
    const uint2* prim_meta;    // {order, material (int32 bits)}
    const float4* rects;       // 4 float4 per rectangle: row z, row x, row y of object<-world (xyz, translation), (hx, hy, 0, 0)
    const float4* rect_nrm;    // per rectangle: world-space normal M (0,0,1) (forward matrix, not renormalised: SURVEY A.5)
    const RtPrim* gen;         // general records (non translation-only-sphere primitives), then the boundary surfaces of the RT_PRIM_MEDIUM
                               // records (RtSceneDesc.boundary_prims); a medium record's `aux` is the distance IN RECORDS to its first surface
    const float4* boxes;       // 2 float4 per box primitive (csrc/box_prims.h): (lo.xyz, first face primitive) (hi.xyz, 0); the six faces are the
                               // device primitives first .. first + 5 in the order x-, x+, y-, y+, z-, z+ — behind the leaf primitives, never in a leaf
    const DMaterial* mats;
    const RtTexture* texs;
    const float4* texels;
    // Blobs staged into shared memory by ONE bulk copy each (csrc/bulk_copy.cuh), laid out by rt_scene_create exactly as the kernels read them:
    //   oct_blob   (k_extend<., true>; null when the tree is not staged whole): [octant-ready pair records, 7 float4 per sibling pair
    //              (intersect.cuh node_step_oct)][prim_head n_leaf_prims][rects 4 n_leaf_rects][boxes 2 n_boxes]
    //   shade_blob (k_shade): [materials 2 float4 x min(n_mats, RT_MAX_STAGED_MATS)][prim_head n_staged_prims][rect_nrm][prim_meta uint2, padded to 16 B]
    const float4* oct_blob;
    const float4* shade_blob;
    uint32_t oct_blob4, shade_blob4;  // sizes in float4
    const uint32_t* order_to_prim;  // DFS rank (RtPrim.order) -> device primitive index, RT_NO_PRIM for ranks without a primitive
    uint32_t n_orders;
    uint32_t n_nodes, n_prims, n_always, n_gen, n_mats, n_texs;
    uint32_t root_is_leaf;     // BVH over < 2 primitives: node 0 is a leaf
    uint32_t bvh_depth;        // depth of the tree (levels below the root)
    uint32_t n_staged_nodes;   // prefix of `nodes` staged in shared memory
    uint32_t has_media;
    uint32_t has_general;      // some primitive is reached through its general 128-byte record (rotated / scaled / negative-radius sphere, medium);
                               // rectangles own a general record too but are tested through `rects` / `rect_nrm` and never read it
    uint32_t n_rects;
    uint32_t n_leaf_prims;     // primitives the tree (or the list) references: [0, n_leaf_prims); the faces of the box primitives follow
    uint32_t n_leaf_rects;     // rectangles among them: their compact records come first in `rects`
    uint32_t n_boxes;
    uint32_t n_staged_prims;   // n_prims when heads (and all rects) are staged in shared memory, else 0
    uint32_t n_small;          // > 0: no tree, test primitives [0, n_small) in order
    uint32_t small_all_rects;  // n_small > 0, every primitive is a rectangle and the records are staged: specialised loop
    float always_k[RT_MAX_ALWAYS];  // |centre|^2 - r^2 of the always-list spheres, rounded from f64 (big_sphere_hit)
    // Frame-rect list (k_extend_rects; small all-rectangle scenes whose transforms are rigid): the rectangles are grouped
    // by the orthonormal frame they are axis-aligned in (cornell-box: world + one frame per rotated cube).  Inside a frame,
    // rectangles that lie on the faces of one common box (each within its face) form a FRAME BOX: a Cube::new (geometry.rs:250-287)
    // in its own frame, or the four walls of a room.  A line crosses the surface of a convex box at most twice, so the closest hit
    // among the faces is the entry crossing if that face exists and the point is inside ITS rectangle, else the exit crossing —
    // one slab test instead of up to six rectangle tests, with every rectangle keeping its own plane, extents and `t >= 1e-6`.
    // The rest stay LOOSE rectangles, grouped by normal axis.
    // fr_blob = [3 float4 rows per frame]
    //           [fr_boxes x 10 float4: lo.xyz, hi.xyz, 6 x (centre a, centre b, half a, half b) for the faces x-, x+, y-, y+, z-, z+,
    //                                   6 primitive ids (RT_NO_PRIM: no rectangle on that face) + 2 pad]
    //           [fr_loose x float4 (plane w, centre a, centre b, primitive)] [fr_loose x float2 (half extent a, half extent b)]
    // boxes of frame f = fr_box_range[f] .. fr_box_range[f+1]; loose group (f, k) = records fr_range[3f+k] .. fr_range[3f+k+1].
    const float4* fr_blob;
    uint32_t fr_n;         // rectangles in the list (boxed + loose); > 0: extend uses k_extend_rects
    uint32_t fr_frames;
    uint32_t fr_identity;  // frame 0 is the world frame: no rotation of the ray
    uint32_t fr_boxes, fr_loose;
    uint32_t fr_blob4;     // float4 count of the blob
    uint32_t fr_box_range[RT_MAX_FRAMES + 1];
    uint32_t fr_range[RT_MAX_FRAMES * 3 + 1];
};

// One queue buffer of live rays, SoA of three float4 streams (48 B per ray):
//   a = origin.xyz, prev primitive (uint bits; RT_NO_PRIM = none)
//   b = direction.xyz, local pixel index (uint bits)
//   c = throughput.rgb, (sample << 8 | segment) (uint bits)
struct RayQueue {
    float4* a;
    float4* b;
    float4* c;
};

// Per-iteration counters live in a ring so no kernel ever has to reset a counter it also reads.
struct WaveCounters {
    uint32_t surv[RT_RING];    // rays appended by shade of iteration k-1 into queue k
    uint32_t total[RT_RING];   // surv + regenerated camera rays = size of queue k
    uint32_t done[RT_RING];    // finished-path records produced by shade k
    uint64_t issued[RT_RING];  // camera samples issued before iteration k
    uint32_t cursor[RT_RING];  // ray-index allocator of the persistent extend kernel of iteration k (chunks of RT_FEED_RANGE)
    // status block copied to the host when polling for termination
    uint64_t st_issued;
    uint64_t st_rays;
    uint32_t st_total;
    uint32_t st_iter;
    unsigned long long st_box_tests, st_prim_tests;  // RT_RENDER_COUNT_WORK only
};

struct RenderConst {
    RtCamera cam;
    uint32_t width, height, spp, max_depth;
    uint32_t seed_lo, seed_hi;
    uint32_t row_begin, row_stride, rows_local;
    uint32_t capacity;
    uint32_t pixel_major;   // path index = pixel * spp + sample (1) or sample * pixels_local + pixel (0)
    uint64_t paths_total;   // width * rows_local * spp
    uint64_t pixels_local;  // width * rows_local
    const RtRayIn* user_rays;  // rt_color_rays: path `pixel` starts from user_rays[pixel] instead of a camera ray (else null)
};

}  // namespace rtb
