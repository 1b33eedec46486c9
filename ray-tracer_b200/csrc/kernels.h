// Launch wrappers of the wavefront kernels (kernels.cu), called by the C ABI (capi.cu).
#pragma once
#include <cuda_runtime.h>

#include "device_scene.h"

#ifndef RT_EXTEND_BLOCK
#define RT_EXTEND_BLOCK 256
// k_extend<.., true> (tree in shared memory, octant-ready records): the tree copy dominates a block's shared memory, so larger blocks
// buy resident warps: 2 x 512 threads at <= 64 registers = 32 warps per SM against 24 for 3 x 256 at 74 registers
#ifndef RT_EXTEND_ALL_BLOCK
#define RT_EXTEND_ALL_BLOCK 512
#endif
// ... and ONE block of 1024 threads (the same 32 warps, one copy of the tables) for scenes whose tables need more than half an SM's
// shared memory (book-two with its ground blocks as box primitives: ~1 800 nodes, 1 400 heads, 400 boxes).  Both forms run the same
// kernel: __launch_bounds__(1024, 1) and (512, 2) cap it at the same 64 registers.
#define RT_EXTEND_ALL_BIG_BLOCK 1024
#endif
// k_shade: 8 blocks of 128 threads per SM.  The two barriers per round wait for the slowest of a block's warps (divergent material
// code) and for the slot-allocation atomics: 4 warps per barrier instead of 8 is +3.5 % (book-one) to +5 % (cornell) on the kernel,
// 2 warps per barrier (64 x 16) loses to the doubled atomics and staging prologues (profiles/r3c_*, r3d_*).
#ifndef RT_SHADE_BLOCK
#define RT_SHADE_BLOCK 128
#endif
#ifndef RT_SHADE_MIN_BLOCKS
#define RT_SHADE_MIN_BLOCKS 8
#endif
#ifndef RT_REFILL_MIN
#define RT_REFILL_MIN 16     // persistent extend: idle lanes of a warp fetch new rays once at least this many are idle
#endif
#ifndef RT_PERSISTENT_MIN_NODES
#define RT_PERSISTENT_MIN_NODES 65536u  // BVH size from which extend runs in its persistent, lane-refilling form
#endif
#ifndef RT_LEAF_MIN
#define RT_LEAF_MIN 8        // persistent extend: pending leaf ranges are intersected once this many lanes hold one
#endif
#ifndef RT_FEED_RANGE
#define RT_FEED_RANGE 128    // ray indices a warp claims per global atomic (multiple of 32)
#endif
#ifndef RT_BVH_MIN_BLOCKS
#define RT_BVH_MIN_BLOCKS 3
#endif
#ifndef RT_RECTS_MIN_BLOCKS
#define RT_RECTS_MIN_BLOCKS 6
#endif
#ifndef RT_EXTEND_MIN_BLOCKS
#define RT_EXTEND_MIN_BLOCKS 3
#endif

namespace rtb {

constexpr uint32_t RT_WORK_UNKNOWN = 0xFFFFFFFFu;

// Tuning options of a context.  rt_init reads the RT_B200_* environment variables ONCE into this struct (RtInitOptions of the C ABI
// overrides them); nothing is re-read per render.
#ifndef RT_TAIL_MODE
#define RT_TAIL_MODE 1  // 0: keep launching k_generate while the render drains (A/B of the tail iterations)
#endif
// shared memory of k_extend<., true> at `block` threads; whether the scene's tables fit one of its two block forms
size_t extend_all_smem_bytes(const DScene& S, uint32_t block);
bool extend_all_smem_fits(const DScene& S);
struct LaunchOptions {
    int persistent = 2;   // extend in its persistent lane-refilling form: 0 never, 1 always (where a tree exists), 2 from RT_PERSISTENT_MIN_NODES nodes
    int fuse_rects = 1;   // small all-rectangle scenes: extend runs inside k_shade (k_shade<., ., true>) instead of as k_extend_rects; 2 = two rays per round (A/B)
    int shade_rays2 = 1;  // two rays per thread per barrier round in k_shade on small all-rectangle scenes
    int gen_blocks = 4, acc_blocks = 8;  // blocks per SM of k_generate / k_accumulate
};

// Launch configuration of ONE scene on ONE device: dynamic shared-memory bytes of the tables each kernel stages and the resident
// blocks per SM at those sizes.  Owned by the scene (capi.cu RtScene), filled by kernels_configure_scene.
struct SceneLaunch {
    int device = -1, num_sms = 0;
    size_t scene_smem = 0, extend_smem = 0, trace_smem = 0, bvh_smem = 0, shade_smem = 0, rects_smem = 0, all_smem = 0;
    size_t shade_fr_smem = 0;
    int shade_fr_blocks = 1, shade2_fr_blocks = 1;
    bool fused_rects = false;  // the scene can run extend fused into shade (frame-rect list, no general records)
    int all_block = RT_EXTEND_ALL_BLOCK;  // threads per block of k_extend<., true>: 512 (two blocks per SM) or RT_EXTEND_ALL_BIG_BLOCK (one)
    int extend_blocks = 1, trace_blocks = 1, bvh_blocks = 1, shade_blocks = 1, shade2_blocks = 1, rects_blocks = 1, all_blocks = 1;
    bool gen = true;          // the scene has general 128-byte records: GEN = true instantiations of extend / shade
    bool all_staged = false;  // k_extend<., true>: whole tree + heads in shared memory, octant-ready records
};

// One launch site: whose scene, which options, on which stream, and (tail of a render) an upper bound of the rays left in the queue.
struct Launch {
    const SceneLaunch* sl;
    const LaunchOptions* opt;
    uint32_t work_hint;  // RT_WORK_UNKNOWN: full grids
    cudaStream_t st;
};

// `device` must be current in the calling thread.  staged_bytes = shared-memory bytes of the scene prefix staged by extend / trace,
// shade_bytes = material / primitive tables staged by shade, fr_bytes = frame-rect table of k_extend_rects (0: none).
cudaError_t kernels_configure_scene(int device, const DScene& S, size_t staged_bytes, size_t shade_bytes, size_t fr_bytes, SceneLaunch& out);
cudaError_t kernels_num_sms(int device, int* sms);

void launch_generate(const Launch& L, const RenderConst& rc, const RayQueue& q, WaveCounters* wc, uint32_t k);
void launch_extend(const Launch& L, const DScene& S, const RayQueue& q, float2* hits, WaveCounters* wc, uint32_t k, uint32_t seed_lo, uint32_t seed_hi,
                   bool count_work);
void launch_shade(const Launch& L, const DScene& S, const RenderConst& rc, const RayQueue& qin, const RayQueue& qout, const float2* hits, float4* done,
                  WaveCounters* wc, uint32_t k, bool count_work = false);
// extend runs inside k_shade for this scene / these options (launch_extend is then a no-op)
bool extend_is_fused(const Launch& L, const DScene& S, bool count_work);
void launch_accumulate(const Launch& L, const RenderConst& rc, const float4* done, const WaveCounters* wc, uint32_t k, double* sum, double* sumsq);
void launch_finish(int sms, const double* acc, size_t n, float* out, cudaStream_t st);
void launch_resolve_u8(int sms, const float* sum, uint32_t width, uint32_t height, uint32_t spp, uint8_t* out, cudaStream_t st);
void launch_deinterleave(int sms, const float* gathered, float* full, uint32_t width, uint32_t height, uint32_t n_shards, cudaStream_t st);
void launch_trace(const SceneLaunch& sl, const DScene& S, const RtRayIn* rays, const uint32_t* self_order, uint64_t n, uint32_t seed_lo, uint32_t seed_hi,
                  RtHitOut* out, cudaStream_t st);

}  // namespace rtb
