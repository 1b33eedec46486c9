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
#ifndef RT_EXTEND_ALL_MIN_BLOCKS
#define RT_EXTEND_ALL_MIN_BLOCKS 2
#endif
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
#ifndef RT_SHADE_RAYS
#define RT_SHADE_RAYS 1   // rays shaded per thread per block round (one barrier pair / atomic pair per round)
#endif
#ifndef RT_REFILL_MIN
#define RT_REFILL_MIN 16     // persistent extend: idle lanes of a warp fetch new rays once at least this many are idle
#endif
#ifndef RT_PERSISTENT_MIN_NODES
#define RT_PERSISTENT_MIN_NODES 65536u  // BVH size from which extend runs in its persistent, lane-refilling form
#endif
#ifndef RT_SHADE_PIPE_DEFAULT
#define RT_SHADE_PIPE_DEFAULT 0  // 1: shade runs in its warp-specialised form (k_shade_pipe); RT_B200_SHADE_PIPE=0|1 overrides
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

// Must be called before the launches of a scene; `staged_bytes` = shared-memory bytes of the scene prefix staged by
// extend / trace, `fr_bytes` = bytes of the frame-rect table staged by k_extend_rects (0: none), `shade_bytes` = shared-memory bytes of the material / primitive tables staged by shade.
cudaError_t kernels_configure(int device, size_t staged_bytes, size_t shade_bytes, size_t fr_bytes);
// SM count only (kernels without dynamic shared memory).
cudaError_t kernels_init(int device);
int kernels_num_sms();

// upper bound of the rays left in the queue once every path has been issued (0xFFFFFFFF = unknown: full grids); per host thread
void kernels_set_work_hint(uint32_t rays_upper_bound);
void launch_generate(const RenderConst& rc, const RayQueue& q, WaveCounters* wc, uint32_t k, cudaStream_t st);
void launch_extend(const DScene& S, const RayQueue& q, float2* hits, WaveCounters* wc, uint32_t k, uint32_t seed_lo, uint32_t seed_hi,
                   bool count_work, cudaStream_t st);
void launch_shade(const DScene& S, const RenderConst& rc, const RayQueue& qin, const RayQueue& qout, const float2* hits, float4* done,
                  WaveCounters* wc, uint32_t k, cudaStream_t st);
void launch_accumulate(const RenderConst& rc, const float4* done, const WaveCounters* wc, uint32_t k, double* sum, double* sumsq, cudaStream_t st);
void launch_finish(const double* acc, size_t n, float* out, cudaStream_t st);
void launch_resolve_u8(const float* sum, uint32_t width, uint32_t height, uint32_t spp, uint8_t* out, cudaStream_t st);
void launch_deinterleave(const float* gathered, float* full, uint32_t width, uint32_t height, uint32_t n_shards, cudaStream_t st);
void launch_trace(const DScene& S, const RtRayIn* rays, uint64_t n, uint32_t seed_lo, uint32_t seed_hi, RtHitOut* out, cudaStream_t st);

}  // namespace rtb
