// Wavefront path tracer for sm_100a: one kernel per stage — generate, extend, shade, accumulate —
// plus resolve (tonemap) and trace (closest-hit on caller rays).  Replaces the recursive
// render::color loop (src/render.rs:5-29) and the example drivers' pixel loop
// (examples/book-one.rs:62-78).  See DESIGN.md for the queue protocol and the roofline accounting.
//
// Queue protocol (iteration k, ring slot s = k % RT_RING, ray buffers ping-pong on k % 2):
//   generate_k   appends min(capacity - surv[s], paths left) camera rays behind the surv[s] survivors
//                that shade_{k-1} compacted into buffer k%2; thread 0 publishes total[s], issued[s+1]
//                and zeroes the counters the NEXT iterations will increment.
//   extend_k     closest hit for rays [0, total[s]) -> hits[i] = (t, prim).
//   shade_k      scatter / emit; survivors are compacted into buffer (k+1)%2 through warp-aggregated
//                atomics on surv[s+1]; finished paths with non-zero radiance go to the done list (done[s]).
//   accumulate_k adds the done list into the per-pixel sum / sum-of-squares framebuffer.
// Every kernel is a persistent grid-stride loop over a count it reads from device memory, so the
// host never has to know queue sizes; it polls a status block every few iterations to detect the end.
#include <cuda_runtime.h>

#include "device_scene.h"
#include "intersect.cuh"
#include "kernels.h"
#include "shade.cuh"

namespace rtb {

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

// Warp-aggregated append: one atomic per warp, returns this lane's slot (valid where `pred`).
__device__ __forceinline__ uint32_t warp_append(uint32_t* counter, bool pred) {
    const uint32_t mask = __ballot_sync(0xFFFFFFFFu, pred);
    if (mask == 0u) return 0u;
    const uint32_t lane = lane_id();
    const int leader = __ffs(mask) - 1;
    uint32_t base = 0u;
    if ((int)lane == leader) base = atomicAdd(counter, (uint32_t)__popc(mask));
    base = __shfl_sync(0xFFFFFFFFu, base, leader);
    return base + (uint32_t)__popc(mask & ((1u << lane) - 1u));
}

// Block-aggregated append of two streams (survivors, finished paths): ONE pair of global atomics per block
// iteration.  With one atomic per warp, the 2 x 32 k same-address atomics of a launch serialise in one L2 slice
// and cost about half of the shade kernel (ncu: 20 % of its stall samples sat on that ATOMG).
// Must be called by every thread of the block.  `s_cnt` = [RT_SHADE_WARPS + 1][2] shared words.
__device__ __forceinline__ void block_append2(uint32_t* cnt_a, bool pa, uint32_t* cnt_b, bool pb, uint32_t& slot_a, uint32_t& slot_b,
                                              uint32_t (*s_cnt)[2]) {
    const uint32_t lane = lane_id(), warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    const uint32_t ma = __ballot_sync(0xFFFFFFFFu, pa), mb = __ballot_sync(0xFFFFFFFFu, pb);
    if (lane == 0) { s_cnt[warp][0] = (uint32_t)__popc(ma); s_cnt[warp][1] = (uint32_t)__popc(mb); }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t ta = 0u, tb = 0u;
        for (uint32_t w = 0; w < n_warps; ++w) {
            const uint32_t ca = s_cnt[w][0], cb = s_cnt[w][1];
            s_cnt[w][0] = ta; s_cnt[w][1] = tb;
            ta += ca; tb += cb;
        }
        s_cnt[n_warps][0] = ta ? atomicAdd(cnt_a, ta) : 0u;
        s_cnt[n_warps][1] = tb ? atomicAdd(cnt_b, tb) : 0u;
    }
    __syncthreads();
    const uint32_t lt = (1u << lane) - 1u;
    slot_a = s_cnt[n_warps][0] + s_cnt[warp][0] + (uint32_t)__popc(ma & lt);
    slot_b = s_cnt[n_warps][1] + s_cnt[warp][1] + (uint32_t)__popc(mb & lt);
}

// local linear index -> (x, local row): 8x4 pixel tiles per warp where they fit, scanline remainders
__device__ __forceinline__ void pixel_from_index(uint32_t q, uint32_t W, uint32_t rows, uint32_t& x, uint32_t& yl) {
    const uint32_t W8 = W & ~7u, R4 = rows & ~3u;
    const uint32_t nA = W8 * R4;
    if (q < nA) {
        const uint32_t tile = q >> 5, within = q & 31u, tiles_x = W8 >> 3;
        const uint32_t ty = tile / tiles_x, tx = tile - ty * tiles_x;
        x = tx * 8u + (within & 7u);
        yl = ty * 4u + (within >> 3);
        return;
    }
    q -= nA;
    const uint32_t wB = W - W8, nB = wB * R4;
    if (q < nB) {
        yl = q / wB;
        x = W8 + (q - yl * wB);
        return;
    }
    q -= nB;
    const uint32_t r = q / W;
    yl = R4 + r;
    x = q - r * W;
}

// ---- generate --------------------------------------------------------------------------------------

__global__ void __launch_bounds__(256) k_generate(const RenderConst rc, RayQueue q, WaveCounters* __restrict__ wc, uint32_t k) {
    const uint32_t s = k % RT_RING;
    const uint32_t surv = wc->surv[s];
    const uint64_t issued = wc->issued[s];
    const uint64_t left = rc.paths_total - issued;
    const uint32_t room = rc.capacity - surv;
    const uint32_t n_new = left < (uint64_t)room ? (uint32_t)left : room;
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
    for (uint32_t i = tid; i < n_new; i += stride) {
        const uint64_t path = issued + i;
        const uint32_t sample = (uint32_t)(path / rc.pixels_local);
        const uint32_t qi = (uint32_t)(path - (uint64_t)sample * rc.pixels_local);
        uint32_t x, yl;
        pixel_from_index(qi, rc.width, rc.rows_local, x, yl);
        const uint32_t y = rc.row_begin + yl * rc.row_stride;
        const uint32_t gp = y * rc.width + x;
        const Philox4 rnd = philox4x32_10(0u, RT_DIM_CAMERA, rc.seed_lo, rc.seed_hi, gp, sample);
        float3 o, d;
        camera_ray(rc.cam, x, y, rc.width, rc.height, rnd, o, d);
        const uint32_t slot = surv + i;
        q.a[slot] = make_float4(o.x, o.y, o.z, __uint_as_float(RT_NO_PRIM));
        q.b[slot] = make_float4(d.x, d.y, d.z, __uint_as_float(gp));
        q.c[slot] = make_float4(1.0f, 1.0f, 1.0f, __uint_as_float((sample << 8) | 1u));
    }
    if (tid == 0) {
        const uint32_t total = surv + n_new;
        wc->total[s] = total;
        wc->issued[(k + 1) % RT_RING] = issued + n_new;
        wc->surv[(k + 2) % RT_RING] = 0u;
        wc->done[(k + 1) % RT_RING] = 0u;
        wc->st_issued = issued + n_new;
        wc->st_rays += total;
        wc->st_total = total;
        wc->st_iter = k;
    }
}

// ---- extend ----------------------------------------------------------------------------------------

// Copy the BVH prefix (and, for scenes that fit, the primitive heads and compact rectangle records) into shared
// memory: [stack RT_SMEM_STACK x blockDim words][nodes 2*n_staged_nodes][heads n_staged_prims][rects 4*n_rects].
__device__ __forceinline__ Staged stage_scene(const DScene& S, float4* smem_base) {
    float4* smem4 = smem_base + (RT_SMEM_STACK * blockDim.x) / 4u;
    const uint32_t n_node4 = 2u * S.n_staged_nodes, n_head4 = S.n_staged_prims, n_rect4 = S.n_staged_prims ? 4u * S.n_rects : 0u;
    for (uint32_t i = threadIdx.x; i < n_node4; i += blockDim.x) smem4[i] = S.nodes[i];
    for (uint32_t i = threadIdx.x; i < n_head4; i += blockDim.x) smem4[n_node4 + i] = S.prim_head[i];
    for (uint32_t i = threadIdx.x; i < n_rect4; i += blockDim.x) smem4[n_node4 + n_head4 + i] = S.rects[i];
    __syncthreads();
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(smem_base);
    const uint32_t scene = base + RT_SMEM_STACK * blockDim.x * 4u;
    return Staged{scene, scene + 16u * n_node4, scene + 16u * (n_node4 + n_head4), base + 4u * threadIdx.x};
}

template <bool COUNT>
__global__ void __launch_bounds__(RT_EXTEND_BLOCK, RT_EXTEND_MIN_BLOCKS) k_extend(const DScene S, const RayQueue q, float2* __restrict__ hits, WaveCounters* __restrict__ wc,
                                                          uint32_t k, uint32_t seed_lo, uint32_t seed_hi) {
    extern __shared__ float4 smem4[];
    const Staged sm = stage_scene(S, smem4);
    const uint32_t n = wc->total[k % RT_RING];
    const uint32_t stride = gridDim.x * blockDim.x;
    TraversalWork work{0u, 0u};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float4 a = q.a[i], b = q.b[i];
        MediumKey key{0u, 0u, 0u, seed_lo, seed_hi};
        if (S.has_media) {
            const uint32_t sb = __float_as_uint(q.c[i].w);
            key.k0 = __float_as_uint(b.w);
            key.k1 = sb >> 8;
            key.segment = sb & 0xFFu;
        }
        const ClosestHit h = closest_hit<COUNT>(S, sm, f3(a.x, a.y, a.z), f3(b.x, b.y, b.z), __float_as_uint(a.w), key, &work);
        hits[i] = make_float2(h.t, __uint_as_float(h.prim));
    }
    if (COUNT) {  // instrumented build only (RT_RENDER_COUNT_WORK): one atomic pair per warp
        uint32_t bt = work.box_tests, pt = work.prim_tests;
        for (int off = 16; off > 0; off >>= 1) { bt += __shfl_down_sync(0xFFFFFFFFu, bt, off); pt += __shfl_down_sync(0xFFFFFFFFu, pt, off); }
        if (lane_id() == 0) { atomicAdd(&wc->st_box_tests, (unsigned long long)bt); atomicAdd(&wc->st_prim_tests, (unsigned long long)pt); }
    }
}

// ---- shade -----------------------------------------------------------------------------------------

__global__ void __launch_bounds__(RT_SHADE_BLOCK, RT_SHADE_MIN_BLOCKS) k_shade(const DScene S, const RenderConst rc, const RayQueue qin, RayQueue qout,
                                               const float2* __restrict__ hits, float4* __restrict__ done, WaveCounters* __restrict__ wc, uint32_t k) {
    __shared__ float4 smats[2 * RT_MAX_STAGED_MATS];
    __shared__ uint32_t s_cnt[RT_SHADE_BLOCK / 32 + 1][2];
    const uint32_t n_staged = min(S.n_mats, RT_MAX_STAGED_MATS);
    for (uint32_t i = threadIdx.x; i < 2u * n_staged; i += blockDim.x) smats[i] = reinterpret_cast<const float4*>(S.mats)[i];
    __syncthreads();
    const uint32_t s = k % RT_RING;
    const uint32_t n = wc->total[s];
    uint32_t* surv_next = &wc->surv[(k + 1) % RT_RING];
    uint32_t* done_cnt = &wc->done[s];
    const uint32_t stride = gridDim.x * blockDim.x;
    // whole blocks iterate together: the append below synchronises the block
    for (uint32_t base = blockIdx.x * blockDim.x; base < n; base += stride) {
        const uint32_t i = base + threadIdx.x;
        bool alive = false, record = false;
        float4 na, nb, nc, rec;
        if (i < n) {
            const float4 a = qin.a[i], b = qin.b[i], c = qin.c[i];
            const float2 h = hits[i];
            const uint32_t prim = __float_as_uint(h.y);
            if (prim != RT_NO_PRIM) {  // miss -> black (render.rs:26)
                const uint2 meta = __ldg(&S.prim_meta[prim]);
                const int32_t mi = (int32_t)meta.y;
                if (mi >= 0) {  // no material -> black (render.rs:18-20)
                    DMaterial m;
                    {
                        const float4* src = ((uint32_t)mi < n_staged) ? &smats[2 * mi] : reinterpret_cast<const float4*>(S.mats) + 2 * (size_t)mi;
                        const float4 m0 = src[0], m1 = src[1];
                        m.kind = __float_as_uint(m0.x); m.texture = (int32_t)__float_as_uint(m0.y); m.fuzz = m0.z; m.ior = m0.w;
                        m.r = m1.x; m.g = m1.y; m.b = m1.z; m.pad = 0.0f;
                    }
                    const uint32_t sb = __float_as_uint(c.w), gp = __float_as_uint(b.w);
                    const uint32_t sample = sb >> 8, segment = sb & 0xFFu;
                    const float3 o = f3(a.x, a.y, a.z), d = f3(b.x, b.y, b.z), thr = f3(c.x, c.y, c.z);
                    const MediumKey key{gp, sample, segment, rc.seed_lo, rc.seed_hi};
                    const HitDetail hd = hit_detail(S, prim, o, d, h.x, m.texture >= 0, key);
                    const Philox4 rnd = philox4x32_10(segment, RT_DIM_SHADE, rc.seed_lo, rc.seed_hi, gp, sample);
                    const ScatterResult sr = shade_material(S, m, d, hd, rnd);
                    if (sr.scattered) {
                        if (segment < rc.max_depth) {  // color(.., 0) returns black without tracing (render.rs:6-8)
                            const float3 t2 = thr * sr.attenuation;
                            const float4 head = __ldg(&S.prim_head[prim]);  // general records carry their kind in .y
                            const bool is_medium = head.w <= 0.0f && __float_as_uint(head.y) == (uint32_t)RT_PRIM_MEDIUM_SPHERE;
                            na = make_float4(hd.p.x, hd.p.y, hd.p.z, __uint_as_float(is_medium ? RT_NO_PRIM : prim));
                            nb = make_float4(sr.dir.x, sr.dir.y, sr.dir.z, b.w);
                            nc = make_float4(t2.x, t2.y, t2.z, __uint_as_float((sample << 8) | (segment + 1u)));
                            alive = true;
                        }
                    } else if (m.kind == RT_MAT_DIFFUSE_LIGHT) {
                        const float3 L = thr * sr.attenuation;
                        if (L.x != 0.0f || L.y != 0.0f || L.z != 0.0f) {
                            rec = make_float4(L.x, L.y, L.z, b.w);
                            record = true;
                        }
                    }
                }
            }
        }
        uint32_t slot, dslot;
        block_append2(surv_next, alive, done_cnt, record, slot, dslot, s_cnt);
        if (alive) { qout.a[slot] = na; qout.b[slot] = nb; qout.c[slot] = nc; }
        if (record) done[dslot] = rec;
    }
}

// ---- accumulate ------------------------------------------------------------------------------------

__global__ void __launch_bounds__(256) k_accumulate(const RenderConst rc, const float4* __restrict__ done, const WaveCounters* __restrict__ wc,
                                                    uint32_t k, double* __restrict__ sum, double* __restrict__ sumsq) {
    // f64 accumulators (native 64-bit atomics): an f32 running sum of 1024+ equal sky samples rounds the same way on
    // every add and drifts by 1e-4 relative, which the 3-sigma image test sees on zero-variance pixels.
    const uint32_t n = wc->done[k % RT_RING];
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float4 r = done[i];
        const uint32_t gp = __float_as_uint(r.w);
        const uint32_t y = gp / rc.width, x = gp - y * rc.width;
        const uint32_t yl = (y - rc.row_begin) / rc.row_stride;
        const size_t o = ((size_t)yl * rc.width + x) * 3;
        atomicAdd(&sum[o + 0], (double)r.x);
        atomicAdd(&sum[o + 1], (double)r.y);
        atomicAdd(&sum[o + 2], (double)r.z);
        if (sumsq) {
            atomicAdd(&sumsq[o + 0], (double)r.x * (double)r.x);
            atomicAdd(&sumsq[o + 1], (double)r.y * (double)r.y);
            atomicAdd(&sumsq[o + 2], (double)r.z * (double)r.z);
        }
    }
}

// last step of the accumulate stage: f64 accumulators -> the caller's f32 sum buffers
__global__ void __launch_bounds__(256) k_finish(const double* __restrict__ acc, size_t n, float* __restrict__ out) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) out[i] = (float)acc[i];
}

// ---- resolve (examples/book-one.rs:90-100) -----------------------------------------------------------

__global__ void k_resolve_u8(const float* __restrict__ sum, uint32_t width, uint32_t height, float inv_spp, uint8_t* __restrict__ out) {
    const size_t n = (size_t)width * height * 3;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t c = (uint32_t)(i % 3);
        const size_t pix = i / 3;
        const uint32_t row = (uint32_t)(pix / width), x = (uint32_t)(pix - (size_t)row * width);
        const uint32_t y = height - 1u - row;  // rows top -> bottom
        float v = sqrtf(sum[((size_t)y * width + x) * 3 + c] * inv_spp) * 255.0f;
        v = (v != v) ? 255.0f : fminf(v, 255.0f);  // f64::min keeps the non-NaN operand (A.12)
        out[i] = (uint8_t)fmaxf(v, 0.0f);
    }
}

// ---- de-interleave of gathered row shards (multi-GPU framebuffer reassembly) ----------------------------

__global__ void __launch_bounds__(256) k_deinterleave(const float* __restrict__ gathered, float* __restrict__ full, uint32_t width, uint32_t height,
                                                      uint32_t n_shards, uint32_t rows_padded) {
    const size_t row_elems = (size_t)width * 3, n = row_elems * height;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t y = (uint32_t)(i / row_elems);
        const size_t within = i - (size_t)y * row_elems;
        const uint32_t shard = y % n_shards, k = y / n_shards;
        full[i] = gathered[((size_t)shard * rows_padded + k) * row_elems + within];
    }
}

// ---- trace (parity harness; batched `world.hit(&ray)`) -------------------------------------------------

__global__ void __launch_bounds__(RT_EXTEND_BLOCK) k_trace(const DScene S, const RtRayIn* __restrict__ rays, uint64_t n, uint32_t seed_lo,
                                                         uint32_t seed_hi, RtHitOut* __restrict__ out) {
    extern __shared__ float4 smem4[];
    const Staged sm = stage_scene(S, smem4);
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const RtRayIn r = rays[i];
        const float3 o = f3(r.origin[0], r.origin[1], r.origin[2]), d = f3(r.direction[0], r.direction[1], r.direction[2]);
        const MediumKey key{(uint32_t)i, (uint32_t)(i >> 32), 1u, seed_lo, seed_hi};
        const ClosestHit h = closest_hit<false>(S, sm, o, d, RT_NO_PRIM, key);
        RtHitOut res;
        res.hit = 0; res.prim = -1; res.t = 0.0f; res.material = -1;
        res.p[0] = res.p[1] = res.p[2] = res.n[0] = res.n[1] = res.n[2] = res.uv[0] = res.uv[1] = 0.0f;
        if (h.prim != RT_NO_PRIM) {
            const uint2 meta = __ldg(&S.prim_meta[h.prim]);
            const HitDetail hd = hit_detail(S, h.prim, o, d, h.t, true, key);
            res.hit = 1; res.prim = (int32_t)meta.x; res.t = hd.t; res.material = (int32_t)meta.y;
            res.p[0] = hd.p.x; res.p[1] = hd.p.y; res.p[2] = hd.p.z;
            res.n[0] = hd.n.x; res.n[1] = hd.n.y; res.n[2] = hd.n.z;
            res.uv[0] = hd.u; res.uv[1] = hd.v;
        }
        out[i] = res;
    }
}

// ---- launch wrappers ---------------------------------------------------------------------------------

static int g_num_sms = 0;
static int g_extend_blocks_per_sm = 0, g_trace_blocks_per_sm = 0, g_shade_blocks_per_sm = 0;
static size_t g_extend_smem = ~(size_t)0;

cudaError_t kernels_init(int device) {
    if (g_num_sms > 0) return cudaSuccess;
    int sms = 0;
    const cudaError_t e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess) return e;
    g_num_sms = sms;
    const cudaError_t e2 = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_shade_blocks_per_sm, k_shade, RT_SHADE_BLOCK, 0);
    if (e2 != cudaSuccess) return e2;
    if (g_shade_blocks_per_sm < 1) g_shade_blocks_per_sm = 1;
    return cudaSuccess;
}

cudaError_t kernels_configure(int device, size_t scene_bytes) {
    cudaError_t e = kernels_init(device);
    if (e != cudaSuccess) return e;
    const size_t staged_bytes = scene_bytes + (size_t)RT_SMEM_STACK * RT_EXTEND_BLOCK * sizeof(uint32_t);
    if (staged_bytes != g_extend_smem) {
        e = cudaFuncSetAttribute(k_extend<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)staged_bytes);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(k_extend<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)staged_bytes);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(k_trace, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)staged_bytes);
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_extend_blocks_per_sm, k_extend<false>, RT_EXTEND_BLOCK, staged_bytes);
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_trace_blocks_per_sm, k_trace, RT_EXTEND_BLOCK, staged_bytes);
        if (e != cudaSuccess) return e;
        if (g_extend_blocks_per_sm < 1) g_extend_blocks_per_sm = 1;
        if (g_trace_blocks_per_sm < 1) g_trace_blocks_per_sm = 1;
        g_extend_smem = staged_bytes;
    }
    return cudaSuccess;
}

int kernels_num_sms() { return g_num_sms; }

void launch_generate(const RenderConst& rc, const RayQueue& q, WaveCounters* wc, uint32_t k, cudaStream_t st) {
    k_generate<<<g_num_sms * 4, 256, 0, st>>>(rc, q, wc, k);
}
void launch_extend(const DScene& S, const RayQueue& q, float2* hits, WaveCounters* wc, uint32_t k, uint32_t seed_lo, uint32_t seed_hi,
                   bool count_work, cudaStream_t st) {
    if (count_work) k_extend<true><<<g_num_sms * g_extend_blocks_per_sm, RT_EXTEND_BLOCK, g_extend_smem, st>>>(S, q, hits, wc, k, seed_lo, seed_hi);
    else k_extend<false><<<g_num_sms * g_extend_blocks_per_sm, RT_EXTEND_BLOCK, g_extend_smem, st>>>(S, q, hits, wc, k, seed_lo, seed_hi);
}
void launch_shade(const DScene& S, const RenderConst& rc, const RayQueue& qin, const RayQueue& qout, const float2* hits, float4* done,
                  WaveCounters* wc, uint32_t k, cudaStream_t st) {
    k_shade<<<g_num_sms * g_shade_blocks_per_sm, RT_SHADE_BLOCK, 0, st>>>(S, rc, qin, qout, hits, done, wc, k);
}
void launch_accumulate(const RenderConst& rc, const float4* done, const WaveCounters* wc, uint32_t k, double* sum, double* sumsq, cudaStream_t st) {
    k_accumulate<<<g_num_sms * 2, 256, 0, st>>>(rc, done, wc, k, sum, sumsq);
}
void launch_finish(const double* acc, size_t n, float* out, cudaStream_t st) { k_finish<<<g_num_sms * 4, 256, 0, st>>>(acc, n, out); }
void launch_resolve_u8(const float* sum, uint32_t width, uint32_t height, uint32_t spp, uint8_t* out, cudaStream_t st) {
    k_resolve_u8<<<g_num_sms * 4, 256, 0, st>>>(sum, width, height, 1.0f / (float)spp, out);
}
void launch_deinterleave(const float* gathered, float* full, uint32_t width, uint32_t height, uint32_t n_shards, cudaStream_t st) {
    const uint32_t rows_padded = (height + n_shards - 1) / n_shards;
    k_deinterleave<<<g_num_sms * 4, 256, 0, st>>>(gathered, full, width, height, n_shards, rows_padded);
}
void launch_trace(const DScene& S, const RtRayIn* rays, uint64_t n, uint32_t seed_lo, uint32_t seed_hi, RtHitOut* out, cudaStream_t st) {
    k_trace<<<g_num_sms * g_trace_blocks_per_sm, RT_EXTEND_BLOCK, g_extend_smem, st>>>(S, rays, n, seed_lo, seed_hi, out);
}

}  // namespace rtb
