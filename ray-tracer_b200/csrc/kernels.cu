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

#include <mutex>

#include "bulk_copy.cuh"
#include "device_scene.h"
#include "intersect.cuh"
#include "kernels.h"
#include "shade.cuh"

namespace rtb {

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

// Tail of a render (every path has been issued; the host knows it from its status poll): k_generate would only roll the counters, and
// as the first of three dependent launches per iteration it is a third of the latency chain that the ~50 draining iterations of a
// depth-50 render consist of.  The host then stops launching it and passes k | RT_K_TAIL: the queue size IS the survivor count, and
// thread 0 of block 0 of the extend kernel rolls the ring (it runs after shade of k - 1 and before shade of k, where k_generate sat).
__device__ __forceinline__ uint32_t queue_size(const WaveCounters* __restrict__ wc, uint32_t kf) {
    const uint32_t s = (kf & ~RT_K_TAIL) % RT_RING;
    return (kf & RT_K_TAIL) ? wc->surv[s] : wc->total[s];
}
__device__ __forceinline__ void tail_roll(WaveCounters* __restrict__ wc, uint32_t kf, uint32_t n) {
    if (!(kf & RT_K_TAIL) || blockIdx.x != 0u || threadIdx.x != 0u) return;
    const uint32_t k = kf & ~RT_K_TAIL;
    wc->issued[(k + 1) % RT_RING] = wc->issued[k % RT_RING];
    wc->surv[(k + 2) % RT_RING] = 0u;
    wc->done[(k + 1) % RT_RING] = 0u;
    wc->cursor[(k + 1) % RT_RING] = 0u;
    wc->st_rays += n;
    wc->st_total = n;
    wc->st_iter = k;
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

// L2 prefetch of the queue entries a persistent grid-stride loop will read in its NEXT round (the 4 M-ray queues stream from
// HBM): one lane per 128-byte line issues `prefetch.global.L2`, no registers are held across the round.  k_shade +2 % (r1s);
// k_extend_rects, whose 48 resident warps already cover the latency, lost 2 % and does not prefetch.
#ifndef RT_PREFETCH_L2
#define RT_PREFETCH_L2 1
#endif
__device__ __forceinline__ void prefetch_l2(const void* p) {
#if RT_PREFETCH_L2
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#endif
}

// ---- generate --------------------------------------------------------------------------------------

__global__ void __launch_bounds__(256) k_generate(const RenderConst rc, RayQueue q, WaveCounters* __restrict__ wc, uint32_t k) {
    const uint32_t s = k % RT_RING;
    const uint32_t surv = wc->surv[s];
    const uint64_t issued = wc->issued[s];
    const uint64_t left = rc.paths_total - issued;
    const uint32_t room = rc.capacity - surv;
    const uint32_t n_new = left < (uint64_t)room ? (uint32_t)left : room;
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    // Path order.  pixel-major (default): path = pixel * spp + sample, so the 32 lanes of a warp start as 32 samples of ONE pixel —
    // near-identical camera rays (coherent traversal and material branches) whose finished records stay adjacent in the `done`
    // queue, where k_accumulate folds them into one atomic per run.  sample-major (RT_RENDER_SAMPLE_MAJOR): path = sample * pixels
    // + pixel.
    // Work split: every WARP takes one contiguous slice of the new paths and walks it 32 at a time (coalesced 512-byte stores per
    // queue word); a lane's path index advances by 32 per step, so in pixel-major order its pixel changes only every spp / 32 steps
    // and the tile decode (two integer divisions) is redone only then.  One 64-bit division per thread.
    const uint32_t mod = rc.pixel_major ? rc.spp : (uint32_t)rc.pixels_local;
    const uint32_t lane = threadIdx.x & 31u, n_warps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t per_warp = ((n_new + n_warps - 1u) / n_warps + 31u) & ~31u;
    const uint32_t begin = min((tid >> 5) * per_warp, n_new), end = min(begin + per_warp, n_new);
    uint32_t major = 0u, minor = 0u;
    if (begin + lane < end) {
        const uint64_t path0 = issued + begin + lane;
        major = (uint32_t)(path0 / mod);
        minor = (uint32_t)(path0 - (uint64_t)major * mod);
    }
    uint32_t cur_q = 0xFFFFFFFFu, x = 0u, y = 0u, gp = 0u;
    for (uint32_t i = begin + lane; i < end; i += 32u) {
        const uint32_t sample = rc.pixel_major ? minor : major, qi = rc.pixel_major ? major : minor;
        if (qi != cur_q) {
            uint32_t yl;
            pixel_from_index(qi, rc.width, rc.rows_local, x, yl);
            y = rc.row_begin + yl * rc.row_stride;
            gp = y * rc.width + x;
            cur_q = qi;
        }
        const Philox4 rnd = philox4x32_10(0u, RT_DIM_CAMERA, rc.seed_lo, rc.seed_hi, gp, sample);
        float3 o, d;
        if (rc.user_rays) {  // batched render::color(&ray, ..): the caller's ray as it is (direction not normalised, src/render.rs:5)
            const RtRayIn r = rc.user_rays[gp];
            o = f3(r.origin[0], r.origin[1], r.origin[2]);
            d = f3(r.direction[0], r.direction[1], r.direction[2]);
        } else {
            camera_ray(rc.cam, x, y, rc.width, rc.height, rnd, o, d);
        }
        const uint32_t slot = surv + i;
        RT_ASSERT(slot < rc.capacity && sample < rc.spp && gp < rc.width * rc.height);
        q.a[slot] = make_float4(o.x, o.y, o.z, __uint_as_float(RT_NO_PRIM));
        q.b[slot] = make_float4(d.x, d.y, d.z, __uint_as_float(gp));
        q.c[slot] = make_float4(1.0f, 1.0f, 1.0f, __uint_as_float((sample << 8) | 1u));
        minor += 32u;
        while (minor >= mod) { minor -= mod; ++major; }
    }
    if (tid == 0) {
        const uint32_t total = surv + n_new;
        wc->total[s] = total;
        wc->issued[(k + 1) % RT_RING] = issued + n_new;
        wc->surv[(k + 2) % RT_RING] = 0u;
        wc->done[(k + 1) % RT_RING] = 0u;
        wc->cursor[(k + 1) % RT_RING] = 0u;
        wc->st_issued = issued + n_new;
        wc->st_rays += total;
        wc->st_total = total;
        wc->st_iter = k;
    }
}

// ---- extend ----------------------------------------------------------------------------------------

// Stage the BVH prefix (and, for scenes that fit, the primitive heads and compact rectangle records) into shared memory:
// [stack RT_SMEM_STACK x blockDim words][nodes 2*n_staged_nodes][heads n_staged_prims][rects 4*n_rects] — three bulk copies on the
// copy engine, one mbarrier; the caller waits (mbar_wait(mbar, 0)) after it has issued its first ray loads.
__device__ __forceinline__ Staged stage_scene(const DScene& S, float4* smem_base, uint32_t mbar) {
    const uint32_t n_node4 = 2u * S.n_staged_nodes, n_head4 = S.n_staged_prims, n_rect4 = S.n_staged_prims ? 4u * S.n_rects : 0u;
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(smem_base);
    const uint32_t scene = base + RT_SMEM_STACK * blockDim.x * 4u;
    bulk_stage(mbar, BulkPart{S.nodes, scene, 16u * n_node4}, BulkPart{S.prim_head, scene + 16u * n_node4, 16u * n_head4},
               BulkPart{S.rects, scene + 16u * (n_node4 + n_head4), 16u * n_rect4});
    return Staged{scene, scene + 16u * n_node4, scene + 16u * (n_node4 + n_head4), 0u, base + 4u * threadIdx.x};
}

// Staging of the all-in-shared-memory variant: [stack (depth + 2) x blockDim words][pairs, 7 float4 each (intersect.cuh: octant-ready
// records)][heads n_leaf_prims][rects 4*n_leaf_rects][boxes 2*n_boxes] — the blob is laid out once by rt_scene_create (S.oct_blob), so this
// is ONE bulk copy.
__device__ __forceinline__ Staged stage_scene_oct(const DScene& S, float4* smem_base, uint32_t mbar) {
    const uint32_t stack_words = (S.bvh_depth + 2u) * blockDim.x;
    const uint32_t n_pairs = (S.n_nodes + 1u) / 2u, n_node4 = 7u * n_pairs, n_head4 = S.n_leaf_prims, n_rect4 = 4u * S.n_leaf_rects;
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(smem_base);
    const uint32_t scene = base + stack_words * 4u;
    bulk_stage(mbar, BulkPart{S.oct_blob, scene, 16u * S.oct_blob4});
    return Staged{scene, scene + 16u * n_node4, scene + 16u * (n_node4 + n_head4), scene + 16u * (n_node4 + n_head4 + n_rect4), base + 4u * threadIdx.x};
}
size_t extend_all_smem_bytes(const DScene& S, uint32_t block) {
    return (size_t)(S.bvh_depth + 2u) * block * 4u + (size_t)RT_OCT_PAIR_BYTES * ((S.n_nodes + 1u) / 2u) + 16u * (size_t)S.n_leaf_prims + 64u * (size_t)S.n_leaf_rects +
           32u * (size_t)S.n_boxes;
}

// ALL: the whole tree and every primitive head are in shared memory (book-one: 836 nodes, 487 heads) — the per-access
// "staged or global" range checks compile away.
template <bool COUNT, bool ALL, bool GEN>
__global__ void __launch_bounds__(ALL ? RT_EXTEND_ALL_BIG_BLOCK : RT_EXTEND_BLOCK, ALL ? 1 : RT_EXTEND_MIN_BLOCKS) k_extend(const DScene S, const RayQueue q, float2* __restrict__ hits, WaveCounters* __restrict__ wc,
                                                          uint32_t k, uint32_t seed_lo, uint32_t seed_hi) {
    extern __shared__ float4 smem4[];
    __shared__ __align__(8) uint64_t s_mbar;
    const uint32_t mbar = (uint32_t)__cvta_generic_to_shared(&s_mbar);
    const Staged sm = ALL ? stage_scene_oct(S, smem4, mbar) : stage_scene(S, smem4, mbar);
    // the tables stream in on the copy engine while every thread requests its first ray from HBM
    const uint32_t n = queue_size(wc, k);
    tail_roll(wc, k, n);
    const uint32_t stride = gridDim.x * blockDim.x;
    TraversalWork work{0u, 0u};
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
    uint32_t sb = 0u;
    if (i < n) { a = q.a[i]; b = q.b[i]; if (S.has_media) sb = __float_as_uint(q.c[i].w); }
    mbar_wait(mbar, 0u);
    while (i < n) {
        MediumKey key{0u, 0u, 0u, seed_lo, seed_hi};
        if (S.has_media) {
            key.k0 = __float_as_uint(b.w);
            key.k1 = sb >> 8;
            key.segment = sb & 0xFFu;
        }
        const ClosestHit h = closest_hit<COUNT, ALL, GEN>(S, sm, f3(a.x, a.y, a.z), f3(b.x, b.y, b.z), __float_as_uint(a.w), key, &work);
        hits[i] = make_float2(h.t, __uint_as_float(h.prim));
        i += stride;
        if (i < n) { a = q.a[i]; b = q.b[i]; if (S.has_media) sb = __float_as_uint(q.c[i].w); }
    }
    if (COUNT) {  // instrumented build only (RT_RENDER_COUNT_WORK): one atomic pair per warp
        uint32_t bt = work.box_tests, pt = work.prim_tests;
        for (int off = 16; off > 0; off >>= 1) { bt += __shfl_down_sync(0xFFFFFFFFu, bt, off); pt += __shfl_down_sync(0xFFFFFFFFu, pt, off); }
        if (lane_id() == 0) { atomicAdd(&wc->st_box_tests, (unsigned long long)bt); atomicAdd(&wc->st_prim_tests, (unsigned long long)pt); }
    }
}

// ---- extend, persistent form for BVH scenes ------------------------------------------------------------------
//
// k_extend gives every thread one ray at a time, so a warp is as slow as its longest ray: the r1 profiles show 20.8
// (book-one) and 9.2 (1 M spheres) active threads per instruction.  Here the warps are persistent and their lanes are
// refilled: as soon as RT_REFILL_MIN lanes have finished their rays, those lanes store their hits and take the next
// rays, while the other lanes keep their traversal state (Aila & Laine's persistent threads with dynamic fetch, done
// per lane with __ballot_sync / popc ranks).  Rays are not read from HBM by the idle lanes — that would stall the
// whole warp for a DRAM round trip at every refill — but from a per-warp shared-memory double buffer that cp.async
// fills one 32-ray chunk ahead; chunks come from a per-launch cursor in ranges of RT_FEED_RANGE (one global atomic per
// 128 rays), so a warp's first 32 rays are consecutive queue entries (a coherent 8x4 camera tile).
__device__ __forceinline__ void cp_async16(uint32_t smem, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t smem, const void* gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

constexpr uint32_t RT_FEED_BUF_BYTES = 32u * (16u + 16u + 4u);  // a, b, c.w of 32 rays
constexpr uint32_t RT_FEED_WARP_BYTES = 2u * RT_FEED_BUF_BYTES;

struct WarpFeed {  // warp-uniform
    uint32_t buf;                  // shared-window address of this warp's two buffers
    uint32_t cur, cur_base, cur_cnt, cur_pos;
    uint32_t nxt_base, nxt_cnt;    // chunk in flight into buffer cur ^ 1 (nxt_cnt = 0: none)
    uint32_t r_next, r_end;        // claimed ray indices not yet requested
    bool more;                     // the cursor may still hold rays
};

// claim the next chunk of up to 32 rays; cnt = 0 when the queue is exhausted
__device__ __forceinline__ void feed_claim(WarpFeed& F, uint32_t* cursor, uint32_t n, uint32_t& base, uint32_t& cnt) {
    if (F.r_next >= F.r_end) {
        if (F.more) {
            uint32_t b = 0u;
            if (lane_id() == 0) b = atomicAdd(cursor, (uint32_t)RT_FEED_RANGE);
            b = __shfl_sync(0xFFFFFFFFu, b, 0);
            if (b >= n) { F.more = false; }
            else { F.r_next = b; F.r_end = min(b + (uint32_t)RT_FEED_RANGE, n); }
        }
        if (!F.more) { base = 0u; cnt = 0u; return; }
    }
    base = F.r_next;
    cnt = min(32u, F.r_end - F.r_next);
    F.r_next += cnt;
}

__device__ __forceinline__ void feed_prefetch(const WarpFeed& F, uint32_t which, const RayQueue& q, bool has_media, uint32_t base, uint32_t cnt) {
    const uint32_t lane = lane_id();
    if (lane < cnt) {
        const uint32_t b = F.buf + which * RT_FEED_BUF_BYTES;
        cp_async16(b + 16u * lane, &q.a[base + lane]);
        cp_async16(b + 512u + 16u * lane, &q.b[base + lane]);
        if (has_media) cp_async4(b + 1024u + 4u * lane, &q.c[base + lane].w);
    }
    cp_async_commit();
}

template <bool COUNT, bool GEN>
__global__ void __launch_bounds__(RT_EXTEND_BLOCK, RT_BVH_MIN_BLOCKS) k_extend_bvh(const DScene S, const RayQueue q, float2* __restrict__ hits,
                                                                                  WaveCounters* __restrict__ wc, uint32_t k, uint32_t seed_lo, uint32_t seed_hi,
                                                                                  uint32_t staged_bytes) {
    extern __shared__ float4 smem4[];
    __shared__ __align__(8) uint64_t s_mbar;
    const uint32_t mbar = (uint32_t)__cvta_generic_to_shared(&s_mbar);
    const Staged sm = stage_scene(S, smem4, mbar);
    const uint32_t n = queue_size(wc, k);
    tail_roll(wc, k, n);
    uint32_t* cursor = &wc->cursor[(k & ~RT_K_TAIL) % RT_RING];
    const uint32_t lane = lane_id(), lt = (1u << lane) - 1u;
    const bool has_media = S.has_media != 0u;
    TraversalWork work{0u, 0u};

    WarpFeed F;
    F.buf = (uint32_t)__cvta_generic_to_shared(smem4) + RT_SMEM_STACK * blockDim.x * 4u + staged_bytes + (threadIdx.x >> 5) * RT_FEED_WARP_BYTES;
    F.cur = 0u; F.cur_pos = 0u; F.r_next = F.r_end = 0u; F.more = true;
    feed_claim(F, cursor, n, F.cur_base, F.cur_cnt);
    if (F.cur_cnt) feed_prefetch(F, 0u, q, has_media, F.cur_base, F.cur_cnt);
    feed_claim(F, cursor, n, F.nxt_base, F.nxt_cnt);
    if (F.nxt_cnt) feed_prefetch(F, 1u, q, has_media, F.nxt_base, F.nxt_cnt);
    cp_async_wait_all();  // (the first refill needs buffer 0; waiting for both keeps the code simple)
    __syncwarp();
    mbar_wait(mbar, 0u);  // the staged tree prefix (bulk copy issued by stage_scene, overlapped with the first two ray chunks)

    // per-lane traversal state
    RayCtx R = make_ray(f3(0.f, 0.f, 0.f), f3(1.f, 0.f, 0.f));
    ClosestHit best{CUDART_INF_F, RT_NO_PRIM};
    MediumKey key{0u, 0u, 0u, seed_lo, seed_hi};
    uint32_t stack_spill[RT_STACK_DEPTH - RT_SMEM_STACK];
    uint32_t ray_i = RT_NO_PRIM, self_prim = RT_NO_PRIM, pair = RT_NO_PRIM;
    uint32_t leaf_begin = 0u, leaf_end = 0u;  // pending primitive range of this lane (while-while: tested together with the other lanes')
    int sp = 0;
    const uint32_t root_pair = __float_as_uint(load_node(S, sm, 0).w);

    for (;;) {
        const bool idle = pair == RT_NO_PRIM && leaf_end == leaf_begin;
        const uint32_t m_idle = __ballot_sync(0xFFFFFFFFu, idle);
        if (m_idle == 0xFFFFFFFFu || (uint32_t)__popc(m_idle) >= (uint32_t)RT_REFILL_MIN) {
            if (idle && ray_i != RT_NO_PRIM) {
                hits[ray_i] = make_float2(best.t, __uint_as_float(best.prim));
                ray_i = RT_NO_PRIM;
            }
            uint32_t want = m_idle;  // lanes still without a ray
            while (want != 0u) {
                if (F.cur_pos == F.cur_cnt) {  // current buffer drained: switch to the chunk in flight, request the one after
                    if (F.nxt_cnt == 0u) break;  // queue exhausted
                    cp_async_wait_all();
                    __syncwarp();
                    F.cur ^= 1u; F.cur_base = F.nxt_base; F.cur_cnt = F.nxt_cnt; F.cur_pos = 0u;
                    feed_claim(F, cursor, n, F.nxt_base, F.nxt_cnt);
                    if (F.nxt_cnt) feed_prefetch(F, F.cur ^ 1u, q, has_media, F.nxt_base, F.nxt_cnt);
                }
                const uint32_t avail = F.cur_cnt - F.cur_pos;
                const uint32_t rank = (uint32_t)__popc(want & lt);
                const bool mine = ((want >> lane) & 1u) && rank < avail;
                if (mine) {
                    const uint32_t slot = F.cur_pos + rank;
                    RT_ASSERT(slot < 32u && F.cur_base + slot < n);
                    const uint32_t b = F.buf + F.cur * RT_FEED_BUF_BYTES;
                    const float4 ra = lds128(b + 16u * slot), rb = lds128(b + 512u + 16u * slot);
                    ray_i = F.cur_base + slot;
                    self_prim = __float_as_uint(ra.w);
                    if (has_media) {
                        const uint32_t sb = lds32(b + 1024u + 4u * slot);
                        key.k0 = __float_as_uint(rb.w);
                        key.k1 = sb >> 8;
                        key.segment = sb & 0xFFu;
                    }
                    R = make_ray(f3(ra.x, ra.y, ra.z), f3(rb.x, rb.y, rb.z));
                    best.t = CUDART_INF_F;
                    best.prim = RT_NO_PRIM;
                    for (uint32_t i = 0; i < S.n_always; ++i) test_prim<true, false, false, GEN>(S, sm, i, R, self_prim, key, best);
                    if (COUNT) work.prim_tests += S.n_always;
                    sp = 0;
                    pair = root_pair;
                    leaf_begin = leaf_end = 0u;
                }
                const uint32_t got = __ballot_sync(0xFFFFFFFFu, mine);
                want &= ~got;
                F.cur_pos += (uint32_t)__popc(got);
                __syncwarp();  // every lane has read its slot before the buffer can be refilled
            }
            if (__ballot_sync(0xFFFFFFFFu, pair != RT_NO_PRIM || leaf_end != leaf_begin) == 0u) break;  // no lane has work and the queue is exhausted
        }
        // pending leaves: intersected once RT_LEAF_MIN lanes hold one, or when no lane can take a node step
        const bool pend = leaf_end != leaf_begin;
        const uint32_t m_pend = __ballot_sync(0xFFFFFFFFu, pend);
        const uint32_t m_trav = __ballot_sync(0xFFFFFFFFu, pair != RT_NO_PRIM && !pend);
        if (m_pend != 0u && ((uint32_t)__popc(m_pend) >= (uint32_t)RT_LEAF_MIN || m_trav == 0u)) {
            if (pend) {
                leaf_tests<COUNT, false, GEN>(S, sm, R, self_prim, key, best, leaf_begin, leaf_end, &work);
                leaf_begin = leaf_end = 0u;
            }
        }
        if (pair != RT_NO_PRIM && leaf_end == leaf_begin) pair = node_step<COUNT>(S, sm, R, best, pair, sp, stack_spill, leaf_begin, leaf_end, &work);
    }
    if (COUNT) {
        uint32_t bt = work.box_tests, pt = work.prim_tests;
        for (int off = 16; off > 0; off >>= 1) { bt += __shfl_down_sync(0xFFFFFFFFu, bt, off); pt += __shfl_down_sync(0xFFFFFFFFu, pt, off); }
        if (lane == 0) { atomicAdd(&wc->st_box_tests, (unsigned long long)bt); atomicAdd(&wc->st_prim_tests, (unsigned long long)pt); }
    }
}

// Extend for small all-rectangle scenes (cornell-box): the frame-rect list of device_scene.h.  Per ray: one rotation
// into each frame (none for the world frame) and three reciprocals, then ~15 instructions per rectangle
// (t = (w - o_k) / d_k, two in-plane coordinates against their extents) instead of the ~37 of the general record
// (r1b profile: 111 SASS instructions per 3 rectangles, 68 % issue-active -> issue bound).  No traversal stack and
// few registers: twice the resident warps of k_extend.  Rectangle::hit semantics as in rect_hit (geometry.rs:152-181:
// t >= 1e-6, inclusive extents); closest hit with the reference's tie rule (strict <, DFS-earlier wins: the groups
// are not in DFS order, so an exact tie compares primitive indices, which ARE DFS ranks in a small scene).
template <int K>
__device__ __forceinline__ void frame_axis_rects(const DScene& S, uint32_t recs, uint32_t exts, uint32_t f, const float3 o, const float3 d, const float3 inv,
                                                 uint32_t self_prim, float& best_t, uint32_t& best_i) {
    const float ok = K == 0 ? o.x : (K == 1 ? o.y : o.z), ik = K == 0 ? inv.x : (K == 1 ? inv.y : inv.z);
    const float oa = K == 0 ? o.y : o.x, da = K == 0 ? d.y : d.x;
    const float ob = K == 2 ? o.y : o.z, db = K == 2 ? d.y : d.z;
    const uint32_t j0 = S.fr_range[3u * f + K], j1 = S.fr_range[3u * f + K + 1];
    // groups hold one to three rectangles: a rolled loop (3 uniform instructions per trip) beats ptxas' unroll-by-4
    // with its remainder ladder, which cost ~25 instructions per group (r1d profile: 689 lane-instructions per ray)
#pragma unroll 1
    for (uint32_t j = j0; j < j1; ++j) {
        const float4 A = lds128(recs + 16u * j);
        float2 B;
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(B.x), "=f"(B.y) : "r"(exts + 8u * j));
        const float t = (A.x - ok) * ik;
        const float pa = fmaf(t, da, oa) - A.y, pb = fmaf(t, db, ob) - A.z;
        const uint32_t idx = __float_as_uint(A.w);
        // best_t starts at FLT_MAX, not +inf, so t = +inf (d_k = 0) can neither be closer nor tie; NaN fails t >= eps
        const bool closer = (t < best_t) | ((t == best_t) & (idx < best_i));
        const bool hit = (t >= RT_T_EPS) & (fabsf(pa) <= B.x) & (fabsf(pb) <= B.y) & (idx != self_prim) & closer;
        best_t = hit ? t : best_t;
        best_i = hit ? idx : best_i;
    }
}

__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// One frame box (device_scene.h): slab test against [lo, hi] in the frame's coordinates, then the entry crossing — or, when the
// entry face carries no rectangle, the point lies outside its rectangle, the ray starts on it or inside the box — the exit
// crossing.  t is (plane - o_k) * (1 / d_k) exactly as for a loose rectangle, the plane being the face rectangle's own.
__device__ __forceinline__ void frame_box(uint32_t box, const float3 o, const float3 d, const float3 inv, uint32_t self_prim, float& best_t, uint32_t& best_i) {
    const float4 L = lds128(box), H = lds128(box + 16u);
    const float x0 = (L.x - o.x) * inv.x, x1 = (H.x - o.x) * inv.x;
    const float y0 = (L.y - o.y) * inv.y, y1 = (H.y - o.y) * inv.y;
    const float z0 = (L.z - o.z) * inv.z, z1 = (H.z - o.z) * inv.z;
    // fminf / fmaxf drop a NaN operand (0 * inf: origin on a slab plane of a zero direction component): that axis then does not constrain
    const float nx = fminf(x0, x1), ny = fminf(y0, y1), nz = fminf(z0, z1);
    const float fx = fmaxf(x0, x1), fy = fmaxf(y0, y1), fz = fmaxf(z0, z1);
    const float t_en = fmaxf(fmaxf(nx, ny), nz), t_ex = fminf(fminf(fx, fy), fz);
    const bool crosses = t_en <= t_ex;  // false (also for NaN): the line misses the box, it crosses none of its faces inside their bounds
    if (__ballot_sync(__activemask(), crosses) == 0u) return;  // warp-uniform skip; everything below is branch-free (32 of 32 lanes per instruction)
    // face slot = 2 * axis + (hi plane ? 1 : 0); the entry plane of axis k is lo when d_k > 0
    const uint32_t ke = (nx == t_en) ? 0u : ((ny == t_en) ? 1u : 2u), kx = (fx == t_ex) ? 0u : ((fy == t_ex) ? 1u : 2u);
    const float de = ke == 0u ? d.x : (ke == 1u ? d.y : d.z), dx = kx == 0u ? d.x : (kx == 1u ? d.y : d.z);
    const uint32_t slot_en = 2u * ke + (de > 0.0f ? 0u : 1u), slot_ex = 2u * kx + (dx > 0.0f ? 1u : 0u);
    // both crossings are evaluated; the entry crossing wins when it is a hit (it is the nearer one)
    const uint32_t id_en = lds32(box + 128u + 4u * slot_en), id_ex = lds32(box + 128u + 4u * slot_ex);
    bool hit_en = crosses & (t_en >= RT_T_EPS) & (id_en != RT_NO_PRIM) & (id_en != self_prim);
    bool hit_ex = crosses & (t_ex >= RT_T_EPS) & (id_ex != RT_NO_PRIM) & (id_ex != self_prim);
    if (L.w == 0.0f) {  // (warp-uniform) some face rectangle is smaller than its face: test the crossing points against THEIR rectangles
        const float4 Fe = lds128(box + 32u + 16u * slot_en), Fx = lds128(box + 32u + 16u * slot_ex);  // centre a, centre b, half a, half b
        const float ex = fmaf(t_en, d.x, o.x), ey = fmaf(t_en, d.y, o.y), ez = fmaf(t_en, d.z, o.z);
        const float xx = fmaf(t_ex, d.x, o.x), xy = fmaf(t_ex, d.y, o.y), xz = fmaf(t_ex, d.z, o.z);
        const float ea = (ke == 0u ? ey : ex) - Fe.x, eb = (ke == 2u ? ey : ez) - Fe.y;
        const float xa = (kx == 0u ? xy : xx) - Fx.x, xb = (kx == 2u ? xy : xz) - Fx.y;
        hit_en &= (fabsf(ea) <= Fe.z) & (fabsf(eb) <= Fe.w);
        hit_ex &= (fabsf(xa) <= Fx.z) & (fabsf(xb) <= Fx.w);
    }
    const float t = hit_en ? t_en : t_ex;
    const uint32_t idx = hit_en ? id_en : id_ex;
    const bool closer = (t < best_t) | ((t == best_t) & (idx < best_i));
    const bool hit = (hit_en | hit_ex) & closer;
    best_t = hit ? t : best_t;
    best_i = hit ? idx : best_i;
}

// Closest hit of one ray against the frame-rect list staged at shared-window address `base` (k_extend_rects and, for scenes that use
// the list, k_trace — so that the fixed-ray parity tests exercise exactly what renders run).
__device__ __forceinline__ ClosestHit frame_rects_closest(const DScene& S, uint32_t base, const float3 o, const float3 d, uint32_t self_prim) {
    const uint32_t boxes = base + 48u * S.fr_frames, recs = boxes + 160u * S.fr_boxes, exts = recs + 16u * S.fr_loose;
    float best_t = 3.402823466e38f;  // a miss is prim == RT_NO_PRIM; its t is never read
    uint32_t best_i = RT_NO_PRIM;
#pragma unroll 1
    for (uint32_t f = 0; f < S.fr_frames; ++f) {
        float3 of = o, df = d;
        if (f > 0u || !S.fr_identity) {
            const float4 u0 = lds128(base + 48u * f), u1 = lds128(base + 48u * f + 16u), u2 = lds128(base + 48u * f + 32u);
            of = f3(dot3(u0, o), dot3(u1, o), dot3(u2, o));
            df = f3(dot3(u0, d), dot3(u1, d), dot3(u2, d));
        }
        const float3 inv = f3(rcp_approx(df.x), rcp_approx(df.y), rcp_approx(df.z));
#pragma unroll 1
        for (uint32_t bx = S.fr_box_range[f]; bx < S.fr_box_range[f + 1]; ++bx) frame_box(boxes + 160u * bx, of, df, inv, self_prim, best_t, best_i);
        if (S.fr_range[3u * f] != S.fr_range[3u * f + 3u]) {  // frames made of boxes only (a rotated cube) have no loose rectangles
            frame_axis_rects<0>(S, recs, exts, f, of, df, inv, self_prim, best_t, best_i);
            frame_axis_rects<1>(S, recs, exts, f, of, df, inv, self_prim, best_t, best_i);
            frame_axis_rects<2>(S, recs, exts, f, of, df, inv, self_prim, best_t, best_i);
        }
    }
    return ClosestHit{best_i == RT_NO_PRIM ? CUDART_INF_F : best_t, best_i};
}

__global__ void __launch_bounds__(RT_EXTEND_BLOCK, RT_RECTS_MIN_BLOCKS) k_extend_rects(const DScene S, const RayQueue q, float2* __restrict__ hits,
                                                                                      WaveCounters* __restrict__ wc, uint32_t k) {
    extern __shared__ float4 smem4[];
    __shared__ __align__(8) uint64_t s_mbar;
    const uint32_t mbar = (uint32_t)__cvta_generic_to_shared(&s_mbar);
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(smem4);
    bulk_stage(mbar, BulkPart{S.fr_blob, base, 16u * S.fr_blob4});  // the frame-rect list streams in while the first rays are requested
    const uint32_t n = queue_size(wc, k);
    tail_roll(wc, k, n);
    const uint32_t stride = gridDim.x * blockDim.x;
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
    if (i < n) { a = q.a[i]; b = q.b[i]; }
    mbar_wait(mbar, 0u);
    while (i < n) {
        const ClosestHit h = frame_rects_closest(S, base, f3(a.x, a.y, a.z), f3(b.x, b.y, b.z), __float_as_uint(a.w));
        hits[i] = make_float2(h.t, __uint_as_float(h.prim));
        i += stride;
        if (i < n) { a = q.a[i]; b = q.b[i]; }
    }
}

// ---- shade -----------------------------------------------------------------------------------------

// Shared memory of k_shade (dynamic): [materials 2 float4 each][prim heads][rect normals][prim meta uint2].  The
// primitive tables are staged when the scene is small (S.n_staged_prims): the r1b profile showed the dependent chain
// hit -> prim_meta -> prim_head -> rect_nrm (three L1/L2 round trips) as ~30 % of the kernel's stall samples.
struct ShadeOne {
    bool alive, record;
    float4 o0, o1, o2;  // alive: the three queue words of the scattered ray; record: o0 = (radiance, pixel)
};

template <bool GEN>
__device__ __forceinline__ ShadeOne shade_one(const DScene& S, const RenderConst& rc, const float4* smats, uint32_t n_staged_mats,
                                              const float4 a, const float4 b, const float4 c, const float2 h) {
    ShadeOne r;
    r.alive = r.record = false;
    r.o0 = r.o1 = r.o2 = make_float4(0.f, 0.f, 0.f, 0.f);
    const uint32_t prim = __float_as_uint(h.y);
    if (prim == RT_NO_PRIM) return r;  // miss -> black (render.rs:26)
    RT_ASSERT(prim < S.n_prims);
    const uint2 meta = S.prim_meta[prim];
    const float4 head = S.prim_head[prim];
    const int32_t mi = (int32_t)meta.y;
    if (mi < 0) return r;  // no material -> black (render.rs:18-20)
    RT_ASSERT((uint32_t)mi < S.n_mats);
    DMaterial m;
    {
        const float4* src = ((uint32_t)mi < n_staged_mats) ? &smats[2 * mi] : reinterpret_cast<const float4*>(S.mats) + 2 * (size_t)mi;
        const float4 m0 = src[0], m1 = src[1];
        m.kind = __float_as_uint(m0.x); m.texture = (int32_t)__float_as_uint(m0.y); m.fuzz = m0.z; m.ior = m0.w;
        m.r = m1.x; m.g = m1.y; m.b = m1.z; m.pad = 0.0f;
    }
    const uint32_t sb = __float_as_uint(c.w), gp = __float_as_uint(b.w);
    const uint32_t sample = sb >> 8, segment = sb & 0xFFu;
    const float3 o = f3(a.x, a.y, a.z), d = f3(b.x, b.y, b.z), thr = f3(c.x, c.y, c.z);
    const MediumKey key{gp, sample, segment, rc.seed_lo, rc.seed_hi};
    const HitDetail hd = hit_detail<GEN>(S, prim, o, d, h.x, m.texture >= 0, key, head);
    const Philox4 rnd = philox4x32_10(segment, RT_DIM_SHADE, rc.seed_lo, rc.seed_hi, gp, sample);
    const ScatterResult sr = shade_material(S, m, d, hd, rnd);
    if (sr.scattered) {
        if (segment < rc.max_depth) {  // color(.., 0) returns black without tracing (render.rs:6-8)
            const float3 t2 = thr * sr.attenuation;
            const bool is_medium = head.w <= 0.0f && __float_as_uint(head.y) >= (uint32_t)RT_PRIM_MEDIUM_SPHERE;  // general records carry their kind in .y
            r.o0 = make_float4(hd.p.x, hd.p.y, hd.p.z, __uint_as_float(is_medium ? RT_NO_PRIM : prim));
            r.o1 = make_float4(sr.dir.x, sr.dir.y, sr.dir.z, b.w);
            r.o2 = make_float4(t2.x, t2.y, t2.z, __uint_as_float((sample << 8) | (segment + 1u)));
            r.alive = true;
        }
    } else if (m.kind == RT_MAT_DIFFUSE_LIGHT) {
        const float3 L = thr * sr.attenuation;
        if (L.x != 0.0f || L.y != 0.0f || L.z != 0.0f) {
            r.o0 = make_float4(L.x, L.y, L.z, b.w);
            r.record = true;
        }
    }
    return r;
}

// Each thread shades RT_SHADE_RAYS rays per block round (their queue words are all requested before the first is
// used: RT_SHADE_RAYS x the memory-level parallelism) and the block allocates output slots for the whole round with
// ONE barrier pair and ONE pair of global atomics (block_append2 generalised to RT_SHADE_RAYS ballots per warp).
// k_shade<2> (two rays per thread per round) at 6 resident blocks of 128 threads (<= 85 registers) instead of 8 (<= 64): at 64 its
// two in-flight rays spill in the hot path — 38.1 -> 44.3 Grays/s on cornell-box (profiles/r2_20_shade2_variants.txt)
#ifndef RT_SHADE2_MIN_BLOCKS
#define RT_SHADE2_MIN_BLOCKS 6
#endif
// FR (small all-rectangle scenes, S.fr_n > 0): the closest hit is computed HERE, from the frame-rect list (frame_rects_closest, the body of
// k_extend_rects), instead of being read back from the hit queue — extend + shade as one kernel.  The ray is read once instead of
// twice and the 8-byte hit record is neither written nor read: 96 + 8 instead of 144 + 8 bytes of queue traffic per segment
// (cornell-box is bound by exactly that traffic), and a wavefront iteration is three dependent launches instead of four.
template <int R, bool GEN, bool FR = false>
#ifndef RT_SHADE_FR_MIN_BLOCKS
#define RT_SHADE_FR_MIN_BLOCKS 8
#endif
__global__ void __launch_bounds__(RT_SHADE_BLOCK, R == 2 ? RT_SHADE2_MIN_BLOCKS : (FR ? RT_SHADE_FR_MIN_BLOCKS : RT_SHADE_MIN_BLOCKS)) k_shade(const DScene S0, const RenderConst rc, const RayQueue qin, RayQueue qout,
                                               const float2* __restrict__ hits, float4* __restrict__ done, WaveCounters* __restrict__ wc, uint32_t k) {
    extern __shared__ float4 shade_smem[];
    __shared__ uint32_t s_cnt[RT_SHADE_BLOCK / 32 + 1][2];
    const uint32_t n_staged = min(S0.n_mats, RT_MAX_STAGED_MATS);
    const uint32_t n_sp = S0.n_staged_prims, n_sr = n_sp ? S0.n_rects : 0u;
    float4* smats = shade_smem;
    float4* s_head = smats + 2u * n_staged;
    float4* s_nrm = s_head + n_sp;
    uint2* s_meta = reinterpret_cast<uint2*>(s_nrm + n_sr);
#ifndef RT_SHADE_BULK
#define RT_SHADE_BULK 1
#endif

#if RT_SHADE_BULK
    // one bulk copy of the blob rt_scene_create laid out in exactly this order; it lands while the first round's rays are requested
    __shared__ __align__(8) uint64_t s_mbar;
    const uint32_t mbar = (uint32_t)__cvta_generic_to_shared(&s_mbar);
    const uint32_t fr_base = (uint32_t)__cvta_generic_to_shared(shade_smem) + 16u * S0.shade_blob4;  // FR: the frame-rect list behind the shade tables
    if (FR) bulk_stage(mbar, BulkPart{S0.shade_blob, (uint32_t)__cvta_generic_to_shared(shade_smem), 16u * S0.shade_blob4}, BulkPart{S0.fr_blob, fr_base, 16u * S0.fr_blob4});
    else bulk_stage(mbar, BulkPart{S0.shade_blob, (uint32_t)__cvta_generic_to_shared(shade_smem), 16u * S0.shade_blob4});
    bool tables_ready = false;
#else
    for (uint32_t i = threadIdx.x; i < S0.shade_blob4; i += blockDim.x) shade_smem[i] = S0.shade_blob[i];
    __syncthreads();
    const uint32_t mbar = 0u, fr_base = 0u;
    static_assert(!FR, "the fused form stages its tables with the bulk copy");
    bool tables_ready = true;
#endif
    DScene S = S0;
    if (n_sp) { S.prim_head = s_head; S.prim_meta = s_meta; S.rect_nrm = s_nrm; }
    const uint32_t s = (k & ~RT_K_TAIL) % RT_RING;
    const uint32_t n = queue_size(wc, k);
    if (FR) tail_roll(wc, k, n);  // no extend launch in front of this kernel: it rolls the ring counters of a tail iteration itself
    uint32_t* surv_next = &wc->surv[((k & ~RT_K_TAIL) + 1) % RT_RING];
    uint32_t* done_cnt = &wc->done[s];
    const uint32_t lane = lane_id(), warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    const uint32_t stride = gridDim.x * blockDim.x * R;
    // whole blocks iterate together: the slot allocation below synchronises the block
    for (uint32_t base = blockIdx.x * blockDim.x * R; base < n; base += stride) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const uint32_t nx = base + stride + r * blockDim.x + threadIdx.x;
            if (nx < n) {
                if ((lane & 7u) == 0u) { prefetch_l2(&qin.a[nx]); prefetch_l2(&qin.b[nx]); prefetch_l2(&qin.c[nx]); }
                if (!FR && (lane & 15u) == 0u) prefetch_l2(&hits[nx]);
            }
        }
        float4 a[R], b[R], c[R];
        float2 h[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const uint32_t i = base + r * blockDim.x + threadIdx.x;
            h[r] = make_float2(0.f, __uint_as_float(RT_NO_PRIM));
            if (i < n) { if (!FR) h[r] = hits[i]; a[r] = qin.a[i]; b[r] = qin.b[i]; c[r] = qin.c[i]; }
            else { a[r] = b[r] = c[r] = make_float4(0.f, 0.f, 0.f, 0.f); }
        }
        if (!tables_ready) { mbar_wait(mbar, 0u); tables_ready = true; }  // block-uniform
        if (FR) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if (base + r * blockDim.x + threadIdx.x < n) {
                    const ClosestHit ch = frame_rects_closest(S0, fr_base, f3(a[r].x, a[r].y, a[r].z), f3(b[r].x, b[r].y, b[r].z), __float_as_uint(a[r].w));
                    h[r] = make_float2(ch.t, __uint_as_float(ch.prim));
                }
            }
        }
        ShadeOne res[R];
        uint32_t ma[R], mb[R], ca = 0u, cb = 0u;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            res[r] = shade_one<GEN>(S, rc, smats, n_staged, a[r], b[r], c[r], h[r]);
            ma[r] = __ballot_sync(0xFFFFFFFFu, res[r].alive);
            mb[r] = __ballot_sync(0xFFFFFFFFu, res[r].record);
            ca += (uint32_t)__popc(ma[r]);
            cb += (uint32_t)__popc(mb[r]);
        }
        if (lane == 0) { s_cnt[warp][0] = ca; s_cnt[warp][1] = cb; }
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t ta = 0u, tb = 0u;
            for (uint32_t w = 0; w < n_warps; ++w) {
                const uint32_t wa = s_cnt[w][0], wb = s_cnt[w][1];
                s_cnt[w][0] = ta; s_cnt[w][1] = tb;
                ta += wa; tb += wb;
            }
            s_cnt[n_warps][0] = ta ? atomicAdd(surv_next, ta) : 0u;
            s_cnt[n_warps][1] = tb ? atomicAdd(done_cnt, tb) : 0u;
        }
        __syncthreads();
        uint32_t slot = s_cnt[n_warps][0] + s_cnt[warp][0], dslot = s_cnt[n_warps][1] + s_cnt[warp][1];
        __syncwarp();
        const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (res[r].alive) {
                const uint32_t sl = slot + (uint32_t)__popc(ma[r] & lt);
                RT_ASSERT(sl < rc.capacity);
                qout.a[sl] = res[r].o0; qout.b[sl] = res[r].o1; qout.c[sl] = res[r].o2;
            }
            RT_ASSERT(!res[r].record || dslot + (uint32_t)__popc(mb[r] & lt) < rc.capacity);
            if (res[r].record) done[dslot + (uint32_t)__popc(mb[r] & lt)] = res[r].o0;
            slot += (uint32_t)__popc(ma[r]);
            dslot += (uint32_t)__popc(mb[r]);
        }
    }
}

// ---- accumulate ------------------------------------------------------------------------------------

// SUMSQ: the caller asked for per-pixel sums of squares too (standard errors); renders that do not (bench, rt_render with a null
// out_sumsq) run the instantiation with half the shuffles.
template <bool SUMSQ>
__global__ void __launch_bounds__(256) k_accumulate(const RenderConst rc, const float4* __restrict__ done, const WaveCounters* __restrict__ wc,
                                                    uint32_t k, double* __restrict__ sum, double* __restrict__ sumsq) {
    // f64 accumulators in the framebuffer (native 64-bit atomics): an f32 RUNNING sum of 1024+ equal sky samples rounds the same way
    // on every add and drifts by 1e-4 relative, which the 3-sigma image test sees on zero-variance pixels.
    const uint32_t n = wc->done[k % RT_RING];
    const uint32_t stride = gridDim.x * blockDim.x, lane = threadIdx.x & 31u;
    constexpr int NV = SUMSQ ? 6 : 3;
    // Runs of records of one pixel (pixel-major path order keeps the samples of a pixel together) are summed in the warp first — a
    // segmented shuffle reduction over the run heads — and only the head of a run touches the framebuffer: 3-6 f64 atomics per run
    // instead of per path (the L2 atomic units were the limit of this kernel: 164 G atomics/s on book-one 4K).  The in-warp tree is f32
    // (<= 32 addends, 5 levels: <= 3e-7 relative, far below the f64 accumulation it feeds): a third of the shuffles and adds of an
    // f64 tree, which had become the cost of the kernel (~150 instructions per record).
    for (uint32_t base = blockIdx.x * blockDim.x; base < n; base += stride) {  // warp-uniform trip count
        const uint32_t i = base + threadIdx.x;
        const bool valid = i < n;
        const float4 r = valid ? done[i] : make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(0xffffffffu));
        const uint32_t gp = __float_as_uint(r.w);
        float v[NV];
        v[0] = r.x; v[1] = r.y; v[2] = r.z;
        if (SUMSQ) { v[NV - 3] = r.x * r.x; v[NV - 2] = r.y * r.y; v[NV - 1] = r.z * r.z; }
        const uint32_t gp_prev = __shfl_up_sync(0xffffffffu, gp, 1);
        const bool head = lane == 0u || gp != gp_prev;
        const uint32_t heads = __ballot_sync(0xffffffffu, head);
        if (heads != 0xffffffffu) {
            const uint32_t after = lane == 31u ? 0u : heads >> (lane + 1u);  // bit j: lane + 1 + j starts a new run
#pragma unroll
            for (uint32_t delta = 1u; delta < 32u; delta <<= 1) {
                const bool same_run = lane + delta < 32u && (after & ((1u << delta) - 1u)) == 0u;
#pragma unroll
                for (int c = 0; c < NV; ++c) {
                    const float other = __shfl_down_sync(0xffffffffu, v[c], delta);
                    if (same_run) v[c] += other;
                }
            }
        }
        if (!head || !valid) continue;
        const uint32_t y = gp / rc.width, x = gp - y * rc.width;
        const uint32_t yl = (y - rc.row_begin) / rc.row_stride;
        const size_t o = ((size_t)yl * rc.width + x) * 3;
        RT_ASSERT(y >= rc.row_begin && (y - rc.row_begin) % rc.row_stride == 0u && yl < rc.rows_local && x < rc.width);
        atomicAdd(&sum[o + 0], (double)v[0]);
        atomicAdd(&sum[o + 1], (double)v[1]);
        atomicAdd(&sum[o + 2], (double)v[2]);
        if (SUMSQ) {
            atomicAdd(&sumsq[o + 0], (double)v[NV - 3]);
            atomicAdd(&sumsq[o + 1], (double)v[NV - 2]);
            atomicAdd(&sumsq[o + 2], (double)v[NV - 1]);
        }
    }
}

// last step of the accumulate stage: f64 accumulators -> the caller's f32 sum buffers
__global__ void __launch_bounds__(256) k_finish(const double* __restrict__ acc, size_t n, float* __restrict__ out) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) out[i] = (float)acc[i];
}

// ---- resolve (examples/book-one.rs:90-100) -----------------------------------------------------------

// f64 like the reference (mean = sum / spp, sqrt, * 255, min, truncation): bytes are integer work and must come out bit-exact; IEEE
// f64 division and square root are correctly rounded on the device, so this equals the host computation for every input.
__global__ void k_resolve_u8(const float* __restrict__ sum, uint32_t width, uint32_t height, double spp, uint8_t* __restrict__ out) {
    const size_t n = (size_t)width * height * 3;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t c = (uint32_t)(i % 3);
        const size_t pix = i / 3;
        const uint32_t row = (uint32_t)(pix / width), x = (uint32_t)(pix - (size_t)row * width);
        const uint32_t y = height - 1u - row;  // rows top -> bottom
        double v = sqrt((double)sum[((size_t)y * width + x) * 3 + c] / spp) * 255.0;
        v = (v != v) ? 255.0 : fmin(v, 255.0);  // f64::min keeps the non-NaN operand (A.12)
        out[i] = (uint8_t)fmax(v, 0.0);         // `as usize` saturates at 0
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

// `self_order` (nullable): per ray the DFS rank of the primitive the ray STARTS on (RT_NO_PRIM = none) — the state a scattered ray
// carries through the wavefront (queue word a.w), so that the f32 self-intersection rule of intersect.cuh is testable ray by ray.
__global__ void __launch_bounds__(RT_EXTEND_BLOCK) k_trace(const DScene S, const RtRayIn* __restrict__ rays, const uint32_t* __restrict__ self_order,
                                                         uint64_t n, uint32_t seed_lo, uint32_t seed_hi, RtHitOut* __restrict__ out) {
    extern __shared__ float4 smem4[];
    __shared__ __align__(8) uint64_t s_mbar;
    const uint32_t mbar = (uint32_t)__cvta_generic_to_shared(&s_mbar);
    const Staged sm = stage_scene(S, smem4, mbar);
    mbar_wait(mbar, 0u);
    // scenes that render through the frame-rect list are traced through it too; its blob sits behind the staged scene prefix
    float4* fr_smem = smem4 + (RT_SMEM_STACK * blockDim.x) / 4u + 2u * S.n_staged_nodes + S.n_staged_prims + (S.n_staged_prims ? 4u * S.n_rects : 0u);
    if (S.fr_n) {
        for (uint32_t i = threadIdx.x; i < S.fr_blob4; i += blockDim.x) fr_smem[i] = S.fr_blob[i];
        __syncthreads();
    }
    const uint32_t fr_base = (uint32_t)__cvta_generic_to_shared(fr_smem);
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const RtRayIn r = rays[i];
        const float3 o = f3(r.origin[0], r.origin[1], r.origin[2]), d = f3(r.direction[0], r.direction[1], r.direction[2]);
        const MediumKey key{(uint32_t)i, (uint32_t)(i >> 32), 1u, seed_lo, seed_hi};
        uint32_t self_prim = RT_NO_PRIM;
        if (self_order) {
            const uint32_t ord = self_order[i];
            if (ord < S.n_orders) self_prim = __ldg(&S.order_to_prim[ord]);
        }
        const ClosestHit h = S.fr_n ? frame_rects_closest(S, fr_base, o, d, self_prim) : closest_hit<false>(S, sm, o, d, self_prim, key);
        RtHitOut res;
        res.hit = 0; res.prim = -1; res.t = 0.0f; res.material = -1;
        res.p[0] = res.p[1] = res.p[2] = res.n[0] = res.n[1] = res.n[2] = res.uv[0] = res.uv[1] = 0.0f;
        if (h.prim != RT_NO_PRIM) {
            const uint2 meta = __ldg(&S.prim_meta[h.prim]);
            const HitDetail hd = hit_detail(S, h.prim, o, d, h.t, true, key, __ldg(&S.prim_head[h.prim]));
            res.hit = 1; res.prim = (int32_t)meta.x; res.t = hd.t; res.material = (int32_t)meta.y;
            res.p[0] = hd.p.x; res.p[1] = hd.p.y; res.p[2] = hd.p.z;
            res.n[0] = hd.n.x; res.n[1] = hd.n.y; res.n[2] = hd.n.z;
            res.uv[0] = hd.u; res.uv[1] = hd.v;
        }
        out[i] = res;
    }
}

// ---- launch configuration ---------------------------------------------------------------------------------
//
// Nothing here is per process.  What depends on the DEVICE (SM count, the dynamic shared-memory ceiling each kernel has been
// granted there — cudaFuncSetAttribute is per device) lives in DeviceState, keyed by device id; what depends on the SCENE
// (shared-memory bytes of its staged tables, resident blocks per SM at those sizes, which extend kernel it gets) lives in the
// SceneLaunch the scene owns; tuning options live in the context (LaunchOptions, read once by rt_init).  Several contexts on
// several devices, and several scenes on one device, can therefore render concurrently from different host threads.

namespace {

constexpr int RT_MAX_DEVICES = 64;
struct DeviceState {
    std::mutex mu;
    int sms = 0;
    size_t granted[16] = {0};  // largest MaxDynamicSharedMemorySize set so far, per kernel slot
};
DeviceState g_dev[RT_MAX_DEVICES];
// (kernels that exist in a GEN = false and a GEN = true instantiation take two consecutive slots: slot + gen)
enum KernelSlot { SLOT_EXTEND = 0, SLOT_EXTEND_ALL = 2, SLOT_BVH = 4, SLOT_SHADE1 = 6, SLOT_SHADE2 = 8, SLOT_EXTEND_COUNT = 10, SLOT_BVH_COUNT, SLOT_RECTS, SLOT_TRACE, SLOT_SHADE_FR1, SLOT_SHADE_FR2 };

// the ceiling only ever grows: a launch may use any size up to it, so scenes of different sizes never invalidate each other
template <class K>
cudaError_t grant(DeviceState& d, int slot, K kernel, size_t bytes) {
    if (bytes <= d.granted[slot] || bytes <= 48u * 1024u) return cudaSuccess;
    const cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e == cudaSuccess) d.granted[slot] = bytes;
    return e;
}
template <class K>
cudaError_t occupancy(int& blocks, K kernel, int threads, size_t bytes) {
    const cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks, kernel, threads, bytes);
    if (e == cudaSuccess && blocks < 1) blocks = 1;
    return e;
}

}  // namespace

cudaError_t kernels_num_sms(int device, int* sms) {
    if (device < 0 || device >= RT_MAX_DEVICES) return cudaErrorInvalidDevice;
    DeviceState& d = g_dev[device];
    std::lock_guard<std::mutex> lock(d.mu);
    if (d.sms == 0) {
        const cudaError_t e = cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, device);
        if (e != cudaSuccess) { d.sms = 0; return e; }
    }
    *sms = d.sms;
    return cudaSuccess;
}

bool extend_all_smem_fits(const DScene& S) {
    return extend_all_smem_bytes(S, RT_EXTEND_ALL_BLOCK) <= 100u * 1024u || extend_all_smem_bytes(S, RT_EXTEND_ALL_BIG_BLOCK) <= 224u * 1024u;
}
constexpr size_t RT_EXTEND_ALL_MAX_SMEM = 100u * 1024u;
constexpr size_t RT_EXTEND_ALL_BIG_MAX_SMEM = 224u * 1024u;  // of the 227 KB a block may own (the kernel adds 16 bytes of static shared memory)

// The calling thread must have `device` current.
cudaError_t kernels_configure_scene(int device, const DScene& S, size_t scene_bytes, size_t shade_bytes, size_t fr_bytes, SceneLaunch& L) {
    L = SceneLaunch();
    cudaError_t e = kernels_num_sms(device, &L.num_sms);
    if (e != cudaSuccess) return e;
    DeviceState& d = g_dev[device];
    std::lock_guard<std::mutex> lock(d.mu);
    L.device = device;
    L.scene_smem = scene_bytes;
    L.extend_smem = scene_bytes + (size_t)RT_SMEM_STACK * RT_EXTEND_BLOCK * sizeof(uint32_t);
    L.bvh_smem = L.extend_smem + (size_t)(RT_EXTEND_BLOCK / 32) * RT_FEED_WARP_BYTES;
    L.shade_smem = shade_bytes;
    L.rects_smem = fr_bytes;
#define RT_TRY(x) do { e = (x); if (e != cudaSuccess) return e; } while (0)
    const bool gen = S.has_general != 0u;  // primitives reached through their general 128-byte record: the kernels that carry that path
    L.gen = gen;
    auto* const extend_k = gen ? k_extend<false, false, true> : k_extend<false, false, false>;
    auto* const all_k = gen ? k_extend<false, true, true> : k_extend<false, true, false>;
    auto* const bvh_k = gen ? k_extend_bvh<false, true> : k_extend_bvh<false, false>;
    auto* const shade1_k = gen ? k_shade<1, true> : k_shade<1, false>;
    auto* const shade2_k = gen ? k_shade<2, true> : k_shade<2, false>;
    RT_TRY(grant(d, SLOT_EXTEND + gen, extend_k, L.extend_smem));
    RT_TRY(grant(d, SLOT_EXTEND_COUNT, k_extend<true, false, true>, L.extend_smem));
    L.trace_smem = L.extend_smem + fr_bytes;
    RT_TRY(grant(d, SLOT_TRACE, k_trace, L.trace_smem));
    RT_TRY(grant(d, SLOT_BVH + gen, bvh_k, L.bvh_smem));
    RT_TRY(grant(d, SLOT_BVH_COUNT, k_extend_bvh<true, true>, L.bvh_smem));
    RT_TRY(occupancy(L.extend_blocks, extend_k, RT_EXTEND_BLOCK, L.extend_smem));
    RT_TRY(occupancy(L.trace_blocks, k_trace, RT_EXTEND_BLOCK, L.trace_smem));
    RT_TRY(occupancy(L.bvh_blocks, bvh_k, RT_EXTEND_BLOCK, L.bvh_smem));
    if (fr_bytes) {
        RT_TRY(grant(d, SLOT_RECTS, k_extend_rects, fr_bytes));
        RT_TRY(occupancy(L.rects_blocks, k_extend_rects, RT_EXTEND_BLOCK, fr_bytes));
    }
    RT_TRY(grant(d, SLOT_SHADE1 + gen, shade1_k, shade_bytes));
    RT_TRY(grant(d, SLOT_SHADE2 + gen, shade2_k, shade_bytes));
    if (fr_bytes && !gen) {  // extend fused into shade (k_shade<., ., true>): the shade tables plus the frame-rect list
        L.shade_fr_smem = shade_bytes + fr_bytes;
        RT_TRY(grant(d, SLOT_SHADE_FR1, k_shade<1, false, true>, L.shade_fr_smem));
        RT_TRY(grant(d, SLOT_SHADE_FR2, k_shade<2, false, true>, L.shade_fr_smem));
        RT_TRY(occupancy(L.shade_fr_blocks, k_shade<1, false, true>, RT_SHADE_BLOCK, L.shade_fr_smem));
        RT_TRY(occupancy(L.shade2_fr_blocks, k_shade<2, false, true>, RT_SHADE_BLOCK, L.shade_fr_smem));
        L.fused_rects = true;
    }
    RT_TRY(occupancy(L.shade_blocks, shade1_k, RT_SHADE_BLOCK, shade_bytes));
    RT_TRY(occupancy(L.shade2_blocks, shade2_k, RT_SHADE_BLOCK, shade_bytes));
    // whole tree + heads in shared memory in the octant-ready layout, as long as that keeps the occupancy of the general kernel
    L.all_staged = S.oct_blob != nullptr && S.n_nodes > 0u && !S.n_small && !S.root_is_leaf;
    if (L.all_staged) {
        L.all_block = RT_EXTEND_ALL_BLOCK;
        L.all_smem = extend_all_smem_bytes(S, RT_EXTEND_ALL_BLOCK);
        if (L.all_smem > RT_EXTEND_ALL_MAX_SMEM) {  // more than half an SM: one block of twice the threads
            L.all_block = RT_EXTEND_ALL_BIG_BLOCK;
            L.all_smem = extend_all_smem_bytes(S, RT_EXTEND_ALL_BIG_BLOCK);
        }
        if (L.all_smem > RT_EXTEND_ALL_BIG_MAX_SMEM) L.all_staged = false;
        else {
            RT_TRY(grant(d, SLOT_EXTEND_ALL + gen, all_k, L.all_smem));
            RT_TRY(occupancy(L.all_blocks, all_k, L.all_block, L.all_smem));
            if (L.all_blocks * L.all_block < L.extend_blocks * RT_EXTEND_BLOCK) L.all_staged = false;
        }
    }
#undef RT_TRY
    return cudaSuccess;
}

// Tail of a render: once every path has been issued the queue only shrinks, and the host knows an upper bound of what is left from
// its last status poll (Launch::work_hint).  A full persistent grid (hundreds of blocks, each staging its tables) for a few thousand
// rays costs ~2 us more per launch than the blocks the rays need; the ~50 iterations of a depth-50 tail are all of that kind.
static inline int fit_grid(const Launch& L, int full_grid, int threads) {
    if (L.work_hint == RT_WORK_UNKNOWN) return full_grid;
    const uint64_t need = ((uint64_t)L.work_hint + (uint32_t)threads - 1u) / (uint32_t)threads;
    return (int)(need < 1u ? 1u : (need < (uint64_t)full_grid ? need : (uint64_t)full_grid));
}

void launch_generate(const Launch& L, const RenderConst& rc, const RayQueue& q, WaveCounters* wc, uint32_t k) {
    // nothing left to issue: thread 0 rolls the counters
    k_generate<<<L.work_hint == RT_WORK_UNKNOWN ? L.sl->num_sms * L.opt->gen_blocks : 1, 256, 0, L.st>>>(rc, q, wc, k);
}
bool extend_is_fused(const Launch& L, const DScene& S, bool count_work) { return S.fr_n && !count_work && L.sl->fused_rects && L.opt->fuse_rects; }
void launch_extend(const Launch& L, const DScene& S, const RayQueue& q, float2* hits, WaveCounters* wc, uint32_t k, uint32_t seed_lo, uint32_t seed_hi,
                   bool count_work) {
    const SceneLaunch& sl = *L.sl;
    const int sms = sl.num_sms;
    if (extend_is_fused(L, S, count_work)) return;  // k_shade<., ., true> intersects
    if (S.fr_n && !count_work) {
        k_extend_rects<<<fit_grid(L, sms * sl.rects_blocks, RT_EXTEND_BLOCK), RT_EXTEND_BLOCK, sl.rects_smem, L.st>>>(S, q, hits, wc, k);
        return;
    }
    // The persistent form pays where rays are long and uneven (1 M spheres: 69 box tests per ray, +20 %); on trees that
    // live in shared memory (book-one: 18 box tests) the per-ray set-up at partial warp occupancy costs more than the
    // idle lanes it saves (r1e: 7.1 vs 12.6 Grays/s), so those keep one ray per thread.
    const bool persistent = L.opt->persistent == 1 || (L.opt->persistent == 2 && S.n_nodes >= RT_PERSISTENT_MIN_NODES);
    if (persistent && S.n_small == 0u && !S.root_is_leaf && S.n_nodes > 0u) {
        const int grid = sms * sl.bvh_blocks;
        if (count_work) k_extend_bvh<true, true><<<grid, RT_EXTEND_BLOCK, sl.bvh_smem, L.st>>>(S, q, hits, wc, k, seed_lo, seed_hi, (uint32_t)sl.scene_smem);
        else if (sl.gen) k_extend_bvh<false, true><<<grid, RT_EXTEND_BLOCK, sl.bvh_smem, L.st>>>(S, q, hits, wc, k, seed_lo, seed_hi, (uint32_t)sl.scene_smem);
        else k_extend_bvh<false, false><<<grid, RT_EXTEND_BLOCK, sl.bvh_smem, L.st>>>(S, q, hits, wc, k, seed_lo, seed_hi, (uint32_t)sl.scene_smem);
        return;
    }
    const int grid = sms * sl.extend_blocks;
    if (count_work) k_extend<true, false, true><<<grid, RT_EXTEND_BLOCK, sl.extend_smem, L.st>>>(S, q, hits, wc, k, seed_lo, seed_hi);
    else if (sl.all_staged) {
        const int g = fit_grid(L, sms * sl.all_blocks, sl.all_block);
        if (sl.gen) k_extend<false, true, true><<<g, sl.all_block, sl.all_smem, L.st>>>(S, q, hits, wc, k, seed_lo, seed_hi);
        else k_extend<false, true, false><<<g, sl.all_block, sl.all_smem, L.st>>>(S, q, hits, wc, k, seed_lo, seed_hi);
    } else {
        const int g = fit_grid(L, grid, RT_EXTEND_BLOCK);
        if (sl.gen) k_extend<false, false, true><<<g, RT_EXTEND_BLOCK, sl.extend_smem, L.st>>>(S, q, hits, wc, k, seed_lo, seed_hi);
        else k_extend<false, false, false><<<g, RT_EXTEND_BLOCK, sl.extend_smem, L.st>>>(S, q, hits, wc, k, seed_lo, seed_hi);
    }
}
void launch_shade(const Launch& L, const DScene& S, const RenderConst& rc, const RayQueue& qin, const RayQueue& qout, const float2* hits, float4* done,
                  WaveCounters* wc, uint32_t k, bool count_work) {
    const SceneLaunch& sl = *L.sl;
    if (extend_is_fused(L, S, count_work)) {
        // one ray per thread per round at 8 resident blocks (51 registers): 4776 Mpaths/s on cornell-box against 4422 for two rays per
        // round at 6 blocks (74 registers) — with the intersection in the kernel, resident warps count for more than loads in flight
        // (profiles/r2_42_fused_extend_shade.txt); 10 / 12 blocks (48 / 40 registers) are slower again
        if (L.opt->fuse_rects == 2) k_shade<2, false, true><<<fit_grid(L, sl.num_sms * sl.shade2_fr_blocks, 2 * RT_SHADE_BLOCK), RT_SHADE_BLOCK, sl.shade_fr_smem, L.st>>>(S, rc, qin, qout, hits, done, wc, k);
        else k_shade<1, false, true><<<fit_grid(L, sl.num_sms * sl.shade_fr_blocks, RT_SHADE_BLOCK), RT_SHADE_BLOCK, sl.shade_fr_smem, L.st>>>(S, rc, qin, qout, hits, done, wc, k);
        return;
    }
    // two rays per thread per barrier round where shading is light on registers (small all-rectangle scenes: no f64 sphere re-solve);
    // +5 % there, -7 % on the sphere scenes (r1f), hence per scene
    if (S.fr_n && L.opt->shade_rays2) {
        const int g = fit_grid(L, sl.num_sms * sl.shade2_blocks, 2 * RT_SHADE_BLOCK);
        if (sl.gen) k_shade<2, true><<<g, RT_SHADE_BLOCK, sl.shade_smem, L.st>>>(S, rc, qin, qout, hits, done, wc, k);
        else k_shade<2, false><<<g, RT_SHADE_BLOCK, sl.shade_smem, L.st>>>(S, rc, qin, qout, hits, done, wc, k);
    } else {
        const int g = fit_grid(L, sl.num_sms * sl.shade_blocks, RT_SHADE_BLOCK);
        if (sl.gen) k_shade<1, true><<<g, RT_SHADE_BLOCK, sl.shade_smem, L.st>>>(S, rc, qin, qout, hits, done, wc, k);
        else k_shade<1, false><<<g, RT_SHADE_BLOCK, sl.shade_smem, L.st>>>(S, rc, qin, qout, hits, done, wc, k);
    }
}
void launch_accumulate(const Launch& L, const RenderConst& rc, const float4* done, const WaveCounters* wc, uint32_t k, double* sum, double* sumsq) {
    const int g = fit_grid(L, L.sl->num_sms * L.opt->acc_blocks, 256);
    if (sumsq) k_accumulate<true><<<g, 256, 0, L.st>>>(rc, done, wc, k, sum, sumsq);
    else k_accumulate<false><<<g, 256, 0, L.st>>>(rc, done, wc, k, sum, sumsq);
}
void launch_finish(int sms, const double* acc, size_t n, float* out, cudaStream_t st) { k_finish<<<sms * 4, 256, 0, st>>>(acc, n, out); }
void launch_resolve_u8(int sms, const float* sum, uint32_t width, uint32_t height, uint32_t spp, uint8_t* out, cudaStream_t st) {
    k_resolve_u8<<<sms * 4, 256, 0, st>>>(sum, width, height, (double)spp, out);
}
void launch_deinterleave(int sms, const float* gathered, float* full, uint32_t width, uint32_t height, uint32_t n_shards, cudaStream_t st) {
    const uint32_t rows_padded = (height + n_shards - 1) / n_shards;
    k_deinterleave<<<sms * 4, 256, 0, st>>>(gathered, full, width, height, n_shards, rows_padded);
}
void launch_trace(const SceneLaunch& sl, const DScene& S, const RtRayIn* rays, const uint32_t* self_order, uint64_t n, uint32_t seed_lo, uint32_t seed_hi,
                  RtHitOut* out, cudaStream_t st) {
    k_trace<<<sl.num_sms * sl.trace_blocks, RT_EXTEND_BLOCK, sl.trace_smem, st>>>(S, rays, self_order, n, seed_lo, seed_hi, out);
}

}  // namespace rtb
