// C ABI of the backend (include/rt_b200.h): context / scene lifetime, the wavefront render loop,
// trace and resolve.  No CPU fallback: every entry point needs a CUDA device.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>  // types and prototypes only: the library is loaded with dlopen (rt_init_multi)

#include <algorithm>
#include <cmath>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "box_prims.h"
#include "bvh_build.h"
#include "device_scene.h"
#include "kernels.h"
#include "lbvh.h"

using namespace rtb;

namespace {

thread_local std::string g_error;

int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_error = buf;
    return code;
}

#define RT_CUDA(call)                                                                                   \
    do {                                                                                                \
        cudaError_t e_ = (call);                                                                        \
        if (e_ != cudaSuccess) return fail(RT_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

template <class T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    cudaError_t ensure(size_t count) {
        if (count <= n && p) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        n = 0;
        cudaError_t e = cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T));
        if (e == cudaSuccess) n = count;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
};

// Persistent host threads.  A render with L lanes (or G devices) used to spawn and join L (G) std::threads per call: 60-90 us each
// on the VM-hosted boxes this runs on — 0.2 ms on a 10 ms render, 1-2 ms for the 8 x 3 threads of an rt_render_multi.  The workers
// are created on first use and parked on a condition variable between renders.
class WorkerPool {
public:
    WorkerPool() = default;
    WorkerPool(const WorkerPool&) = delete;
    WorkerPool& operator=(const WorkerPool&) = delete;
    ~WorkerPool() {
        {
            std::lock_guard<std::mutex> lk(mu_);
            stop_ = true;
        }
        cv_work_.notify_all();
        for (auto& t : threads_) t.join();
    }
    // job(i) for i in [0, n), each on its own worker; returns when all have finished.  One caller at a time.
    void run(uint32_t n, const std::function<void(uint32_t)>& job) {
        std::unique_lock<std::mutex> lk(mu_);
        while (threads_.size() < n) {
            const uint32_t id = (uint32_t)threads_.size();
            threads_.emplace_back([this, id] { loop(id); });  // (blocks on mu_ until this call waits below)
        }
        job_ = &job;
        n_ = n;
        pending_ = n;
        ++generation_;
        cv_work_.notify_all();
        cv_done_.wait(lk, [&] { return pending_ == 0; });
        job_ = nullptr;
    }

private:
    void loop(uint32_t id) {
        uint64_t seen = 0;
        std::unique_lock<std::mutex> lk(mu_);
        for (;;) {
            cv_work_.wait(lk, [&] { return stop_ || generation_ != seen; });
            if (stop_) return;
            seen = generation_;
            if (id >= n_) continue;  // this round needs fewer workers
            const std::function<void(uint32_t)>* job = job_;
            lk.unlock();
            (*job)(id);
            lk.lock();
            if (--pending_ == 0) cv_done_.notify_one();
        }
    }
    std::mutex mu_;
    std::condition_variable cv_work_, cv_done_;
    std::vector<std::thread> threads_;
    const std::function<void(uint32_t)>* job_ = nullptr;
    uint32_t n_ = 0, pending_ = 0;
    uint64_t generation_ = 0;
    bool stop_ = false;
};

}  // namespace

struct RtContext {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t side = nullptr;  // accumulate_k runs here, overlapped with generate/extend of iteration k+1
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_shaded[2] = {nullptr, nullptr}, ev_accumulated[2] = {nullptr, nullptr};
    std::vector<cudaEvent_t> stage_events;  // 6 per wavefront iteration when stage timing is on (no sync until the render ends)
    bool overlap_accumulate = true;
    // wavefront state (sized on demand for the requested queue capacity)
    DevBuf<float4> qa[2], qb[2], qc[2], done[2];
    DevBuf<float2> hits;
    DevBuf<WaveCounters> counters;
    DevBuf<float> fb_sum, fb_sumsq;
    DevBuf<double> acc;  // f64 per-pixel accumulators (sum, then sum of squares)
    DevBuf<uint8_t> scratch;
    WaveCounters* h_status = nullptr;  // pinned
    bool stage_timing = false;
    LaunchOptions lopt;              // kernel-level tuning, fixed at rt_init_ex
    RtSceneOptions scene_defaults{}; // what rt_scene_create (no explicit options) uses
    int num_sms = 0;
    WorkerPool pool;                // host threads that drive the lanes
    std::vector<RtContext*> lanes;  // sub-contexts (own streams, queues, accumulators) of the concurrent wavefronts of one render
    cudaEvent_t ev_lane_begin = nullptr;
    DevBuf<float> lane_sum, lane_sumsq;  // the lanes' compact row shards before they are interleaved into the caller's buffer
    uint32_t n_lanes = 2;
    bool capacity_auto = true;  // RtContextOptions.queue_capacity was 0
    bool lanes_auto = true;   // RtContextOptions.lanes was 0: render_lanes picks by render size
    uint32_t default_capacity = 1u << 22;  // 4 M live rays (570 MB of wavefront state): per-launch fixed costs amortise, +20 % over 1 M (r1c)
};

struct RtScene {
    RtContext* ctx = nullptr;
    DScene d{};
    DevBuf<float4> nodes, prim_head, texels, rects, rect_nrm, boxes, fr_blob, oct_blob, shade_blob;
    DevBuf<uint2> prim_meta;
    DevBuf<uint32_t> order_to_prim;
    DevBuf<RtPrim> gen;
    DevBuf<DMaterial> mats;
    DevBuf<RtTexture> texs;
    uint32_t bvh_depth = 0;
    size_t staged_bytes = 0;
    size_t fr_bytes = 0;     // k_extend_rects' staged frame-rect table (0: the scene has none)
    size_t shade_bytes = 0;  // k_shade's staged tables: materials, and for small scenes heads / meta / rect normals
    SceneLaunch launch;      // shared-memory sizes / occupancy of this scene's kernels on ctx->device
};

namespace {

// World-space box of a flattened primitive: the 8 corners of the local box pushed through the forward
// matrix (what Bound for Sprite does, src/optimize.rs:128-185), padded for f32 traversal.
Box3 primBox(const RtSceneDesc* desc, const RtPrim& p) {
    if (p.kind == RT_PRIM_MEDIUM) {  // Bound for ConstantMedium = its boundary's box (optimize.rs:508-516)
        Box3 b{{INFINITY, INFINITY, INFINITY}, {-INFINITY, -INFINITY, -INFINITY}};
        const uint32_t first = p.aux, count = (uint32_t)p.param[2];
        for (uint32_t i = first; i < first + count; ++i) {
            const Box3 c = primBox(desc, desc->boundary_prims[i]);
            for (int a = 0; a < 3; ++a) { b.lo[a] = std::min(b.lo[a], c.lo[a]); b.hi[a] = std::max(b.hi[a], c.hi[a]); }
        }
        return b;
    }
    double lx, ly, lz;
    if (p.kind == RT_PRIM_RECT) { lx = 0.5 * std::fabs((double)p.param[0]); ly = 0.5 * std::fabs((double)p.param[1]); lz = 1e-6; }  // optimize.rs:115-126
    else { lx = ly = lz = std::fabs((double)p.param[0]); }                                                                            // optimize.rs:104-113
    double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    const float* m = p.world_from_local;
    for (int c = 0; c < 8; ++c) {
        const double x = (c & 1) ? lx : -lx, y = (c & 2) ? ly : -ly, z = (c & 4) ? lz : -lz;
        for (int a = 0; a < 3; ++a) {
            const double v = (double)m[4 * a] * x + (double)m[4 * a + 1] * y + (double)m[4 * a + 2] * z + (double)m[4 * a + 3];
            lo[a] = std::min(lo[a], v);
            hi[a] = std::max(hi[a], v);
        }
    }
    Box3 b;
    for (int a = 0; a < 3; ++a) {
        const double pad = 4e-6 * std::max(1.0, std::max(std::fabs(lo[a]), std::fabs(hi[a])));
        b.lo[a] = std::nextafter((float)(lo[a] - pad), -INFINITY);
        b.hi[a] = std::nextafter((float)(hi[a] + pad), INFINITY);
    }
    return b;
}

bool translationOnly(const RtPrim& p) {
    const float* m = p.local_from_world;
    const float* w = p.world_from_local;
    return m[0] == 1.f && m[1] == 0.f && m[2] == 0.f && m[4] == 0.f && m[5] == 1.f && m[6] == 0.f && m[8] == 0.f && m[9] == 0.f && m[10] == 1.f &&
           w[0] == 1.f && w[1] == 0.f && w[2] == 0.f && w[4] == 0.f && w[5] == 1.f && w[6] == 0.f && w[8] == 0.f && w[9] == 0.f && w[10] == 1.f;
}

int validateDesc(const RtSceneDesc* d) {
    if (!d) return fail(RT_ERR_INVALID, "scene descriptor is null");
    if (d->abi_version != RT_ABI_VERSION) return fail(RT_ERR_INVALID, "ABI version %u, library is %u", d->abi_version, RT_ABI_VERSION);
    if (d->n_prims == 0) return fail(RT_ERR_INVALID, "scene has no primitives");
    if (!d->prims) return fail(RT_ERR_INVALID, "prims is null");
    if (d->n_materials && !d->materials) return fail(RT_ERR_INVALID, "materials is null");
    if (d->n_textures && !d->textures) return fail(RT_ERR_INVALID, "textures is null");
    if (d->n_texels && !d->texels) return fail(RT_ERR_INVALID, "texels is null");
    if (d->n_boundary_prims && !d->boundary_prims) return fail(RT_ERR_INVALID, "boundary_prims is null");
    for (uint32_t i = 0; i < d->n_boundary_prims; ++i) {
        const RtPrim& p = d->boundary_prims[i];
        if (p.kind != RT_PRIM_SPHERE && p.kind != RT_PRIM_RECT) return fail(RT_ERR_UNSUPPORTED, "boundary prim %u: kind %u (only spheres and rectangles bound a medium)", i, p.kind);
        for (int k = 0; k < 12; ++k)
            if (!std::isfinite(p.local_from_world[k]) || !std::isfinite(p.world_from_local[k])) return fail(RT_ERR_INVALID, "boundary prim %u: non-finite transform", i);
    }
    for (uint32_t i = 0; i < d->n_prims; ++i) {
        const RtPrim& p = d->prims[i];
        if (p.kind > RT_PRIM_MEDIUM) return fail(RT_ERR_UNSUPPORTED, "prim %u: unknown kind %u", i, p.kind);
        if (p.material >= (int32_t)d->n_materials) return fail(RT_ERR_INVALID, "prim %u: material %d out of range", i, p.material);
        if ((p.kind == RT_PRIM_MEDIUM_SPHERE || p.kind == RT_PRIM_MEDIUM) && !(p.param[1] > 0.0f)) return fail(RT_ERR_INVALID, "prim %u: medium density must be > 0", i);
        if (p.kind == RT_PRIM_MEDIUM) {
            const float cnt = p.param[2];
            if (!(cnt >= 1.0f) || cnt >= 16777216.0f || cnt != std::floor(cnt) || (uint64_t)p.aux + (uint64_t)cnt > d->n_boundary_prims)
                return fail(RT_ERR_INVALID, "prim %u: medium boundary range [%u, +%g) out of bounds", i, p.aux, (double)cnt);
        }
        for (int k = 0; k < 12; ++k)
            if (!std::isfinite(p.local_from_world[k]) || !std::isfinite(p.world_from_local[k])) return fail(RT_ERR_INVALID, "prim %u: non-finite transform", i);
    }
    for (uint32_t i = 0; i < d->n_materials; ++i) {
        const RtMaterial& m = d->materials[i];
        if (m.kind > RT_MAT_ISOTROPIC) return fail(RT_ERR_UNSUPPORTED, "material %u: unknown kind %u", i, m.kind);
        if (m.kind != RT_MAT_DIELECTRIC && (m.texture < 0 || m.texture >= (int32_t)d->n_textures))
            return fail(RT_ERR_INVALID, "material %u: texture %d out of range", i, m.texture);
    }
    for (uint32_t i = 0; i < d->n_textures; ++i) {
        const RtTexture& t = d->textures[i];
        if (t.kind > RT_TEX_IMAGE) return fail(RT_ERR_UNSUPPORTED, "texture %u: unknown kind %u", i, t.kind);
        if (t.kind == RT_TEX_CHECKER && (t.a < 0 || t.b < 0 || t.a >= (int32_t)d->n_textures || t.b >= (int32_t)d->n_textures))
            return fail(RT_ERR_INVALID, "texture %u: checker child out of range", i);
        if (t.kind == RT_TEX_IMAGE && (t.width == 0 || t.height == 0 || t.texel_offset + (uint64_t)t.width * t.height > d->n_texels))
            return fail(RT_ERR_INVALID, "texture %u: texel range out of bounds", i);
    }
    return RT_OK;
}

// Frame-rect table of a small all-rectangle scene (device_scene.h "fr_*").  Rigid transforms only: every row of the
// object<-world matrix must be +-(a row of the frame) to 1e-6, else the scene keeps the general 64-byte rectangle records.
// With local = s * (U_k . p) + t for a frame row U_k and sign s:  plane z = 0  <=>  U_kz . p = -s_z t_z;
// |local x| <= hx  <=>  |U_ka . p - (-s_x t_x)| <= hx.
struct FrameRects {
    std::vector<float4> blob;
    uint32_t n = 0, frames = 0, identity = 0, boxes = 0, loose = 0;
    uint32_t box_range[RT_MAX_FRAMES + 1] = {0};
    uint32_t range[RT_MAX_FRAMES * 3 + 1] = {0};
};
bool buildFrameRects(const RtPrim* prims, const std::vector<uint32_t>& order, bool find_boxes, FrameRects& out) {
    struct Rec { uint32_t frame, axis, prim; float w, ca, cb, ha, hb; };
    struct Frame { double u[3][3]; bool identity; };
    std::vector<Frame> frames;
    std::vector<Rec> recs;
    const uint32_t n = (uint32_t)order.size();
    auto rowOf = [](const float* m, int r, double v[3]) { v[0] = m[4 * r]; v[1] = m[4 * r + 1]; v[2] = m[4 * r + 2]; };
    // match row v against the rows of frame F: returns axis and sign, or -1
    auto matchRow = [](const Frame& F, const double v[3], double& sign) -> int {
        for (int k = 0; k < 3; ++k) {
            for (double sg : {1.0, -1.0}) {
                double err = 0;
                for (int c = 0; c < 3; ++c) err = std::max(err, std::fabs(v[c] - sg * F.u[k][c]));
                if (err <= 1e-6) { sign = sg; return k; }
            }
        }
        return -1;
    };
    for (uint32_t i = 0; i < n; ++i) {
        const RtPrim& p = prims[order[i]];
        if (p.kind != RT_PRIM_RECT) return false;
        double rx[3], ry[3], rz[3];
        rowOf(p.local_from_world, 0, rx); rowOf(p.local_from_world, 1, ry); rowOf(p.local_from_world, 2, rz);
        const double* rows[3] = {rx, ry, rz};
        for (int a = 0; a < 3; ++a)
            for (int b = a; b < 3; ++b) {
                double dot = rows[a][0] * rows[b][0] + rows[a][1] * rows[b][1] + rows[a][2] * rows[b][2];
                if (std::fabs(dot - (a == b ? 1.0 : 0.0)) > 1e-6) return false;  // not rigid
            }
        int f = -1, kx = -1, ky = -1, kz = -1;
        double sx = 1, sy = 1, sz = 1;
        for (uint32_t j = 0; j < frames.size() && f < 0; ++j) {
            kx = matchRow(frames[j], rx, sx); ky = matchRow(frames[j], ry, sy); kz = matchRow(frames[j], rz, sz);
            if (kx >= 0 && ky >= 0 && kz >= 0 && kx != ky && kx != kz && ky != kz) f = (int)j;
        }
        if (f < 0) {
            Frame F{};
            Frame I{{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}, true};
            double t;
            const bool world = matchRow(I, rx, t) >= 0 && matchRow(I, ry, t) >= 0 && matchRow(I, rz, t) >= 0;
            if (world) F = I;
            else { for (int c = 0; c < 3; ++c) { F.u[0][c] = rx[c]; F.u[1][c] = ry[c]; F.u[2][c] = rz[c]; } F.identity = false; }
            if (frames.size() >= RT_MAX_FRAMES) return false;
            frames.push_back(F);
            f = (int)frames.size() - 1;
            kx = matchRow(frames[f], rx, sx); ky = matchRow(frames[f], ry, sy); kz = matchRow(frames[f], rz, sz);
            if (kx < 0 || ky < 0 || kz < 0 || kx == ky || kx == kz || ky == kz) return false;
        }
        const float* m = p.local_from_world;
        const double hx = 0.5 * (double)p.param[0], hy = 0.5 * (double)p.param[1];
        const double cx = -sx * (double)m[3], cy = -sy * (double)m[7];
        Rec r{};
        r.frame = (uint32_t)f; r.axis = (uint32_t)kz; r.prim = i;
        r.w = (float)(-sz * (double)m[11]);
        if (kx < ky) { r.ca = (float)cx; r.ha = (float)hx; r.cb = (float)cy; r.hb = (float)hy; }
        else { r.ca = (float)cy; r.ha = (float)hy; r.cb = (float)cx; r.hb = (float)hx; }
        recs.push_back(r);
    }
    // the world frame first (its rays need no rotation), then groups by (frame, axis), DFS order inside a group
    int world = -1;
    for (uint32_t j = 0; j < frames.size(); ++j) if (frames[j].identity) world = (int)j;
    std::vector<uint32_t> slot(frames.size());
    uint32_t next = world >= 0 ? 1u : 0u;
    for (uint32_t j = 0; j < frames.size(); ++j) slot[j] = ((int)j == world) ? 0u : next++;
    for (auto& r : recs) r.frame = slot[r.frame];
    std::stable_sort(recs.begin(), recs.end(), [](const Rec& a, const Rec& b) { return a.frame != b.frame ? a.frame < b.frame : a.axis < b.axis; });
    // ---- frame boxes: rectangles of one frame that lie on the faces of a common box, each inside its face -------------------
    // In-plane axes of a rectangle with normal axis k: a = (k == 0 ? 1 : 0), b = (k == 2 ? 1 : 2) (the kernel's convention).
    struct Box { uint32_t frame; double lo[3], hi[3]; int face[6]; };
    std::vector<Box> boxes;
    std::vector<int> owner(recs.size(), -1);
    auto axA = [](uint32_t k) { return k == 0 ? 1 : 0; };
    auto axB = [](uint32_t k) { return k == 2 ? 1 : 2; };
    auto near = [](double x, double y) { return std::fabs(x - y) <= 4e-7 * std::max(1.0, std::max(std::fabs(x), std::fabs(y))); };  // the records are f32: a few ulps
    auto within = [](double lo, double hi, double blo, double bhi) {
        const double tol = 4e-7 * std::max(1.0, std::max(std::fabs(blo), std::fabs(bhi)));
        return lo >= blo - tol && hi <= bhi + tol;
    };
    if (find_boxes) {
        for (size_t i = 0; i < recs.size(); ++i) {
            if (owner[i] >= 0) continue;
            for (size_t j = i + 1; j < recs.size() && owner[i] < 0; ++j) {
                if (owner[j] >= 0 || recs[j].frame != recs[i].frame || recs[j].axis != recs[i].axis || near(recs[i].w, recs[j].w)) continue;
                const uint32_t k = recs[i].axis, a = axA(k), b = axB(k);
                Box B{};
                B.frame = recs[i].frame;
                for (int s = 0; s < 6; ++s) B.face[s] = -1;
                const size_t lo_i = recs[i].w < recs[j].w ? i : j, hi_i = lo_i == i ? j : i;
                B.lo[k] = recs[lo_i].w; B.hi[k] = recs[hi_i].w;
                B.lo[a] = std::min((double)recs[i].ca - recs[i].ha, (double)recs[j].ca - recs[j].ha);
                B.hi[a] = std::max((double)recs[i].ca + recs[i].ha, (double)recs[j].ca + recs[j].ha);
                B.lo[b] = std::min((double)recs[i].cb - recs[i].hb, (double)recs[j].cb - recs[j].hb);
                B.hi[b] = std::max((double)recs[i].cb + recs[i].hb, (double)recs[j].cb + recs[j].hb);
                B.face[2 * k] = (int)lo_i; B.face[2 * k + 1] = (int)hi_i;
                uint32_t count = 2;
                for (size_t r = 0; r < recs.size(); ++r) {
                    if (owner[r] >= 0 || r == i || r == j || recs[r].frame != B.frame || recs[r].axis == k) continue;
                    const uint32_t k2 = recs[r].axis, a2 = axA(k2), b2 = axB(k2);
                    int slot = -1;
                    if (near(recs[r].w, B.lo[k2])) slot = 2 * (int)k2;
                    else if (near(recs[r].w, B.hi[k2])) slot = 2 * (int)k2 + 1;
                    if (slot < 0 || B.face[slot] >= 0) continue;
                    if (!within((double)recs[r].ca - recs[r].ha, (double)recs[r].ca + recs[r].ha, B.lo[a2], B.hi[a2])) continue;
                    if (!within((double)recs[r].cb - recs[r].hb, (double)recs[r].cb + recs[r].hb, B.lo[b2], B.hi[b2])) continue;
                    B.face[slot] = (int)r;
                    ++count;
                }
                if (count < 3) continue;  // a slab test costs about two rectangle tests
                for (int s = 0; s < 6; ++s)
                    if (B.face[s] >= 0) {
                        owner[(size_t)B.face[s]] = (int)boxes.size();
                        (s & 1 ? B.hi : B.lo)[s >> 1] = recs[(size_t)B.face[s]].w;  // every face keeps its OWN plane: t is the rectangle's t
                    }
                boxes.push_back(B);
            }
        }
    }
    std::vector<Rec> loose;
    for (size_t i = 0; i < recs.size(); ++i) if (owner[i] < 0) loose.push_back(recs[i]);
    const uint32_t nb = (uint32_t)boxes.size(), nl = (uint32_t)loose.size(), nf = (uint32_t)frames.size();
    out.n = n; out.frames = nf; out.identity = world >= 0 ? 1u : 0u; out.boxes = nb; out.loose = nl;
    out.blob.assign(3 * (size_t)nf + 10 * (size_t)nb + nl + (nl + 1) / 2, make_float4(0, 0, 0, 0));
    for (uint32_t j = 0; j < nf; ++j)
        for (int k = 0; k < 3; ++k) out.blob[3 * slot[j] + k] = make_float4((float)frames[j].u[k][0], (float)frames[j].u[k][1], (float)frames[j].u[k][2], 0.0f);
    std::stable_sort(boxes.begin(), boxes.end(), [](const Box& a, const Box& b) { return a.frame < b.frame; });
    {
        uint32_t g = 0;
        for (uint32_t i = 0; i < nb; ++i) {
            while (g <= boxes[i].frame) out.box_range[g++] = i;
            float4* rec = out.blob.data() + 3 * (size_t)nf + 10 * (size_t)i;
            // lo.w = 1: every rectangle of this box covers its WHOLE face (a Cube::new, the walls of a closed room) — a crossing of the box
            // surface then lies inside the face's rectangle by construction and the kernel skips the per-face extent test
            bool full = true;
            for (int s2 = 0; s2 < 6 && full; ++s2) {
                if (boxes[i].face[s2] < 0) continue;
                const Rec& r = recs[(size_t)boxes[i].face[s2]];
                const uint32_t k2 = (uint32_t)s2 >> 1, a2 = axA(k2), b2 = axB(k2);
                full = near((double)r.ca - r.ha, boxes[i].lo[a2]) && near((double)r.ca + r.ha, boxes[i].hi[a2]) && near((double)r.cb - r.hb, boxes[i].lo[b2]) &&
                       near((double)r.cb + r.hb, boxes[i].hi[b2]);
            }
            rec[0] = make_float4((float)boxes[i].lo[0], (float)boxes[i].lo[1], (float)boxes[i].lo[2], full ? 1.0f : 0.0f);
            rec[1] = make_float4((float)boxes[i].hi[0], (float)boxes[i].hi[1], (float)boxes[i].hi[2], 0.0f);
            uint32_t* ids = reinterpret_cast<uint32_t*>(rec + 8);
            for (int s = 0; s < 8; ++s) ids[s] = RT_NO_PRIM;
            for (int s = 0; s < 6; ++s) {
                if (boxes[i].face[s] < 0) continue;
                const Rec& r = recs[(size_t)boxes[i].face[s]];
                rec[2 + s] = make_float4(r.ca, r.cb, r.ha, r.hb);
                ids[s] = r.prim;
            }
        }
        while (g <= RT_MAX_FRAMES) out.box_range[g++] = nb;
    }
    float4* lrec = out.blob.data() + 3 * (size_t)nf + 10 * (size_t)nb;
    float2* ext = reinterpret_cast<float2*>(lrec + nl);
    uint32_t g = 0;
    for (uint32_t i = 0; i < nl; ++i) {
        const Rec& r = loose[i];
        while (g <= r.frame * 3 + r.axis) out.range[g++] = i;
        float fi;
        std::memcpy(&fi, &r.prim, 4);
        lrec[i] = make_float4(r.w, r.ca, r.cb, fi);
        ext[i] = make_float2(r.ha, r.hb);
    }
    while (g <= RT_MAX_FRAMES * 3) out.range[g++] = nl;
    return true;
}

// Leaf items of the tree: every primitive that is not a face of a box primitive, then one item per box (csrc/box_prims.h).
struct TreeItems {
    std::vector<BoxGroup> groups;
    std::vector<uint32_t> ids;   // < n_prims: that input primitive; else box group ids[i] - n_prims
    std::vector<Box3> boxes;
};
void planItems(const RtSceneDesc* desc, bool find_boxes, TreeItems& it) {
    const uint32_t n = desc->n_prims;
    std::vector<int> group_of;
    if (find_boxes) it.groups = find_box_groups(desc->prims, n, group_of);
    it.ids.clear(); it.boxes.clear();
    it.ids.reserve(n); it.boxes.reserve(n);
    for (uint32_t i = 0; i < n; ++i)
        if (it.groups.empty() || group_of[i] < 0) { it.ids.push_back(i); it.boxes.push_back(primBox(desc, desc->prims[i])); }
    for (uint32_t g = 0; g < it.groups.size(); ++g) {
        Box3 b{{INFINITY, INFINITY, INFINITY}, {-INFINITY, -INFINITY, -INFINITY}};
        for (int f = 0; f < 6; ++f) {
            const Box3 c = primBox(desc, desc->prims[it.groups[g].face[f]]);
            for (int a = 0; a < 3; ++a) { b.lo[a] = std::min(b.lo[a], c.lo[a]); b.hi[a] = std::max(b.hi[a], c.hi[a]); }
        }
        it.ids.push_back(n + g);
        it.boxes.push_back(b);
    }
}

template <class T>
cudaError_t upload(DevBuf<T>& buf, const T* src, size_t n) {
    cudaError_t e = buf.ensure(n);
    if (e != cudaSuccess || n == 0) return e;
    return cudaMemcpy(buf.p, src, n * sizeof(T), cudaMemcpyHostToDevice);
}

}  // namespace

extern "C" {

const char* rt_last_error(void) { return g_error.c_str(); }
uint32_t rt_abi_version(void) { return RT_ABI_VERSION; }

static int envInt(const char* name, int lo, int hi, int dflt) {
    const char* v = std::getenv(name);
    if (!v || !*v) return dflt;
    const long x = std::atol(v);
    return (x < lo || x > hi) ? dflt : (int)x;
}

// The ONLY place the library reads its environment (rt_init / rt_init_multi call it when no options are passed).
void rt_options_from_env(RtContextOptions* co, RtSceneOptions* so) {
    if (co) {
        std::memset(co, 0, sizeof(*co));
        co->lanes = (uint32_t)envInt("RT_B200_LANES", 1, 4, 0);
        co->queue_capacity = (uint32_t)envInt("RT_B200_QUEUE_CAPACITY", 1, 1 << 26, 0);
        if (envInt("RT_B200_OVERLAP_ACCUMULATE", 0, 1, 1) == 0) co->flags |= RT_CTX_NO_ACCUMULATE_OVERLAP;
        if (envInt("RT_B200_STAGE_TIMING", 0, 1, 0) == 1) co->flags |= RT_CTX_STAGE_TIMING;
        if (envInt("RT_B200_SHADE_RAYS2", 0, 1, 1) == 0) co->flags |= RT_CTX_SHADE_ONE_RAY;
        if (envInt("RT_B200_FUSE_RECTS", 0, 1, 1) == 0) co->flags |= RT_CTX_NO_FUSED_RECTS;
        co->gen_blocks = (uint32_t)envInt("RT_B200_GEN_BLOCKS", 1, 16, 0);
        co->acc_blocks = (uint32_t)envInt("RT_B200_ACC_BLOCKS", 1, 16, 0);
    }
    if (so) {
        std::memset(so, 0, sizeof(*so));
        if (const char* bb = std::getenv("RT_B200_BVH_BUILDER")) so->bvh_builder = bb[0] == 'd' ? RT_BVH_DEVICE_LBVH : (bb[0] == 'h' ? RT_BVH_HOST_SAH : RT_BVH_AUTO);
        so->bvh_max_leaf = (uint32_t)envInt("RT_B200_BVH_MAX_LEAF", 1, 16, 0);
        so->bvh_treelet_nodes = (uint32_t)envInt("RT_B200_BVH_TREELET", 2, 1 << 20, 0);
        so->bvh_threads = (uint32_t)envInt("RT_B200_BVH_THREADS", 1, 64, 0);
        const int sn = envInt("RT_B200_STAGED_NODES", 0, 1 << 20, -1);
        so->staged_nodes = sn < 0 ? 0u : (uint32_t)sn + 1u;
        if (envInt("RT_B200_RECT_FRAMES", 0, 1, 1) == 0) so->flags |= RT_SCENE_NO_FRAME_RECTS;
        if (envInt("RT_B200_RECT_BOXES", 0, 1, 1) == 0) so->flags |= RT_SCENE_NO_FRAME_BOXES;
        if (envInt("RT_B200_BOX_PRIMS", 0, 1, 1) == 0) so->flags |= RT_SCENE_NO_BOX_PRIMS;
        if (envInt("RT_B200_WHOLE_TREE", 0, 1, 1) == 0) so->flags |= RT_SCENE_NO_WHOLE_TREE;
        if (envInt("RT_B200_STAGE_PRIMS", 0, 1, 1) == 0) so->flags |= RT_SCENE_NO_STAGED_PRIMS;
        if (envInt("RT_B200_SMALL_SCENE", 0, 1, 1) == 0) so->flags |= RT_SCENE_NO_SMALL_LIST;
    }
}

static int context_create(int device_id, const RtContextOptions& co, const RtSceneOptions& so, RtContext** out) {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) return fail(RT_ERR_NO_DEVICE, "no CUDA device (%s); this backend has no CPU fallback", cudaGetErrorString(e));
    if (device_id < 0 || device_id >= count) return fail(RT_ERR_INVALID, "device %d out of range (have %d)", device_id, count);
    RT_CUDA(cudaSetDevice(device_id));
    cudaDeviceProp prop;
    RT_CUDA(cudaGetDeviceProperties(&prop, device_id));
    if (prop.major < 10) return fail(RT_ERR_NO_DEVICE, "device %d is sm_%d%d; the kernels are built for sm_100a only", device_id, prop.major, prop.minor);
    RtContext* ctx = new (std::nothrow) RtContext();
    if (!ctx) return fail(RT_ERR_OOM, "out of host memory");
    ctx->device = device_id;
    ctx->n_lanes = co.lanes >= 1 && co.lanes <= 4 ? co.lanes : 2u;
    ctx->lanes_auto = !(co.lanes >= 1 && co.lanes <= 4);
    if (co.queue_capacity) { ctx->default_capacity = co.queue_capacity; ctx->capacity_auto = false; }
    ctx->overlap_accumulate = !(co.flags & RT_CTX_NO_ACCUMULATE_OVERLAP);
    ctx->stage_timing = (co.flags & RT_CTX_STAGE_TIMING) != 0;
    ctx->lopt.shade_rays2 = (co.flags & RT_CTX_SHADE_ONE_RAY) ? 0 : 1;
    ctx->lopt.fuse_rects = (co.flags & RT_CTX_NO_FUSED_RECTS) ? 0 : 1;
    if (co.gen_blocks >= 1 && co.gen_blocks <= 16) ctx->lopt.gen_blocks = (int)co.gen_blocks;
    if (co.acc_blocks >= 1 && co.acc_blocks <= 16) ctx->lopt.acc_blocks = (int)co.acc_blocks;
    ctx->scene_defaults = so;
    // every failure below releases what was created so far (rt_shutdown tolerates a half-built context)
    auto bail = [&](cudaError_t err, const char* what) {
        rt_shutdown(ctx);
        return fail(RT_ERR_CUDA, "%s failed: %s", what, cudaGetErrorString(err));
    };
    if ((e = kernels_num_sms(device_id, &ctx->num_sms)) != cudaSuccess) return bail(e, "cudaDeviceGetAttribute");
    if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
    if ((e = cudaStreamCreateWithFlags(&ctx->side, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
    if ((e = cudaEventCreate(&ctx->ev0)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = cudaEventCreate(&ctx->ev1)) != cudaSuccess) return bail(e, "cudaEventCreate");
    for (int i = 0; i < 2; ++i) {
        if ((e = cudaEventCreateWithFlags(&ctx->ev_shaded[i], cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
        if ((e = cudaEventCreateWithFlags(&ctx->ev_accumulated[i], cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
    }
    if ((e = cudaMallocHost(&ctx->h_status, sizeof(WaveCounters))) != cudaSuccess) return bail(e, "cudaMallocHost");
    if ((e = cudaEventCreateWithFlags(&ctx->ev_lane_begin, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
    *out = ctx;
    return RT_OK;
}

int rt_init_ex(int device_id, const RtContextOptions* options, RtContext** out) {
    if (!out) return fail(RT_ERR_INVALID, "out is null");
    *out = nullptr;
    RtContextOptions co;
    RtSceneOptions so;
    rt_options_from_env(&co, &so);  // scene defaults always come from the environment as it is NOW; never re-read later
    if (options) co = *options;
    return context_create(device_id, co, so, out);
}

int rt_init(int device_id, RtContext** out) { return rt_init_ex(device_id, nullptr, out); }

void rt_shutdown(RtContext* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    for (RtContext* lane : ctx->lanes) rt_shutdown(lane);
    ctx->lanes.clear();
    ctx->lane_sum.release(); ctx->lane_sumsq.release();
    if (ctx->ev_lane_begin) cudaEventDestroy(ctx->ev_lane_begin);
    for (int i = 0; i < 2; ++i) { ctx->qa[i].release(); ctx->qb[i].release(); ctx->qc[i].release(); }
    ctx->done[0].release(); ctx->done[1].release(); ctx->hits.release(); ctx->counters.release(); ctx->fb_sum.release(); ctx->fb_sumsq.release(); ctx->acc.release(); ctx->scratch.release();
    if (ctx->h_status) cudaFreeHost(ctx->h_status);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    for (auto& ev : ctx->stage_events) if (ev) cudaEventDestroy(ev);
    for (int i = 0; i < 2; ++i) { if (ctx->ev_shaded[i]) cudaEventDestroy(ctx->ev_shaded[i]); if (ctx->ev_accumulated[i]) cudaEventDestroy(ctx->ev_accumulated[i]); }
    if (ctx->side) cudaStreamDestroy(ctx->side);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int rt_scene_create(RtContext* ctx, const RtSceneDesc* desc, RtScene** out) { return rt_scene_create_ex(ctx, desc, nullptr, out); }

int rt_scene_create_ex(RtContext* ctx, const RtSceneDesc* desc, const RtSceneOptions* options, RtScene** out) {
    if (!ctx || !out) return fail(RT_ERR_INVALID, "ctx/out is null");
    *out = nullptr;
    if (int rc = validateDesc(desc)) return rc;
    RT_CUDA(cudaSetDevice(ctx->device));
    const RtSceneOptions so = options ? *options : ctx->scene_defaults;

    const uint32_t n_in = desc->n_prims;
    const bool small = n_in <= RT_SMALL_SCENE && !(so.flags & RT_SCENE_NO_SMALL_LIST);
    // Large scenes build their tree on the GPU (lbvh.cu: LBVH, one primitive per leaf; the node array never leaves HBM): on
    // 1 M spheres rt_scene_create takes 152 ms instead of 373 ms (16 host threads) and the tree renders within 2.4 % of the
    // SAH one (r1v).  Smaller scenes keep the host SAH build, whose trees are 10-20 % faster to traverse (book-two).
    const bool want_device = !small && (so.bvh_builder == RT_BVH_DEVICE_LBVH || (so.bvh_builder == RT_BVH_AUTO && n_in >= RT_DEVICE_BUILD_MIN_PRIMS));
    // leaf items: the primitives, with the six faces of every axis-aligned box folded into one item (host-built trees only)
    TreeItems items;
    planItems(desc, !small && !want_device && !(so.flags & RT_SCENE_NO_BOX_PRIMS), items);
    const std::vector<Box3>& boxes = items.boxes;
    const uint32_t n = (uint32_t)items.ids.size();  // leaf items
    BvhOptions opt;
    if (so.bvh_max_leaf >= 1 && so.bvh_max_leaf <= 16) opt.max_leaf = so.bvh_max_leaf;
    opt.max_always = RT_MAX_ALWAYS;
    if (so.bvh_treelet_nodes >= 2) opt.treelet_nodes = so.bvh_treelet_nodes;
    opt.top_nodes = RT_TOP_STAGED_NODES;
    opt.threads = so.bvh_threads;
    BvhResult bvh;
    LbvhResult lbvh;
    bool use_lbvh = false;
    if (small) {
        // no tree: primitives sorted by the reference's DFS rank, so a strict `<` scan reproduces its tie rule
        bvh.order.resize(n);
        for (uint32_t i = 0; i < n; ++i) bvh.order[i] = i;
        std::stable_sort(bvh.order.begin(), bvh.order.end(), [&](uint32_t a, uint32_t b) { return desc->prims[a].order < desc->prims[b].order; });
    } else {
        BvhOptions dopt = opt;
        if (!so.bvh_max_leaf) dopt.max_leaf = 1;
        std::vector<uint32_t> always;
        if (want_device) always = select_always(boxes, dopt);
        if (want_device && n - always.size() >= 1024) {
            std::vector<uint8_t> pulled(n, 0);
            for (uint32_t id : always) pulled[id] = 1;
            std::vector<uint32_t> rest_ids;
            std::vector<Box3> rest;
            rest_ids.reserve(n); rest.reserve(n);
            for (uint32_t i = 0; i < n; ++i) if (!pulled[i]) { rest_ids.push_back(i); rest.push_back(boxes[i]); }
            const cudaError_t le = build_lbvh_device(rest, (uint32_t)always.size(), dopt.max_leaf, ctx->stream, lbvh);
            if (le != cudaSuccess) { if (lbvh.d_nodes) cudaFree(lbvh.d_nodes); return fail(RT_ERR_CUDA, "device BVH build failed: %s", cudaGetErrorString(le)); }
            if (lbvh.depth + 2 > (uint32_t)RT_STACK_DEPTH) {
                // many coincident centroids can make the radix tree deeper than the traversal stack: the host builder caps its depth
                cudaFree(lbvh.d_nodes);
                lbvh.d_nodes = nullptr;
            } else {
                use_lbvh = true;
                bvh.order = always;
                for (uint32_t k : lbvh.order) bvh.order.push_back(rest_ids[k]);
                bvh.n_always = (uint32_t)always.size();
                bvh.depth = lbvh.depth;
                bvh.root_is_leaf = false;
            }
        }
        if (!use_lbvh) bvh = build_bvh(boxes, opt);
    }
    // the scene adopts the device array of the LBVH below; until then every early return must free it
    struct LbvhGuard {
        LbvhResult& r; bool armed;
        ~LbvhGuard() { if (armed && r.d_nodes) cudaFree(r.d_nodes); }
    } lbvh_guard{lbvh, use_lbvh};
    if (bvh.depth + 2 > (uint32_t)RT_STACK_DEPTH) return fail(RT_ERR_UNSUPPORTED, "BVH depth %u exceeds the traversal stack", bvh.depth);

    // device primitives: the items in leaf order, then the six faces of every box item (never in a leaf: box_hit answers for them)
    const uint32_t n_boxes = (uint32_t)items.groups.size(), n_dev = n + 6u * n_boxes;
    std::vector<uint32_t> src(n_dev, RT_NO_PRIM);  // input primitive of device primitive i; RT_NO_PRIM: a box item
    std::vector<float4> box_recs;
    box_recs.reserve(2 * (size_t)n_boxes);
    for (uint32_t i = 0; i < n; ++i) {
        const uint32_t id = items.ids[bvh.order[i]];
        if (id < n_in) { src[i] = id; continue; }
        const BoxGroup& g = items.groups[id - n_in];
        const uint32_t first = n + 6u * (uint32_t)(box_recs.size() / 2);
        for (uint32_t f = 0; f < 6; ++f) src[first + f] = g.face[f];
        float ff;
        std::memcpy(&ff, &first, 4);
        box_recs.push_back(make_float4(g.lo[0], g.lo[1], g.lo[2], ff));
        box_recs.push_back(make_float4(g.hi[0], g.hi[1], g.hi[2], 0.0f));
    }
    std::vector<float4> head(n_dev);
    std::vector<uint2> meta(n_dev);
    std::vector<RtPrim> gen;
    std::vector<float4> rects, rect_nrm;
    bool has_media = false, has_general = false;
    float always_k[RT_MAX_ALWAYS] = {0};
    uint32_t n_leaf_rects = 0, next_box = 0;
    for (uint32_t i = 0; i < n_dev; ++i) {
        if (i == n) n_leaf_rects = (uint32_t)(rects.size() / 4);
        if (src[i] == RT_NO_PRIM) {  // a box item: its record index and the device-only kind
            const uint32_t bi = next_box++, kind = RT_DPRIM_BOX;
            float fx, fy;
            std::memcpy(&fx, &bi, 4);
            std::memcpy(&fy, &kind, 4);
            head[i] = make_float4(fx, fy, 0.0f, -1.0f);
            const RtPrim& f0 = desc->prims[items.groups[items.ids[bvh.order[i]] - n_in].face[0]];
            meta[i] = make_uint2(f0.order, (uint32_t)f0.material);  // never read: hits carry the face
            continue;
        }
        const RtPrim& p = desc->prims[src[i]];
        meta[i] = make_uint2(p.order, (uint32_t)p.material);
        if (p.kind == RT_PRIM_SPHERE && p.param[0] > 0.0f && translationOnly(p)) {
            head[i] = make_float4(p.world_from_local[3], p.world_from_local[7], p.world_from_local[11], p.param[0]);
            if (i < bvh.n_always) {
                const double cx = head[i].x, cy = head[i].y, cz = head[i].z, r = head[i].w;
                always_k[i] = (float)(cx * cx + cy * cy + cz * cz - r * r);
            }
        } else {
            uint32_t gi = (uint32_t)gen.size(), kind = p.kind, ri = 0;
            if (p.kind == RT_PRIM_RECT) {
                ri = (uint32_t)(rects.size() / 4);
                const float* m = p.local_from_world;
                rects.push_back(make_float4(m[8], m[9], m[10], m[11]));
                rects.push_back(make_float4(m[0], m[1], m[2], m[3]));
                rects.push_back(make_float4(m[4], m[5], m[6], m[7]));
                rects.push_back(make_float4(0.5f * p.param[0], 0.5f * p.param[1], 0.0f, 0.0f));
                rect_nrm.push_back(make_float4(p.world_from_local[2], p.world_from_local[6], p.world_from_local[10], 0.0f));
            }
            float fx, fy, fz;
            std::memcpy(&fx, &gi, 4);
            std::memcpy(&fy, &kind, 4);
            std::memcpy(&fz, &ri, 4);
            head[i] = make_float4(fx, fy, fz, -1.0f);
            gen.push_back(p);
            has_media |= (p.kind == RT_PRIM_MEDIUM_SPHERE || p.kind == RT_PRIM_MEDIUM);
            has_general |= (p.kind != RT_PRIM_RECT);
        }
    }
    if (n_dev == n) n_leaf_rects = (uint32_t)(rects.size() / 4);
    // materials with the solid colour inlined
    std::vector<DMaterial> mats(desc->n_materials);
    for (uint32_t i = 0; i < desc->n_materials; ++i) {
        const RtMaterial& m = desc->materials[i];
        DMaterial d{};
        d.kind = m.kind; d.fuzz = m.fuzz; d.ior = m.ior; d.texture = m.texture;
        if (m.kind == RT_MAT_DIELECTRIC) { d.texture = -1; d.r = d.g = d.b = 1.0f; }
        else if (desc->textures[m.texture].kind == RT_TEX_SOLID) {
            d.texture = -1;
            d.r = desc->textures[m.texture].rgb[0]; d.g = desc->textures[m.texture].rgb[1]; d.b = desc->textures[m.texture].rgb[2];
        }
        mats[i] = d;
    }

    RtScene* s = new (std::nothrow) RtScene();
    if (!s) return fail(RT_ERR_OOM, "out of host memory");  // (lbvh_guard frees the device tree)
    s->ctx = ctx;
    s->bvh_depth = bvh.depth;
    s->d.bvh_depth = bvh.depth;
    cudaError_t e = cudaSuccess;
    static_assert(sizeof(BvhNode32) == 2 * sizeof(float4), "node layout");
    if (use_lbvh) { s->nodes.p = lbvh.d_nodes; s->nodes.n = (size_t)lbvh.n_nodes * 2; lbvh_guard.armed = false; }  // built in place: the scene adopts the device array
    else if (e == cudaSuccess) e = upload(s->nodes, reinterpret_cast<const float4*>(bvh.nodes.data()), bvh.nodes.size() * 2);
    if (e == cudaSuccess) e = upload(s->prim_head, head.data(), head.size());
    if (e == cudaSuccess) e = upload(s->prim_meta, meta.data(), meta.size());
    std::vector<uint32_t> order_to_prim;
    for (uint32_t i = 0; i < n_dev; ++i) {
        if (src[i] == RT_NO_PRIM) continue;  // a box item has no rank of its own
        if (meta[i].x >= order_to_prim.size()) order_to_prim.resize((size_t)meta[i].x + 1, RT_NO_PRIM);
        if (order_to_prim[meta[i].x] == RT_NO_PRIM) order_to_prim[meta[i].x] = i;
    }
    if (e == cudaSuccess) e = upload(s->order_to_prim, order_to_prim.data(), order_to_prim.size());
    {   // boundary surfaces go behind the general records; each medium points at its own run relative to itself
        const uint32_t first_boundary = (uint32_t)gen.size();
        for (uint32_t gi = 0; gi < first_boundary; ++gi)
            if (gen[gi].kind == RT_PRIM_MEDIUM) gen[gi].aux = first_boundary + gen[gi].aux - gi;
        gen.insert(gen.end(), desc->boundary_prims, desc->boundary_prims + desc->n_boundary_prims);
    }
    if (e == cudaSuccess) e = upload(s->gen, gen.data(), gen.size());
    if (e == cudaSuccess) e = upload(s->rects, rects.data(), rects.size());
    if (e == cudaSuccess) e = upload(s->rect_nrm, rect_nrm.data(), rect_nrm.size());
    if (e == cudaSuccess) e = upload(s->boxes, box_recs.data(), box_recs.size());
    FrameRects fr;
    const bool use_fr = small && !(so.flags & RT_SCENE_NO_FRAME_RECTS) && rects.size() / 4 == n && buildFrameRects(desc->prims, bvh.order, !(so.flags & RT_SCENE_NO_FRAME_BOXES), fr);
    if (e == cudaSuccess && use_fr) e = upload(s->fr_blob, fr.blob.data(), fr.blob.size());
    if (e == cudaSuccess) e = upload(s->mats, mats.data(), mats.size());
    if (e == cudaSuccess) e = upload(s->texs, desc->textures, (size_t)desc->n_textures);
    if (e == cudaSuccess) e = upload(s->texels, reinterpret_cast<const float4*>(desc->texels), (size_t)desc->n_texels);
    if (e != cudaSuccess) {
        rt_scene_destroy(s);
        return fail(e == cudaErrorMemoryAllocation ? RT_ERR_OOM : RT_ERR_CUDA, "scene upload failed: %s", cudaGetErrorString(e));
    }
    DScene& d = s->d;
    d.nodes = s->nodes.p; d.prim_head = s->prim_head.p; d.prim_meta = s->prim_meta.p; d.gen = s->gen.p; d.rects = s->rects.p; d.rect_nrm = s->rect_nrm.p; d.boxes = s->boxes.p;
    d.mats = s->mats.p; d.texs = s->texs.p; d.texels = s->texels.p;
    d.order_to_prim = s->order_to_prim.p; d.n_orders = (uint32_t)order_to_prim.size();
    d.n_nodes = use_lbvh ? lbvh.n_nodes : (uint32_t)bvh.nodes.size(); d.n_prims = n_dev; d.n_leaf_prims = n; d.n_leaf_rects = n_leaf_rects; d.n_boxes = n_boxes; d.n_always = bvh.n_always; d.n_gen = (uint32_t)gen.size();
    d.n_mats = desc->n_materials; d.n_texs = desc->n_textures;
    d.root_is_leaf = bvh.root_is_leaf ? 1u : 0u;
    // whole tree in shared memory when it fits; otherwise only the top levels (every ray walks them) so that
    // shared memory does not cost occupancy and L1 keeps the deeper, ray-dependent nodes
    uint32_t staged = d.n_nodes <= RT_MAX_STAGED_NODES ? d.n_nodes : RT_TOP_STAGED_NODES;
    if (so.staged_nodes) staged = std::min<uint32_t>(d.n_nodes, so.staged_nodes - 1u);
    if (use_lbvh && d.n_nodes > RT_MAX_STAGED_NODES) staged = 2u;  // the LBVH's top levels are not a prefix of its node array: only the root pair
    d.n_staged_nodes = staged & ~1u;
    d.has_media = has_media ? 1u : 0u;
    d.has_general = has_general ? 1u : 0u;
    d.n_rects = (uint32_t)(rects.size() / 4);
    d.n_small = small ? n : 0u;
    d.n_staged_prims = (n_dev <= RT_MAX_STAGED_PRIMS && d.n_rects <= RT_MAX_STAGED_RECTS) ? n_dev : 0u;
    if (so.flags & RT_SCENE_NO_STAGED_PRIMS) d.n_staged_prims = 0u;
    std::memcpy(d.always_k, always_k, sizeof(always_k));
    d.small_all_rects = (small && d.n_staged_prims == n_dev && d.n_rects == n_dev) ? 1u : 0u;
    s->staged_bytes = (size_t)d.n_staged_nodes * 32 + (size_t)d.n_staged_prims * 16 + (d.n_staged_prims ? (size_t)d.n_rects * 64 : 0);
    if (use_fr) {
        d.fr_blob = s->fr_blob.p; d.fr_n = fr.n; d.fr_frames = fr.frames; d.fr_identity = fr.identity;
        d.fr_boxes = fr.boxes; d.fr_loose = fr.loose; d.fr_blob4 = (uint32_t)fr.blob.size();
        std::memcpy(d.fr_range, fr.range, sizeof(fr.range));
        std::memcpy(d.fr_box_range, fr.box_range, sizeof(fr.box_range));
        s->fr_bytes = fr.blob.size() * sizeof(float4);
    }
    s->shade_bytes = (size_t)std::min<uint32_t>(d.n_mats, RT_MAX_STAGED_MATS) * 32 + (size_t)d.n_staged_prims * (16 + 8) +
                     (d.n_staged_prims ? (size_t)d.n_rects * 16 : 0);
    // ---- blobs the kernels stage with ONE bulk copy each (device_scene.h oct_blob / shade_blob) ----
    {
        const uint32_t n_sm = std::min<uint32_t>(d.n_mats, RT_MAX_STAGED_MATS), n_sp = d.n_staged_prims, n_sr = n_sp ? d.n_rects : 0u;
        std::vector<float4> blob((size_t)2 * n_sm + n_sp + n_sr + (n_sp + 1) / 2, make_float4(0, 0, 0, 0));
        if (n_sm) std::memcpy(blob.data(), mats.data(), (size_t)n_sm * sizeof(DMaterial));
        if (n_sp) std::memcpy(blob.data() + 2 * (size_t)n_sm, head.data(), (size_t)n_sp * sizeof(float4));
        if (n_sr) std::memcpy(blob.data() + 2 * (size_t)n_sm + n_sp, rect_nrm.data(), (size_t)n_sr * sizeof(float4));
        if (n_sp) std::memcpy(blob.data() + 2 * (size_t)n_sm + n_sp + n_sr, meta.data(), (size_t)n_sp * sizeof(uint2));
        e = upload(s->shade_blob, blob.data(), blob.size());
        d.shade_blob = s->shade_blob.p; d.shade_blob4 = (uint32_t)blob.size();
        s->shade_bytes = blob.size() * sizeof(float4);
    }
    // whole tree + leaf heads + leaf rectangles + boxes in shared memory (k_extend<., true>) when that fits one block — two blocks of 512
    // threads per SM up to 100 KB, one of 1024 threads up to 224 KB (kernels.cu)
    const bool whole_tree = e == cudaSuccess && d.n_nodes > 0u && !d.n_small && !d.root_is_leaf && !(so.flags & RT_SCENE_NO_WHOLE_TREE) &&
                            !(so.flags & RT_SCENE_NO_STAGED_PRIMS) && extend_all_smem_fits(d);
    if (whole_tree) {
        // octant-ready pair records (intersect.cuh node_step_oct): per sibling pair 7 float4 — per axis (l.lo, l.hi, r.lo, r.hi) and its
        // mirror (l.hi, l.lo, r.hi, r.lo), then (lref, lcnt, rref, rcnt) — laid out once here instead of by every block of every launch
        std::vector<float4> host_nodes;
        const float4* nodes4 = reinterpret_cast<const float4*>(bvh.nodes.data());
        if (use_lbvh) {  // (a small scene forced onto the device builder: its nodes live in HBM only)
            host_nodes.resize((size_t)d.n_nodes * 2);
            e = cudaMemcpy(host_nodes.data(), s->nodes.p, host_nodes.size() * sizeof(float4), cudaMemcpyDeviceToHost);
            nodes4 = host_nodes.data();
        }
        const uint32_t n_pairs = (d.n_nodes + 1u) / 2u;
        std::vector<float4> blob((size_t)7 * n_pairs + n + 4 * (size_t)n_leaf_rects + box_recs.size(), make_float4(0, 0, 0, 0));
        const float4 zero = make_float4(0, 0, 0, 0);
        for (uint32_t j = 0; j < n_pairs && e == cudaSuccess; ++j) {
            const bool has_right = 2u * j + 1u < d.n_nodes;
            const float4 ll = nodes4[4u * j], lh = nodes4[4u * j + 1u], rl = has_right ? nodes4[4u * j + 2u] : zero, rh = has_right ? nodes4[4u * j + 3u] : zero;
            float4* o = blob.data() + 7 * (size_t)j;
            o[0] = make_float4(ll.x, lh.x, rl.x, rh.x); o[1] = make_float4(lh.x, ll.x, rh.x, rl.x);
            o[2] = make_float4(ll.y, lh.y, rl.y, rh.y); o[3] = make_float4(lh.y, ll.y, rh.y, rl.y);
            o[4] = make_float4(ll.z, lh.z, rl.z, rh.z); o[5] = make_float4(lh.z, ll.z, rh.z, rl.z);
            o[6] = make_float4(ll.w, lh.w, rl.w, rh.w);
        }
        std::memcpy(blob.data() + 7 * (size_t)n_pairs, head.data(), (size_t)n * sizeof(float4));
        if (n_leaf_rects) std::memcpy(blob.data() + 7 * (size_t)n_pairs + n, rects.data(), 4 * (size_t)n_leaf_rects * sizeof(float4));
        if (n_boxes) std::memcpy(blob.data() + 7 * (size_t)n_pairs + n + 4 * (size_t)n_leaf_rects, box_recs.data(), box_recs.size() * sizeof(float4));
        if (e == cudaSuccess) e = upload(s->oct_blob, blob.data(), blob.size());
        d.oct_blob = s->oct_blob.p; d.oct_blob4 = (uint32_t)blob.size();
    }
    if (e != cudaSuccess) {
        rt_scene_destroy(s);
        return fail(RT_ERR_CUDA, "scene upload failed: %s", cudaGetErrorString(e));
    }
    e = kernels_configure_scene(ctx->device, d, s->staged_bytes, s->shade_bytes, s->fr_bytes, s->launch);
    if (e != cudaSuccess) {
        rt_scene_destroy(s);
        return fail(RT_ERR_CUDA, "kernel configuration failed: %s", cudaGetErrorString(e));
    }
    *out = s;
    return RT_OK;
}

void rt_scene_destroy(RtScene* s) {
    if (!s) return;
    if (s->ctx) cudaSetDevice(s->ctx->device);
    s->nodes.release(); s->prim_head.release(); s->prim_meta.release(); s->order_to_prim.release(); s->gen.release(); s->rects.release(); s->rect_nrm.release(); s->boxes.release(); s->fr_blob.release(); s->oct_blob.release(); s->shade_blob.release(); s->mats.release(); s->texs.release(); s->texels.release();
    delete s;
}

int rt_scene_bvh_info(const RtScene* s, uint32_t* n_nodes, uint32_t* n_always, uint32_t* max_depth) {
    if (!s) return fail(RT_ERR_INVALID, "scene is null");
    if (n_nodes) *n_nodes = s->d.n_nodes;
    if (n_always) *n_always = s->d.n_always;
    if (max_depth) *max_depth = s->bvh_depth;
    return RT_OK;
}

int rt_scene_frame_rect_info(const RtScene* s, uint32_t* n_frames, uint32_t* n_boxes, uint32_t* n_loose) {
    if (!s) return fail(RT_ERR_INVALID, "scene is null");
    const bool on = s->d.fr_n > 0;
    if (n_frames) *n_frames = on ? s->d.fr_frames : 0u;
    if (n_boxes) *n_boxes = on ? s->d.fr_boxes : 0u;
    if (n_loose) *n_loose = on ? s->d.fr_loose : 0u;
    return RT_OK;
}

int rt_scene_extend_info(const RtScene* s, uint32_t* n_box_prims, uint32_t* n_leaf_items, uint32_t* whole_tree_threads) {
    if (!s) return fail(RT_ERR_INVALID, "scene is null");
    if (n_box_prims) *n_box_prims = s->d.n_boxes;
    if (n_leaf_items) *n_leaf_items = s->d.n_leaf_prims;
    if (whole_tree_threads) *whole_tree_threads = s->launch.all_staged ? (uint32_t)s->launch.all_block : 0u;
    return RT_OK;
}

int rt_scene_plan(const RtSceneDesc* desc, const RtSceneOptions* options, RtScenePlan* out) {
    if (!out) return fail(RT_ERR_INVALID, "out is null");
    if (int rc = validateDesc(desc)) return rc;
    std::memset(out, 0, sizeof(*out));
    RtSceneOptions so{};
    if (options) so = *options;
    const uint32_t n_in = desc->n_prims;
    const bool small = n_in <= RT_SMALL_SCENE && !(so.flags & RT_SCENE_NO_SMALL_LIST);
    const bool want_device = !small && (so.bvh_builder == RT_BVH_DEVICE_LBVH || (so.bvh_builder == RT_BVH_AUTO && n_in >= RT_DEVICE_BUILD_MIN_PRIMS));
    TreeItems items;
    planItems(desc, !small && !want_device && !(so.flags & RT_SCENE_NO_BOX_PRIMS), items);
    out->n_box_prims = (uint32_t)items.groups.size();
    out->n_leaf_items = (uint32_t)items.ids.size();
    if (small || want_device) return RT_OK;
    BvhOptions opt;
    if (so.bvh_max_leaf >= 1 && so.bvh_max_leaf <= 16) opt.max_leaf = so.bvh_max_leaf;
    opt.max_always = RT_MAX_ALWAYS;
    if (so.bvh_treelet_nodes >= 2) opt.treelet_nodes = so.bvh_treelet_nodes;
    opt.top_nodes = RT_TOP_STAGED_NODES;
    opt.threads = so.bvh_threads;
    const BvhResult bvh = build_bvh(items.boxes, opt);
    out->n_nodes = (uint32_t)bvh.nodes.size();
    out->bvh_depth = bvh.depth;
    out->n_always = bvh.n_always;
    DScene d{};
    d.n_nodes = out->n_nodes; d.bvh_depth = bvh.depth; d.n_leaf_prims = out->n_leaf_items; d.n_boxes = out->n_box_prims;
    for (uint32_t id : items.ids) d.n_leaf_rects += (id < n_in && desc->prims[id].kind == RT_PRIM_RECT) ? 1u : 0u;
    if (!bvh.root_is_leaf && d.n_nodes > 0u && !(so.flags & (RT_SCENE_NO_WHOLE_TREE | RT_SCENE_NO_STAGED_PRIMS)) && extend_all_smem_fits(d)) {
        const bool half = extend_all_smem_bytes(d, RT_EXTEND_ALL_BLOCK) <= 100u * 1024u;
        out->whole_tree_threads = half ? RT_EXTEND_ALL_BLOCK : RT_EXTEND_ALL_BIG_BLOCK;
        out->whole_tree_smem = extend_all_smem_bytes(d, out->whole_tree_threads);
    }
    return RT_OK;
}

uint64_t rt_scene_device_bytes(const RtScene* s) {
    if (!s) return 0;
    return (uint64_t)s->nodes.n * sizeof(float4) + (uint64_t)s->prim_head.n * sizeof(float4) + (uint64_t)s->prim_meta.n * sizeof(uint2) +
           (uint64_t)s->gen.n * sizeof(RtPrim) + (uint64_t)s->order_to_prim.n * sizeof(uint32_t) + (uint64_t)(s->rects.n + s->rect_nrm.n + s->boxes.n) * sizeof(float4) + (uint64_t)s->mats.n * sizeof(DMaterial) + (uint64_t)s->texs.n * sizeof(RtTexture) +
           (uint64_t)s->texels.n * sizeof(float4);
}

int rt_deinterleave_rows(RtContext* ctx, const float* d_gathered, float* d_full, uint32_t width, uint32_t height, uint32_t n_shards, void* cuda_stream) {
    if (!ctx || !d_gathered || !d_full || !width || !height || !n_shards) return fail(RT_ERR_INVALID, "bad argument");
    RT_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : ctx->stream;
    launch_deinterleave(ctx->num_sms, d_gathered, d_full, width, height, n_shards, st);
    RT_CUDA(cudaGetLastError());
    if (!cuda_stream) RT_CUDA(cudaStreamSynchronize(st));
    return RT_OK;
}

static int render_core(RtContext* ctx, RtScene* scene, const RtCamera* cam, const RtRayIn* d_user_rays, const RtRenderParams* prm, float* d_sum,
                       float* d_sumsq, void* cuda_stream, RtStats* stats);
static int render_lanes(RtScene* scene, const RtCamera* cam, const RtRenderParams* prm, float* d_sum, float* d_sumsq, void* cuda_stream, RtStats* stats);

int rt_render_device(RtScene* scene, const RtCamera* cam, const RtRenderParams* prm, float* d_sum, float* d_sumsq, void* cuda_stream, RtStats* stats) {
    if (!scene || !cam || !prm || !d_sum) return fail(RT_ERR_INVALID, "null argument");
    return render_lanes(scene, cam, prm, d_sum, d_sumsq, cuda_stream, stats);
}

// Batched `render::color(&ray, world, max_depth)` (src/render.rs:5-29) on caller rays: ray i is "pixel" i of an n x 1 image,
// `spp` independent evaluations of color() per ray (Philox key = (i, sample)), sums returned like rt_render.
int rt_color_rays(RtScene* scene, const RtRayIn* rays, uint64_t n, uint32_t spp, uint32_t max_depth, uint64_t seed, float* out_sum, float* out_sumsq,
                  RtStats* stats) {
    if (!scene || !rays || !out_sum) return fail(RT_ERR_INVALID, "null argument");
    if (n == 0 || n >= (1ull << 32)) return fail(RT_ERR_INVALID, "ray count must be in [1, 2^32)");
    RtContext* ctx = scene->ctx;
    RT_CUDA(cudaSetDevice(ctx->device));
    RT_CUDA(ctx->scratch.ensure(n * sizeof(RtRayIn)));
    RT_CUDA(ctx->fb_sum.ensure(n * 3));
    if (out_sumsq) RT_CUDA(ctx->fb_sumsq.ensure(n * 3));
    RT_CUDA(cudaMemcpyAsync(ctx->scratch.p, rays, n * sizeof(RtRayIn), cudaMemcpyHostToDevice, ctx->stream));
    RtRenderParams prm{};
    prm.width = (uint32_t)n; prm.height = 1; prm.spp = spp; prm.max_depth = max_depth; prm.seed = seed; prm.row_begin = 0; prm.row_stride = 1;
    RtCamera cam{};
    const int rc = render_core(ctx, scene, &cam, reinterpret_cast<const RtRayIn*>(ctx->scratch.p), &prm, ctx->fb_sum.p, out_sumsq ? ctx->fb_sumsq.p : nullptr, nullptr, stats);
    if (rc != RT_OK) return rc;
    RT_CUDA(cudaMemcpy(out_sum, ctx->fb_sum.p, n * 3 * sizeof(float), cudaMemcpyDeviceToHost));
    if (out_sumsq) RT_CUDA(cudaMemcpy(out_sumsq, ctx->fb_sumsq.p, n * 3 * sizeof(float), cudaMemcpyDeviceToHost));
    return RT_OK;
}

}  // extern "C" (render_core is internal)

// Several wavefronts per render.  One wavefront leaves the SMs idle at every kernel boundary (the tail of a persistent grid,
// the launch of the next: ~7-12 % of each launch, ncu cycles_active vs elapsed) and during the drain at the end of the image.
// rt_render_device therefore splits its rows over `n_lanes` interleaved sub-shards — the same partition as the multi-GPU one — and
// runs one complete wavefront per lane, each with its own streams, queues and accumulators, driven by its own host thread; the
// blocks of one lane's kernel start as the other lane's blocks retire.  Measured (profiles/r2b_lanes_sweep.txt):
// cornell-box 2987 -> 3613 Mpaths/s, book-one 1080p 2930 -> 3423.  Philox keys are functions of the global pixel, so the image
// does not depend on the number of lanes.  Instrumented renders (stage timing, work counters) run as one lane so that a timed
// launch has the device to itself.
static int render_lanes(RtScene* scene, const RtCamera* cam, const RtRenderParams* prm, float* d_sum, float* d_sumsq, void* cuda_stream, RtStats* stats) {
    RtContext* ctx = scene->ctx;
    if (prm->row_stride == 0 || prm->row_begin >= prm->height || prm->width == 0) return render_core(ctx, scene, cam, nullptr, prm, d_sum, d_sumsq, cuda_stream, stats);
    const uint32_t rows_local = (prm->height - prm->row_begin + prm->row_stride - 1) / prm->row_stride;
    uint32_t L = ctx->n_lanes;
    // Small all-rectangle scenes run three lanes: their iterations are short (extend is fused into shade: two kernels + accumulate) and
    // the ~50 nearly empty iterations that drain a depth-50 wavefront weigh more — cornell-box 600x600x512: 4776 -> 4833 Mpaths/s, a
    // 1/8 shard (one GPU of eight) 4190 -> 4313 (profiles/r2_41_*, r2_42_*); neutral to slightly negative on the tree scenes
    if (ctx->lanes_auto && scene->d.fr_n) L = 3;
    // Large renders of tree scenes run four lanes (from 2^29 paths: book-two 800x800x1024 2045 -> 2078 Mpaths/s, book-one 1080p x 256
    // 5786 -> 5873; below that the extra lanes' drain costs what their overlap buys), the very large ones (from 2^31 paths: book-one
    // 3840x2160x1024, 8.5 G paths) with 8 Mi-ray queues, 4.4 GB of wavefront state — 5891 -> 5980 Mpaths/s (profiles/r2_44_*)
    const uint64_t n_paths = (uint64_t)prm->width * rows_local * prm->spp;
    const bool huge = !scene->d.fr_n && n_paths >= (1ull << 31);
    if (ctx->lanes_auto && !scene->d.fr_n && n_paths >= (1ull << 29)) L = 4;
    const uint32_t lane_capacity = (huge && ctx->capacity_auto) ? (1u << 23) : ctx->default_capacity;
    if ((prm->flags & (RT_RENDER_STAGE_TIMING | RT_RENDER_COUNT_WORK)) || ctx->stage_timing) L = 1;
    if (rows_local < 4u * L || prm->max_depth == 0) L = 1;
    if ((uint64_t)prm->width * rows_local * prm->spp < (uint64_t)L * (1u << 20)) L = 1;  // too little work to fill two queues
    if ((uint64_t)prm->row_stride * L > 0xFFFFFFFFull) L = 1;
    if (L == 1) return render_core(ctx, scene, cam, nullptr, prm, d_sum, d_sumsq, cuda_stream, stats);

    RT_CUDA(cudaSetDevice(ctx->device));
    while (ctx->lanes.size() < L) {
        RtContext* lane = nullptr;
        RtContextOptions co{};
        co.lanes = 1;
        co.queue_capacity = ctx->default_capacity;
        co.flags = ctx->overlap_accumulate ? 0u : (uint32_t)RT_CTX_NO_ACCUMULATE_OVERLAP;
        const int rc = context_create(ctx->device, co, ctx->scene_defaults, &lane);
        if (rc != RT_OK) return rc;
        lane->lopt = ctx->lopt;
        ctx->lanes.push_back(lane);
    }
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : ctx->stream;
    const uint32_t rows_padded = (rows_local + L - 1) / L;
    const size_t shard = (size_t)rows_padded * prm->width * 3;
    RT_CUDA(ctx->lane_sum.ensure(shard * L));
    if (d_sumsq) RT_CUDA(ctx->lane_sumsq.ensure(shard * L));
    RT_CUDA(cudaEventRecord(ctx->ev0, st));
    RT_CUDA(cudaEventRecord(ctx->ev_lane_begin, st));
    std::vector<RtStats> lane_stats(L);
    std::vector<int> lane_rc(L, RT_OK);
    std::vector<std::string> lane_err(L);
    {
        ctx->pool.run(L, [&](uint32_t l) {
            RtContext* lane = ctx->lanes[l];
            RtRenderParams lp = *prm;
            lp.row_begin = prm->row_begin + l * prm->row_stride;
            lp.row_stride = prm->row_stride * L;
            if (!lp.queue_capacity) lp.queue_capacity = lane_capacity;
            cudaSetDevice(ctx->device);
            cudaError_t e = cudaStreamWaitEvent(lane->stream, ctx->ev_lane_begin, 0);  // after whatever the caller queued before the render
            if (e != cudaSuccess) { lane_rc[l] = RT_ERR_CUDA; lane_err[l] = cudaGetErrorString(e); return; }
            lane_rc[l] = render_core(lane, scene, cam, nullptr, &lp, ctx->lane_sum.p + shard * l, d_sumsq ? ctx->lane_sumsq.p + shard * l : nullptr, nullptr,
                                     &lane_stats[l]);
            if (lane_rc[l] != RT_OK) lane_err[l] = g_error;
        });
    }
    for (uint32_t l = 0; l < L; ++l)
        if (lane_rc[l] != RT_OK) return fail(lane_rc[l], "%s", lane_err[l].c_str());
    // render_core returned after its lane's stream drained: the shards are complete; interleave them on the caller's stream
    launch_deinterleave(ctx->num_sms, ctx->lane_sum.p, d_sum, prm->width, rows_local, L, st);
    if (d_sumsq) launch_deinterleave(ctx->num_sms, ctx->lane_sumsq.p, d_sumsq, prm->width, rows_local, L, st);
    RT_CUDA(cudaEventRecord(ctx->ev1, st));
    RT_CUDA(cudaEventSynchronize(ctx->ev1));
    RT_CUDA(cudaGetLastError());
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        float ms = 0;
        RT_CUDA(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
        for (const RtStats& ls : lane_stats) {
            stats->paths += ls.paths; stats->rays += ls.rays; stats->launches += ls.launches;
            stats->iterations = std::max(stats->iterations, ls.iterations);
        }
        stats->launches += d_sumsq ? 2 : 1;
        stats->ms_total = ms;
        stats->bvh_nodes = scene->d.n_nodes; stats->bvh_always = scene->d.n_always;
    }
    return RT_OK;
}

static int render_core(RtContext* ctx, RtScene* scene, const RtCamera* cam, const RtRayIn* d_user_rays, const RtRenderParams* prm, float* d_sum,
                       float* d_sumsq, void* cuda_stream, RtStats* stats) {
    if (prm->width == 0 || prm->height == 0 || prm->spp == 0) return fail(RT_ERR_INVALID, "width/height/spp must be > 0");
    if (prm->row_stride == 0 || prm->row_begin >= prm->height) return fail(RT_ERR_INVALID, "bad row_begin/row_stride");
    if (prm->max_depth > 255u) return fail(RT_ERR_UNSUPPORTED, "max_depth %u > 255", prm->max_depth);
    if (prm->spp >= (1u << 24)) return fail(RT_ERR_UNSUPPORTED, "spp %u >= 2^24", prm->spp);
    if ((uint64_t)prm->width * prm->height >= (1ull << 32)) return fail(RT_ERR_UNSUPPORTED, "image too large");
    RT_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : ctx->stream;

    RenderConst rc{};
    rc.cam = *cam;
    rc.width = prm->width; rc.height = prm->height; rc.spp = prm->spp; rc.max_depth = prm->max_depth;
    rc.seed_lo = (uint32_t)prm->seed; rc.seed_hi = (uint32_t)(prm->seed >> 32);
    rc.row_begin = prm->row_begin; rc.row_stride = prm->row_stride;
    rc.rows_local = (prm->height - prm->row_begin + prm->row_stride - 1) / prm->row_stride;
    uint32_t cap = prm->queue_capacity ? prm->queue_capacity : ctx->default_capacity;
    cap = std::max(1024u, std::min(cap, 1u << 26)) & ~31u;
    rc.capacity = cap;
    rc.pixels_local = (uint64_t)rc.width * rc.rows_local;
    rc.paths_total = rc.pixels_local * rc.spp;
    rc.user_rays = d_user_rays;
    rc.pixel_major = (prm->flags & RT_RENDER_SAMPLE_MAJOR) ? 0u : 1u;

    for (int i = 0; i < 2; ++i) { RT_CUDA(ctx->qa[i].ensure(cap)); RT_CUDA(ctx->qb[i].ensure(cap)); RT_CUDA(ctx->qc[i].ensure(cap)); }
    RT_CUDA(ctx->hits.ensure(cap));
    RT_CUDA(ctx->done[0].ensure(cap));
    RT_CUDA(ctx->done[1].ensure(cap));
    RT_CUDA(ctx->counters.ensure(1));
    if (scene->launch.device != ctx->device) return fail(RT_ERR_INVALID, "scene lives on device %d, context on device %d", scene->launch.device, ctx->device);
    LaunchOptions lopt = ctx->lopt;
    if (prm->flags & RT_RENDER_EXTEND_PER_THREAD) lopt.persistent = 0;
    if (prm->flags & RT_RENDER_EXTEND_PERSISTENT) lopt.persistent = 1;

    const size_t fb_elems = (size_t)rc.pixels_local * 3;
    RT_CUDA(ctx->acc.ensure(fb_elems * 2));
    double* acc_sum = ctx->acc.p;
    double* acc_sumsq = d_sumsq ? ctx->acc.p + fb_elems : nullptr;
    RT_CUDA(cudaMemsetAsync(ctx->counters.p, 0, sizeof(WaveCounters), st));
    RT_CUDA(cudaMemsetAsync(ctx->acc.p, 0, fb_elems * (d_sumsq ? 2 : 1) * sizeof(double), st));

    if (prm->max_depth == 0) {  // color(ray, world, 0) is black without tracing (src/render.rs:6-8)
        RT_CUDA(cudaMemsetAsync(d_sum, 0, fb_elems * sizeof(float), st));
        if (d_sumsq) RT_CUDA(cudaMemsetAsync(d_sumsq, 0, fb_elems * sizeof(float), st));
        RT_CUDA(cudaStreamSynchronize(st));
        if (stats) { std::memset(stats, 0, sizeof(*stats)); stats->paths = rc.paths_total; stats->bvh_nodes = scene->d.n_nodes; stats->bvh_always = scene->d.n_always; }
        return RT_OK;
    }

    const RayQueue q[2] = {{ctx->qa[0].p, ctx->qb[0].p, ctx->qc[0].p}, {ctx->qa[1].p, ctx->qb[1].p, ctx->qc[1].p}};
    WaveCounters* wc = ctx->counters.p;
    float ms_stage[4] = {0, 0, 0, 0};
    uint64_t launches = 0;
    const uint64_t min_iters = (rc.paths_total + cap - 1) / cap;
    const uint32_t poll = 8;
    uint32_t max_depth_iters = prm->max_depth ? prm->max_depth : 1;
    // hard bound on the number of iterations: every iteration either issues `cap` paths or the queue is draining
    const uint64_t iter_limit = (min_iters + 2) * (uint64_t)(max_depth_iters + 1) + 64;

    const bool timing = ctx->stage_timing || (prm->flags & RT_RENDER_STAGE_TIMING);
    const bool count_work = (prm->flags & RT_RENDER_COUNT_WORK) != 0;
    const bool overlap = ctx->overlap_accumulate;
    cudaStream_t acc_st = overlap ? ctx->side : st;
    // stage timing brackets every `tstride`-th iteration only: six event records per iteration cost ~20 % of a
    // cornell-box render, which would handicap the very step that is being measured
    const uint64_t tstride = (prm->flags & RT_RENDER_TIME_EVERY_ITERATION) ? 1 : 8;
    uint64_t n_timed = 0;
    auto stage_event = [&](uint64_t slot, int which, cudaStream_t on) -> cudaError_t {
        const size_t idx = (size_t)slot * 6 + (size_t)which;
        while (ctx->stage_events.size() <= idx) {
            cudaEvent_t ev = nullptr;
            const cudaError_t e = cudaEventCreate(&ev);
            if (e != cudaSuccess) return e;
            ctx->stage_events.push_back(ev);
        }
        return cudaEventRecord(ctx->stage_events[idx], on);
    };

    RT_CUDA(cudaEventRecord(ctx->ev0, st));
    Launch L{&scene->launch, &lopt, RT_WORK_UNKNOWN, st};
    const bool fused = extend_is_fused(L, scene->d, count_work);  // extend inside k_shade: one launch fewer per iteration
    Launch La = L;  // accumulate's launch site (its own stream when overlapped)
    La.st = acc_st;
    uint64_t k = 0;
    bool finished = false;
    while (!finished) {
        const uint32_t ku = (uint32_t)k;
        const RayQueue& qin = q[k & 1];
        const RayQueue& qout = q[(k + 1) & 1];
        const bool timed = timing && (k % tstride == 0);
        const uint64_t ts = n_timed;
        // tail: every path has been issued (known from a status poll) — k_generate would only roll the counters; extend does that
        // itself (kernels.cu tail_roll) and one of the three dependent launches of a draining iteration is gone
        const bool tail = RT_TAIL_MODE && L.work_hint != RT_WORK_UNKNOWN;
        const uint32_t kx = tail ? (ku | RT_K_TAIL) : ku;
        if (timed) RT_CUDA(stage_event(ts, 0, st));
        if (!tail) launch_generate(L, rc, qin, wc, ku);
        if (timed) RT_CUDA(stage_event(ts, 1, st));
        launch_extend(L, scene->d, qin, ctx->hits.p, wc, kx, rc.seed_lo, rc.seed_hi, count_work);
        if (timed) RT_CUDA(stage_event(ts, 2, st));
        // shade_k refills done[k & 1]: accumulate_{k-2} must be through with it
        if (overlap && k >= 2) RT_CUDA(cudaStreamWaitEvent(st, ctx->ev_accumulated[k & 1], 0));
        launch_shade(L, scene->d, rc, qin, qout, ctx->hits.p, ctx->done[k & 1].p, wc, kx, count_work);
        if (timed) RT_CUDA(stage_event(ts, 3, st));
        if (overlap) {
            RT_CUDA(cudaEventRecord(ctx->ev_shaded[k & 1], st));
            RT_CUDA(cudaStreamWaitEvent(acc_st, ctx->ev_shaded[k & 1], 0));
        }
        if (timed) RT_CUDA(stage_event(ts, 4, acc_st));
        launch_accumulate(La, rc, ctx->done[k & 1].p, wc, ku, acc_sum, acc_sumsq);
        if (timed) { RT_CUDA(stage_event(ts, 5, acc_st)); ++n_timed; }
        if (overlap) RT_CUDA(cudaEventRecord(ctx->ev_accumulated[k & 1], acc_st));
        launches += (tail ? 3 : 4) - (fused ? 1 : 0);
        // a launch that cannot start (shared memory, bad configuration) must surface here, not as "did not terminate" 100 k iterations later
        if (k == 0) RT_CUDA(cudaPeekAtLastError());
        ++k;
        if (k >= min_iters && ((k - min_iters) % poll == 0)) {
            RT_CUDA(cudaPeekAtLastError());
            RT_CUDA(cudaMemcpyAsync(ctx->h_status, wc, sizeof(WaveCounters), cudaMemcpyDeviceToHost, st));
            RT_CUDA(cudaStreamSynchronize(st));
            if (ctx->h_status->st_issued == rc.paths_total && !count_work) L.work_hint = La.work_hint = ctx->h_status->st_total;  // the queue only shrinks from here
            if (ctx->h_status->st_issued == rc.paths_total && ctx->h_status->st_total == 0) finished = true;
            else if (k > iter_limit) return fail(RT_ERR_CUDA, "wavefront did not terminate after %llu iterations", (unsigned long long)k);
        }
    }
    if (overlap) {  // the last two accumulates join the render stream before the accumulators are read
        RT_CUDA(cudaStreamWaitEvent(st, ctx->ev_accumulated[0], 0));
        if (k >= 2) RT_CUDA(cudaStreamWaitEvent(st, ctx->ev_accumulated[1], 0));
    }
    launch_finish(ctx->num_sms, acc_sum, fb_elems, d_sum, st);
    if (d_sumsq) launch_finish(ctx->num_sms, acc_sumsq, fb_elems, d_sumsq, st);
    launches += d_sumsq ? 2 : 1;
    RT_CUDA(cudaEventRecord(ctx->ev1, st));
    RT_CUDA(cudaEventSynchronize(ctx->ev1));
    RT_CUDA(cudaGetLastError());
    if (timing) {
        for (uint64_t it = 0; it < n_timed; ++it) {
            const int from[4] = {0, 1, 2, 4}, to[4] = {1, 2, 3, 5};  // generate, extend, shade on the render stream; accumulate on its own
            for (int i = 0; i < 4; ++i) {
                float ms = 0;
                RT_CUDA(cudaEventElapsedTime(&ms, ctx->stage_events[it * 6 + from[i]], ctx->stage_events[it * 6 + to[i]]));
                ms_stage[i] += ms;
            }
        }
        // estimate of the whole render: the timed iterations are a uniform 1-in-tstride sample
        if (n_timed) for (int i = 0; i < 4; ++i) ms_stage[i] *= (float)((double)k / (double)n_timed);
    }
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        float ms = 0;
        RT_CUDA(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
        stats->paths = rc.paths_total;
        stats->rays = ctx->h_status->st_rays;
        stats->iterations = k;
        stats->launches = launches;
        stats->ms_total = ms;
        stats->ms_generate = ms_stage[0]; stats->ms_extend = ms_stage[1]; stats->ms_shade = ms_stage[2]; stats->ms_accumulate = ms_stage[3];
        stats->bvh_nodes = scene->d.n_nodes; stats->bvh_always = scene->d.n_always;
        stats->box_tests = ctx->h_status->st_box_tests; stats->prim_tests = ctx->h_status->st_prim_tests;
        stats->timed_iterations = n_timed;
    }
    return RT_OK;
}

extern "C" {

int rt_render(RtScene* scene, const RtCamera* cam, const RtRenderParams* prm, float* out_sum, float* out_sumsq, RtStats* stats) {
    if (!scene || !prm || !out_sum) return fail(RT_ERR_INVALID, "null argument");
    if (prm->row_stride == 0 || prm->row_begin >= prm->height) return fail(RT_ERR_INVALID, "bad row_begin/row_stride");
    RtContext* ctx = scene->ctx;
    RT_CUDA(cudaSetDevice(ctx->device));
    const uint32_t rows = (prm->height - prm->row_begin + prm->row_stride - 1) / prm->row_stride;
    const size_t n = (size_t)prm->width * rows * 3;
    RT_CUDA(ctx->fb_sum.ensure(n));
    if (out_sumsq) RT_CUDA(ctx->fb_sumsq.ensure(n));
    const int rc = rt_render_device(scene, cam, prm, ctx->fb_sum.p, out_sumsq ? ctx->fb_sumsq.p : nullptr, nullptr, stats);
    if (rc != RT_OK) return rc;
    RT_CUDA(cudaMemcpy(out_sum, ctx->fb_sum.p, n * sizeof(float), cudaMemcpyDeviceToHost));
    if (out_sumsq) RT_CUDA(cudaMemcpy(out_sumsq, ctx->fb_sumsq.p, n * sizeof(float), cudaMemcpyDeviceToHost));
    return RT_OK;
}

// ---- rt_render_noalloc: a render that allocates nothing ------------------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc (code under
// CUDA graph capture, tight serving loops, harnesses that forbid it).  The wavefront's scratch — two ray queues, the hit and
// finished-path records, the ring counters, the f64 accumulators and the device framebuffers — lives in ONE file-scope device array of
// RT_NOALLOC_SCRATCH_BYTES that every GPU which loads this library holds until exit.  In host code `g_noalloc_scratch` is only a
// stand-in for the symbol: the device address comes from cudaGetSymbolAddress, once per device (noalloc_scratch below).
static __device__ __align__(256) unsigned char g_noalloc_scratch[RT_NOALLOC_SCRATCH_BYTES];

namespace {
std::mutex g_noalloc_mu;
unsigned char* g_noalloc_ptr[64] = {nullptr};  // per device: device address of g_noalloc_scratch

// the calling thread has `device` current
int noalloc_scratch(int device, unsigned char** out) {
    if (device < 0 || device >= 64) return fail(RT_ERR_INVALID, "device %d out of range", device);
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        unsigned char* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) return fail(RT_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch) failed on device %d: %s", device, cudaGetErrorString(e));
        g_noalloc_ptr[device] = p;
    }
    *out = g_noalloc_ptr[device];
    return RT_OK;
}

// Lends slices of the static array to a context's wavefront buffers for the duration of one render and puts the context's own
// (pointer, size) pairs back afterwards, whichever way the render returns.  DevBuf::ensure is a no-op for a buffer that is large enough.
struct NoallocLoan {
    RtContext* ctx;
    DevBuf<float4> qa[2], qb[2], qc[2], done[2];
    DevBuf<float2> hits;
    DevBuf<WaveCounters> counters;
    DevBuf<double> acc;
    bool stage_timing;
    NoallocLoan(RtContext* c, unsigned char* base) : ctx(c) {
        for (int i = 0; i < 2; ++i) { qa[i] = c->qa[i]; qb[i] = c->qb[i]; qc[i] = c->qc[i]; done[i] = c->done[i]; }
        hits = c->hits; counters = c->counters; acc = c->acc; stage_timing = c->stage_timing;
        const size_t cap = RT_NOALLOC_QUEUE_RAYS, q = cap * sizeof(float4);
        size_t off = 0;
        auto take = [&](size_t bytes) { unsigned char* p = base + off; off += (bytes + 255u) & ~(size_t)255u; return p; };
        for (int i = 0; i < 2; ++i) {
            c->qa[i].p = (float4*)take(q); c->qa[i].n = cap;
            c->qb[i].p = (float4*)take(q); c->qb[i].n = cap;
            c->qc[i].p = (float4*)take(q); c->qc[i].n = cap;
            c->done[i].p = (float4*)take(q); c->done[i].n = cap;
        }
        c->hits.p = (float2*)take(cap * sizeof(float2)); c->hits.n = cap;
        c->counters.p = (WaveCounters*)take(sizeof(WaveCounters)); c->counters.n = 1;
        c->acc.p = (double*)take((size_t)RT_NOALLOC_MAX_PIXELS * 6 * sizeof(double)); c->acc.n = (size_t)RT_NOALLOC_MAX_PIXELS * 6;
        fb_sum = (float*)take((size_t)RT_NOALLOC_MAX_PIXELS * 3 * sizeof(float));
        fb_sumsq = (float*)take((size_t)RT_NOALLOC_MAX_PIXELS * 3 * sizeof(float));
        bytes = off;
        c->stage_timing = false;  // (stage timing creates events on demand)
    }
    ~NoallocLoan() {
        for (int i = 0; i < 2; ++i) { ctx->qa[i] = qa[i]; ctx->qb[i] = qb[i]; ctx->qc[i] = qc[i]; ctx->done[i] = done[i]; }
        ctx->hits = hits; ctx->counters = counters; ctx->acc = acc; ctx->stage_timing = stage_timing;
    }
    float* fb_sum;
    float* fb_sumsq;
    size_t bytes;
};
}  // namespace

int rt_render_noalloc(RtScene* scene, const RtCamera* cam, const RtRenderParams* prm, float* out_sum, float* out_sumsq, RtStats* stats) {
    if (!scene || !cam || !prm || !out_sum) return fail(RT_ERR_INVALID, "null argument");
    if (prm->width == 0 || prm->height == 0 || prm->spp == 0) return fail(RT_ERR_INVALID, "width/height/spp must be > 0");
    if (prm->row_stride == 0 || prm->row_begin >= prm->height) return fail(RT_ERR_INVALID, "bad row_begin/row_stride");
    const uint64_t rows = ((uint64_t)prm->height - prm->row_begin + prm->row_stride - 1) / prm->row_stride;
    const uint64_t pixels = (uint64_t)prm->width * rows;
    if (pixels > RT_NOALLOC_MAX_PIXELS)
        return fail(RT_ERR_INVALID, "rt_render_noalloc: %llu pixels exceed the static scratch maximum of %u (RT_NOALLOC_MAX_PIXELS); use rt_render",
                    (unsigned long long)pixels, (unsigned)RT_NOALLOC_MAX_PIXELS);
    RtContext* ctx = scene->ctx;
    RT_CUDA(cudaSetDevice(ctx->device));
    unsigned char* base = nullptr;
    if (int rc = noalloc_scratch(ctx->device, &base)) return rc;
    NoallocLoan loan(ctx, base);
    static_assert(sizeof(WaveCounters) <= 256, "counter block");
    if (loan.bytes > RT_NOALLOC_SCRATCH_BYTES) return fail(RT_ERR_INVALID, "static scratch layout exceeds RT_NOALLOC_SCRATCH_BYTES");
    RtRenderParams p = *prm;
    p.queue_capacity = RT_NOALLOC_QUEUE_RAYS;  // the queues of the static array, whatever the caller asked for
    p.flags &= (RT_RENDER_SAMPLE_MAJOR | RT_RENDER_EXTEND_PER_THREAD | RT_RENDER_EXTEND_PERSISTENT);  // no instrumentation: it creates events
    // one wavefront on the context's own stream (render_core; the lanes of rt_render own device buffers of their own)
    const int rc = render_core(ctx, scene, cam, nullptr, &p, loan.fb_sum, out_sumsq ? loan.fb_sumsq : nullptr, nullptr, stats);
    if (rc != RT_OK) return rc;
    RT_CUDA(cudaMemcpy(out_sum, loan.fb_sum, pixels * 3 * sizeof(float), cudaMemcpyDeviceToHost));
    if (out_sumsq) RT_CUDA(cudaMemcpy(out_sumsq, loan.fb_sumsq, pixels * 3 * sizeof(float), cudaMemcpyDeviceToHost));
    return RT_OK;
}

int rt_trace(RtScene* scene, const RtRayIn* rays, uint64_t n, uint64_t seed, RtHitOut* out) { return rt_trace_from(scene, rays, nullptr, n, seed, out); }

int rt_trace_from(RtScene* scene, const RtRayIn* rays, const uint32_t* self_order, uint64_t n, uint64_t seed, RtHitOut* out) {
    if (!scene || (n && (!rays || !out))) return fail(RT_ERR_INVALID, "null argument");
    if (n == 0) return RT_OK;
    RtContext* ctx = scene->ctx;
    RT_CUDA(cudaSetDevice(ctx->device));
    const size_t in_bytes = n * sizeof(RtRayIn), out_bytes = n * sizeof(RtHitOut), self_bytes = self_order ? n * sizeof(uint32_t) : 0;
    const size_t off_out = (in_bytes + 255) & ~(size_t)255, off_self = (off_out + out_bytes + 255) & ~(size_t)255;
    RT_CUDA(ctx->scratch.ensure(off_self + self_bytes + 256));
    RtRayIn* d_rays = reinterpret_cast<RtRayIn*>(ctx->scratch.p);
    RtHitOut* d_out = reinterpret_cast<RtHitOut*>(ctx->scratch.p + off_out);
    uint32_t* d_self = self_order ? reinterpret_cast<uint32_t*>(ctx->scratch.p + off_self) : nullptr;
    RT_CUDA(cudaMemcpyAsync(d_rays, rays, in_bytes, cudaMemcpyHostToDevice, ctx->stream));
    if (self_order) RT_CUDA(cudaMemcpyAsync(d_self, self_order, self_bytes, cudaMemcpyHostToDevice, ctx->stream));
    launch_trace(scene->launch, scene->d, d_rays, d_self, n, (uint32_t)seed, (uint32_t)(seed >> 32), d_out, ctx->stream);
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaMemcpyAsync(out, d_out, out_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    RT_CUDA(cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}

int rt_resolve_u8(RtContext* ctx, const float* sum_rgb, uint32_t width, uint32_t height, uint32_t spp, uint8_t* out) {
    if (!ctx || !sum_rgb || !out || !width || !height || !spp) return fail(RT_ERR_INVALID, "bad argument");
    RT_CUDA(cudaSetDevice(ctx->device));
    const size_t n = (size_t)width * height * 3;
    RT_CUDA(ctx->fb_sum.ensure(n));
    RT_CUDA(ctx->scratch.ensure(n));
    RT_CUDA(cudaMemcpyAsync(ctx->fb_sum.p, sum_rgb, n * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    launch_resolve_u8(ctx->num_sms, ctx->fb_sum.p, width, height, spp, ctx->scratch.p, ctx->stream);
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaMemcpyAsync(out, ctx->scratch.p, n, cudaMemcpyDeviceToHost, ctx->stream));
    RT_CUDA(cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}

}  // extern "C"

// ---- several GPUs of one process ---------------------------------------------------------------------------------
//
// The reference spawns one thread per core, thread i renders the rows y % n == i and the main thread reassembles the image
// (examples/book-one.rs:54-88).  Here: one host thread per DEVICE (each runs the complete wavefront of rt_render_device on its
// own context, with its own lane threads), scene replicated, rows y % G == g, and the shards go to device 0 with one grouped
// ncclSend / ncclRecv over NVLink; k_deinterleave puts row y back at index y.  NCCL is resolved with dlopen so that the library has
// no link-time dependency on it (a single-GPU user needs none, and a process that already holds an NCCL — PyTorch's — keeps it).

namespace {

struct NcclApi {
    void* handle = nullptr;
    decltype(&ncclCommInitAll) CommInitAll = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclSend) Send = nullptr;
    decltype(&ncclRecv) Recv = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    decltype(&ncclGetVersion) GetVersion = nullptr;
};
std::mutex g_nccl_mutex;
NcclApi g_nccl;

int load_nccl() {
    std::lock_guard<std::mutex> lock(g_nccl_mutex);
    if (g_nccl.handle) return RT_OK;
    void* h = nullptr;
    for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
        h = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) return fail(RT_ERR_NCCL, "cannot load libnccl.so.2: %s", dlerror());
    NcclApi api;
    api.handle = h;
#define RT_SYM(field, sym) api.field = reinterpret_cast<decltype(api.field)>(dlsym(h, #sym)); if (!api.field) return fail(RT_ERR_NCCL, "libnccl lacks %s", #sym)
    RT_SYM(CommInitAll, ncclCommInitAll);
    RT_SYM(CommDestroy, ncclCommDestroy);
    RT_SYM(GroupStart, ncclGroupStart);
    RT_SYM(GroupEnd, ncclGroupEnd);
    RT_SYM(Send, ncclSend);
    RT_SYM(Recv, ncclRecv);
    RT_SYM(GetErrorString, ncclGetErrorString);
    RT_SYM(GetVersion, ncclGetVersion);
#undef RT_SYM
    g_nccl = api;
    return RT_OK;
}

#define RT_NCCL(call)                                                                                               \
    do {                                                                                                            \
        ncclResult_t r_ = (call);                                                                                   \
        if (r_ != ncclSuccess) return fail(RT_ERR_NCCL, "%s failed: %s (%s:%d)", #call, g_nccl.GetErrorString(r_), __FILE__, __LINE__); \
    } while (0)

}  // namespace

struct RtMulti {
    std::vector<RtContext*> ctx;     // one per device, in the caller's order; ctx[0] is the root of the gather
    std::vector<ncclComm_t> comms;   // empty for a single device
    std::vector<DevBuf<float>> shard, shard_sq;  // each device's compact row shard [rows_padded][W][3]
    DevBuf<float> gathered, full;    // on device 0
    WorkerPool pool;                 // one host thread per device
};

struct RtMultiScene {
    RtMulti* multi = nullptr;
    std::vector<RtScene*> scenes;
};

extern "C" {

int rt_init_multi(int n_devices, const int* device_ids, const RtContextOptions* options, RtMulti** out) {
    if (!out) return fail(RT_ERR_INVALID, "out is null");
    *out = nullptr;
    if (n_devices < 1 || n_devices > 64) return fail(RT_ERR_INVALID, "n_devices %d out of range", n_devices);
    std::vector<int> ids(n_devices);
    for (int i = 0; i < n_devices; ++i) {
        ids[i] = device_ids ? device_ids[i] : i;
        for (int j = 0; j < i; ++j) if (ids[j] == ids[i]) return fail(RT_ERR_INVALID, "device %d listed twice", ids[i]);
    }
    RtMulti* m = new (std::nothrow) RtMulti();
    if (!m) return fail(RT_ERR_OOM, "out of host memory");
    for (int i = 0; i < n_devices; ++i) {
        RtContext* c = nullptr;
        const int rc = rt_init_ex(ids[i], options, &c);
        if (rc != RT_OK) { rt_shutdown_multi(m); return rc; }
        m->ctx.push_back(c);
    }
    m->shard.resize(n_devices);
    m->shard_sq.resize(n_devices);
    if (n_devices > 1) {
        const int rc = load_nccl();
        if (rc != RT_OK) { rt_shutdown_multi(m); return rc; }
        m->comms.assign(n_devices, nullptr);
        const ncclResult_t r = g_nccl.CommInitAll(m->comms.data(), n_devices, ids.data());
        if (r != ncclSuccess) {
            m->comms.clear();
            rt_shutdown_multi(m);
            return fail(RT_ERR_NCCL, "ncclCommInitAll failed: %s", g_nccl.GetErrorString(r));
        }
    }
    *out = m;
    return RT_OK;
}

void rt_shutdown_multi(RtMulti* m) {
    if (!m) return;
    for (size_t i = 0; i < m->comms.size(); ++i) {
        if (!m->comms[i]) continue;
        cudaSetDevice(m->ctx[i]->device);
        g_nccl.CommDestroy(m->comms[i]);
    }
    for (size_t i = 0; i < m->ctx.size(); ++i) {
        cudaSetDevice(m->ctx[i]->device);
        if (i < m->shard.size()) { m->shard[i].release(); m->shard_sq[i].release(); }
        if (i == 0) { m->gathered.release(); m->full.release(); }
        rt_shutdown(m->ctx[i]);
    }
    delete m;
}

int rt_multi_device_count(const RtMulti* m) { return m ? (int)m->ctx.size() : 0; }
RtContext* rt_multi_context(RtMulti* m, int index) { return (m && index >= 0 && index < (int)m->ctx.size()) ? m->ctx[index] : nullptr; }

int rt_multi_scene_create(RtMulti* m, const RtSceneDesc* desc, const RtSceneOptions* options, RtMultiScene** out) {
    if (!m || !out) return fail(RT_ERR_INVALID, "multi/out is null");
    *out = nullptr;
    RtMultiScene* ms = new (std::nothrow) RtMultiScene();
    if (!ms) return fail(RT_ERR_OOM, "out of host memory");
    ms->multi = m;
    const size_t G = m->ctx.size();
    ms->scenes.assign(G, nullptr);
    // the replicas are built concurrently (each device builds / uploads its own copy; the host SAH build is deterministic)
    std::vector<int> rcs(G, RT_OK);
    std::vector<std::string> errs(G);
    m->pool.run((uint32_t)G, [&](uint32_t g) {
        rcs[g] = rt_scene_create_ex(m->ctx[g], desc, options, &ms->scenes[g]);
        if (rcs[g] != RT_OK) errs[g] = g_error;
    });
    for (size_t g = 0; g < G; ++g)
        if (rcs[g] != RT_OK) {
            const int rc = rcs[g];
            const std::string msg = errs[g];
            rt_multi_scene_destroy(ms);
            return fail(rc, "device %d: %s", m->ctx[g]->device, msg.c_str());
        }
    *out = ms;
    return RT_OK;
}

void rt_multi_scene_destroy(RtMultiScene* ms) {
    if (!ms) return;
    for (RtScene* s : ms->scenes) rt_scene_destroy(s);
    delete ms;
}

int rt_render_multi(RtMultiScene* ms, const RtCamera* cam, const RtRenderParams* prm, float* out_sum, float* out_sumsq, RtStats* stats) {
    if (!ms || !cam || !prm || !out_sum) return fail(RT_ERR_INVALID, "null argument");
    if (prm->row_begin != 0 || prm->row_stride != 1) return fail(RT_ERR_INVALID, "rt_render_multi shards the rows itself: row_begin / row_stride must be 0 / 1");
    if (prm->width == 0 || prm->height == 0 || prm->spp == 0) return fail(RT_ERR_INVALID, "width/height/spp must be > 0");
    RtMulti* m = ms->multi;
    const uint32_t G = (uint32_t)m->ctx.size();
    const uint32_t W = prm->width, H = prm->height;
    const uint32_t Gact = std::min(G, H);  // devices that own at least one row
    const uint32_t rows_padded = (H + Gact - 1) / Gact;
    const size_t shard = (size_t)rows_padded * W * 3, full = (size_t)H * W * 3;
    RtContext* root = m->ctx[0];
    RT_CUDA(cudaSetDevice(root->device));
    const int n_img = out_sumsq ? 2 : 1;
    RT_CUDA(m->full.ensure(full * n_img));
    if (Gact > 1) RT_CUDA(m->gathered.ensure(shard * Gact * n_img));
    RT_CUDA(cudaEventRecord(root->ev0, root->stream));

    std::vector<RtStats> st(Gact);
    std::vector<int> rcs(Gact, RT_OK);
    std::vector<std::string> errs(Gact);
    {
        m->pool.run(Gact, [&](uint32_t g) {
            RtContext* c = m->ctx[g];
            auto cuda_fail = [&](cudaError_t e) { rcs[g] = RT_ERR_CUDA; errs[g] = cudaGetErrorString(e); };
            cudaError_t e = cudaSetDevice(c->device);
            if (e != cudaSuccess) return cuda_fail(e);
            // a single device renders straight into the full image; otherwise into this device's shard
            float* d_sum = nullptr;
            float* d_sq = nullptr;
            if (Gact == 1) { d_sum = m->full.p; d_sq = out_sumsq ? m->full.p + full : nullptr; }
            else {
                if ((e = m->shard[g].ensure(shard)) != cudaSuccess) return cuda_fail(e);
                if (out_sumsq && (e = m->shard_sq[g].ensure(shard)) != cudaSuccess) return cuda_fail(e);
                d_sum = m->shard[g].p; d_sq = out_sumsq ? m->shard_sq[g].p : nullptr;
            }
            RtRenderParams p = *prm;
            p.row_begin = g; p.row_stride = Gact;
            rcs[g] = rt_render_device(ms->scenes[g], cam, &p, d_sum, d_sq, nullptr, &st[g]);
            if (rcs[g] != RT_OK) errs[g] = g_error;
        });
    }
    for (uint32_t g = 0; g < Gact; ++g)
        if (rcs[g] != RT_OK) return fail(rcs[g], "device %d: %s", m->ctx[g]->device, errs[g].c_str());

    RT_CUDA(cudaSetDevice(root->device));
    if (Gact > 1) {
        // the ONE collective: every shard to device 0 (rt_render_device returned after its stream drained, so the shards are complete)
        for (int img = 0; img < n_img; ++img) {
            float* dst = m->gathered.p + (size_t)img * shard * Gact;
            RT_CUDA(cudaMemcpyAsync(dst, img ? m->shard_sq[0].p : m->shard[0].p, shard * sizeof(float), cudaMemcpyDeviceToDevice, root->stream));
            RT_NCCL(g_nccl.GroupStart());
            for (uint32_t g = 1; g < Gact; ++g) {
                RT_NCCL(g_nccl.Recv(dst + shard * g, shard, ncclFloat, (int)g, m->comms[0], root->stream));
                RT_NCCL(g_nccl.Send(img ? m->shard_sq[g].p : m->shard[g].p, shard, ncclFloat, 0, m->comms[g], m->ctx[g]->stream));
            }
            RT_NCCL(g_nccl.GroupEnd());
            launch_deinterleave(root->num_sms, dst, m->full.p + (size_t)img * full, W, H, Gact, root->stream);
        }
    }
    RT_CUDA(cudaEventRecord(root->ev1, root->stream));
    RT_CUDA(cudaMemcpyAsync(out_sum, m->full.p, full * sizeof(float), cudaMemcpyDeviceToHost, root->stream));
    if (out_sumsq) RT_CUDA(cudaMemcpyAsync(out_sumsq, m->full.p + full, full * sizeof(float), cudaMemcpyDeviceToHost, root->stream));
    RT_CUDA(cudaStreamSynchronize(root->stream));
    for (uint32_t g = 1; g < Gact; ++g) {  // the senders' streams (their buffers are reused by the next render)
        RT_CUDA(cudaSetDevice(m->ctx[g]->device));
        RT_CUDA(cudaStreamSynchronize(m->ctx[g]->stream));
    }
    RT_CUDA(cudaSetDevice(root->device));
    RT_CUDA(cudaGetLastError());
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        float ms_total = 0;
        RT_CUDA(cudaEventElapsedTime(&ms_total, root->ev0, root->ev1));
        for (const RtStats& s : st) {
            stats->paths += s.paths; stats->rays += s.rays; stats->launches += s.launches;
            stats->iterations = std::max(stats->iterations, s.iterations);
        }
        if (Gact > 1) stats->launches += (uint64_t)n_img;
        stats->ms_total = ms_total;
        stats->bvh_nodes = ms->scenes[0]->d.n_nodes; stats->bvh_always = ms->scenes[0]->d.n_always;
    }
    return RT_OK;
}

}  // extern "C"
