// C++ twin of the Rust module `src/gpu.rs`: the two calls that replace the example
// drivers' pixel loop and per-ray `world.hit()`:
//
//   gpu::render_image(world, camera, width, height, spp, maxDepth, seed) -> Vec<Vec3>
//       replaces `for y / for x / for _ in 0..spp { color(&camera.ray(u, v), &world, depth) }`
//       (examples/book-one.rs:62-78); returns per-pixel MEANS, row-major with y up like
//       `buffer[y][x]` (book-one.rs:53,86).
//   gpu::trace_rays(world, rays) -> Vec<Option<TraceHit>>
//       batched `world.hit(&ray)` (src/optimize.rs:469-498).
//
// Both go through the C ABI (include/rt_b200.h); errors become exceptions carrying rt_last_error().
#pragma once
#include <optional>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/rt_b200.h"
#include "flatten.hpp"

namespace rtb {
namespace host {
namespace gpu {

struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string& what) : std::runtime_error(what), code(c) {}
};
inline void check(int rc) {
    if (rc != RT_OK) throw Error(rc, rt_last_error());
}

class Device {
public:
    explicit Device(int cudaDevice = 0) { check(rt_init(cudaDevice, &ctx_)); }
    ~Device() { rt_shutdown(ctx_); }
    Device(const Device&) = delete;
    Device& operator=(const Device&) = delete;
    RtContext* handle() const { return ctx_; }

private:
    RtContext* ctx_ = nullptr;
};

// A flattened, uploaded world (device BVH + primitive / material / texture tables).
class World {
public:
    World(Device& dev, const Bound& root) {
        const FlatScene flat = flattenWorld(root);
        const RtSceneDesc d = flat.desc();
        check(rt_scene_create(dev.handle(), &d, &scene_));
    }
    ~World() { rt_scene_destroy(scene_); }
    World(const World&) = delete;
    World& operator=(const World&) = delete;
    RtScene* handle() const { return scene_; }

private:
    RtScene* scene_ = nullptr;
};

struct Ray {  // src/ray.rs:6-33
    Vec3 origin, direction;
};
struct TraceHit {
    uint32_t prim;  // DFS rank of the leaf in the scene graph
    double t;
    Vec3 intersection, normal;
    double u, v;
    int material;
};

struct RenderOptions {
    uint64_t seed = 1;
    uint32_t rowBegin = 0, rowStride = 1;  // shard: rows y = rowBegin + k * rowStride
    std::vector<float>* sumsq = nullptr;   // optional per-pixel sum of squared path radiance
    RtStats* stats = nullptr;
};

inline std::vector<Vec3> render_image(World& world, const PerspectiveCamera& camera, uint32_t width, uint32_t height, uint32_t spp, uint32_t maxDepth,
                                      const RenderOptions& opt = RenderOptions()) {
    const RtCamera cam = flattenCamera(camera);
    RtRenderParams prm{};
    prm.width = width; prm.height = height; prm.spp = spp; prm.max_depth = maxDepth; prm.seed = opt.seed;
    prm.row_begin = opt.rowBegin; prm.row_stride = opt.rowStride;
    const uint32_t rows = (height - opt.rowBegin + opt.rowStride - 1) / opt.rowStride;
    std::vector<float> sum((size_t)rows * width * 3);
    if (opt.sumsq) opt.sumsq->assign(sum.size(), 0.0f);
    check(rt_render(world.handle(), &cam, &prm, sum.data(), opt.sumsq ? opt.sumsq->data() : nullptr, opt.stats));
    std::vector<Vec3> out((size_t)rows * width);
    for (size_t i = 0; i < out.size(); ++i) out[i] = Vec3(sum[3 * i] / spp, sum[3 * i + 1] / spp, sum[3 * i + 2] / spp);  // pixel /= spp (book-one.rs:76)
    return out;
}

inline std::vector<std::optional<TraceHit>> trace_rays(World& world, const std::vector<Ray>& rays, uint64_t seed = 1) {
    std::vector<RtRayIn> in(rays.size());
    for (size_t i = 0; i < rays.size(); ++i) {
        const Ray& r = rays[i];
        in[i] = RtRayIn{{(float)r.origin.x, (float)r.origin.y, (float)r.origin.z}, {(float)r.direction.x, (float)r.direction.y, (float)r.direction.z}};
    }
    std::vector<RtHitOut> out(rays.size());
    check(rt_trace(world.handle(), in.data(), in.size(), seed, out.data()));
    std::vector<std::optional<TraceHit>> res(rays.size());
    for (size_t i = 0; i < rays.size(); ++i) {
        const RtHitOut& h = out[i];
        if (!h.hit) continue;
        res[i] = TraceHit{(uint32_t)h.prim, h.t, Vec3(h.p[0], h.p[1], h.p[2]), Vec3(h.n[0], h.n[1], h.n[2]), h.uv[0], h.uv[1], h.material};
    }
    return res;
}

// mean -> sqrt -> * 255 -> min(255) -> truncation, rows top to bottom (examples/book-one.rs:90-100), on the pixels render_image
// returned (means, y up).  The device resolves in f64 like the reference, so the bytes equal the reference's arithmetic on the
// same f32-rounded means.
inline std::vector<uint8_t> resolve_u8(Device& dev, const std::vector<Vec3>& pixels, uint32_t width, uint32_t height) {
    std::vector<float> sums(pixels.size() * 3);
    for (size_t i = 0; i < pixels.size(); ++i) { sums[3 * i] = (float)pixels[i].x; sums[3 * i + 1] = (float)pixels[i].y; sums[3 * i + 2] = (float)pixels[i].z; }
    std::vector<uint8_t> out(sums.size());
    check(rt_resolve_u8(dev.handle(), sums.data(), width, height, 1, out.data()));
    return out;
}

// render::color(&ray, world, maxDepth) (src/render.rs:5) for a batch of rays: the mean of `spp` evaluations per ray.
inline std::vector<Vec3> color(World& world, const std::vector<Ray>& rays, uint32_t maxDepth, uint32_t spp = 1, uint64_t seed = 1) {
    std::vector<RtRayIn> in(rays.size());
    for (size_t i = 0; i < rays.size(); ++i) {
        const Ray& r = rays[i];
        in[i] = RtRayIn{{(float)r.origin.x, (float)r.origin.y, (float)r.origin.z}, {(float)r.direction.x, (float)r.direction.y, (float)r.direction.z}};
    }
    std::vector<float> sum(rays.size() * 3);
    check(rt_color_rays(world.handle(), in.data(), in.size(), spp, maxDepth, seed, sum.data(), nullptr, nullptr));
    std::vector<Vec3> out(rays.size());
    for (size_t i = 0; i < out.size(); ++i) out[i] = Vec3(sum[3 * i] / spp, sum[3 * i + 1] / spp, sum[3 * i + 2] / spp);
    return out;
}

}  // namespace gpu
}  // namespace host
}  // namespace rtb
