// examples/book_one.cpp — the reference's examples/book-one.rs through the C++ host mirror: same scene construction calls
// (Sprite::builder, BoundingVolumeHierarchyNode::make), same camera (book-one.rs:35-47), and main()'s thread pool + pixel loop
// + mpsc funnel (book-one.rs:52-88) replaced by ONE call, gpu::render_image; the println! loop (book-one.rs:90-100) becomes
// gpu::resolve_u8 + the PPM P3 encoder — "P3\n<w> <h>\n255\n", then one "r g b" line per pixel, rows top to bottom.
//
//   book_one [width height spp depth seed [out.ppm]]      defaults: 1600 800 100 100 1 (the example's own numbers), stdout
//
// This is the C++ twin of rust/examples/book-one-gpu.rs (no Rust toolchain in this image).  No CPU fallback: without an sm_100
// device rt_init fails and this program exits 1 with the library's message.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../host/api_bundle.hpp"
#include "../host/gpu.hpp"
#include "../host/image_io.hpp"
#include "scenes.hpp"

using namespace rtb::host;

int main(int argc, char** argv) {
    const uint32_t width = argc > 1 ? (uint32_t)std::atoi(argv[1]) : 1600u, height = argc > 2 ? (uint32_t)std::atoi(argv[2]) : 800u;
    const uint32_t spp = argc > 3 ? (uint32_t)std::atoi(argv[3]) : 100u, depth = argc > 4 ? (uint32_t)std::atoi(argv[4]) : 100u;
    const uint64_t seed = argc > 5 ? std::strtoull(argv[5], nullptr, 10) : 1ull;
    const char* out_path = argc > 6 ? argv[6] : nullptr;
    if (!width || !height || !spp) { std::fprintf(stderr, "usage: book_one [width height spp depth seed [out.ppm]]\n"); return 2; }
    try {
        // let world = BoundingVolumeHierarchyNode::new(randomScene()).unwrap();      (book-one.rs:32)
        rtscenes::SceneRng generator(seed);
        BuildAxisStream axes(seed ^ 0xA11CE5EEDull);  // thread_rng() of optimize.rs:374, seeded
        currentAxisStream() = &axes;
        auto world = BoundingVolumeHierarchyNode::make(rtscenes::Scenes<HostApi>::bookOne(generator));
        currentAxisStream() = nullptr;
        if (!world) { std::fprintf(stderr, "empty scene\n"); return 1; }
        // let camera = PerspectiveCamera::new(eye, center, up, 20 deg, width / height, 10.0, 0.05);   (book-one.rs:35-47)
        const rtscenes::CameraSpec c = rtscenes::Scenes<HostApi>::bookOneCamera();
        const PerspectiveCamera camera({c.eye[0], c.eye[1], c.eye[2]}, {c.center[0], c.center[1], c.center[2]}, {c.up[0], c.up[1], c.up[2]},
                                       rtscenes::to_radians(c.fov_degrees), (double)width / (double)height, c.focusDistance, c.lensRadius);
        gpu::Device device(0);
        gpu::World device_world(device, *world);
        RtStats stats{};
        gpu::RenderOptions opt;
        opt.seed = seed;
        opt.stats = &stats;
        const std::vector<Vec3> pixels = gpu::render_image(device_world, camera, width, height, spp, depth, opt);  // buffer[y][x], y up
        std::fprintf(stderr, "%llu paths, %llu rays, %.2f ms on the device (%.1f Mpaths/s)\n", (unsigned long long)stats.paths,
                     (unsigned long long)stats.rays, stats.ms_total, stats.paths / (stats.ms_total * 1e3));
        const std::vector<uint8_t> bytes = gpu::resolve_u8(device, pixels, width, height);  // sqrt, * 255, min, truncation; rows top -> bottom
        const std::string ppm = rth::encode_ppm_p3(bytes.data(), width, height);
        FILE* f = out_path ? std::fopen(out_path, "wb") : stdout;
        if (!f) { std::perror(out_path); return 1; }
        std::fwrite(ppm.data(), 1, ppm.size(), f);
        if (out_path) std::fclose(f);
        return 0;
    } catch (const std::exception& e) {
        std::fprintf(stderr, "book_one: %s\n", e.what());
        return 1;
    }
}
