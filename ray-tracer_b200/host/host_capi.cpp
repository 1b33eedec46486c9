// extern "C" surface of the host mirror (include/rt_b200_host.h): example scenes built through the
// mirrored reference API, flattened, and uploaded through the C ABI.
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "../../include/rt_b200_host.h"
#include "../examples/scenes.hpp"
#include "api_bundle.hpp"
#include "flatten.hpp"
#include "gpu.hpp"
#include "image_io.hpp"

namespace {

thread_local std::string g_host_error;

}  // namespace

struct RthFlatScene {
    rtb::host::FlatScene flat;
    rtscenes::CameraSpec cam;
};

extern "C" {

int rth_example_scene(const char* name, uint64_t seed, RthFlatScene** out) {
    if (!name || !out) return RT_ERR_INVALID;
    *out = nullptr;
    try {
        auto res = std::make_unique<RthFlatScene>();
        std::vector<rtb::host::BoundPtr> list;
        rtb::host::BuildAxisStream axes(seed ^ 0xA11CE5EEDull);
        rtb::host::currentAxisStream() = &axes;
        bool rootIsList = false;
        const bool ok = rtscenes::Scenes<rtb::host::HostApi>::byName(name, seed, &list, &res->cam, &rootIsList);
        rtb::host::BoundPtr world;
        if (ok && rootIsList) {  // the world is the Vec itself (impl Hit / Bound for Vec, src/geometry.rs:76-116)
            auto l = std::make_shared<rtb::host::HitList>();
            l->items = list;
            world = l;
        } else if (ok) world = rtb::host::BoundingVolumeHierarchyNode::make(list);  // examples/book-one.rs:32
        rtb::host::currentAxisStream() = nullptr;
        if (!ok || !world) return RT_ERR_INVALID;
        res->flat = rtb::host::flattenWorld(*world);
        *out = res.release();
        return RT_OK;
    } catch (const rtb::host::FlattenError&) {
        rtb::host::currentAxisStream() = nullptr;
        return RT_ERR_UNSUPPORTED;
    } catch (...) {
        rtb::host::currentAxisStream() = nullptr;
        return RT_ERR_INVALID;
    }
}

void rth_flat_destroy(RthFlatScene* flat) { delete flat; }

int rth_flat_camera_args(const RthFlatScene* flat, double aspect, RthCameraArgs* out) {
    if (!flat || !out) return RT_ERR_INVALID;
    const auto& c = flat->cam;
    for (int i = 0; i < 3; ++i) { out->eye[i] = c.eye[i]; out->center[i] = c.center[i]; out->up[i] = c.up[i]; }
    out->fov = rtscenes::to_radians(c.fov_degrees);
    out->aspect = aspect;
    out->focus_distance = c.focusDistance;
    out->lens_radius = c.lensRadius;
    return RT_OK;
}

int rth_flat_desc(const RthFlatScene* flat, RtSceneDesc* out) {
    if (!flat || !out) return RT_ERR_INVALID;
    *out = flat->flat.desc();
    return RT_OK;
}

int rth_scene_upload(RtContext* ctx, const RthFlatScene* flat, RtScene** out) {
    if (!flat) return RT_ERR_INVALID;
    const RtSceneDesc d = flat->flat.desc();
    return rt_scene_create(ctx, &d, out);
}

int rth_camera(const RthCameraArgs* a, RtCamera* out) {
    if (!a || !out) return RT_ERR_INVALID;
    const rtb::host::PerspectiveCamera cam({a->eye[0], a->eye[1], a->eye[2]}, {a->center[0], a->center[1], a->center[2]}, {a->up[0], a->up[1], a->up[2]},
                                           a->fov, a->aspect, a->focus_distance, a->lens_radius);
    *out = rtb::host::flattenCamera(cam);
    return RT_OK;
}

int rth_decode_image(const uint8_t* bytes, uint64_t n, uint32_t* width, uint32_t* height, uint8_t* out_rgb, uint64_t capacity, uint64_t* needed) {
    if (!bytes || !width || !height || !needed) return RT_ERR_INVALID;
    std::vector<uint8_t> rgb;
    const int rc = rth::decode_image(bytes, (size_t)n, *width, *height, rgb);
    if (rc == -2) return RT_ERR_UNSUPPORTED;
    if (rc != 0) return RT_ERR_INVALID;
    *needed = rgb.size();
    if (!out_rgb) return RT_OK;
    if (capacity < rgb.size()) return RT_ERR_INVALID;
    std::memcpy(out_rgb, rgb.data(), rgb.size());
    return RT_OK;
}

int rth_set_earth_image(const uint8_t* rgb_top_down, uint32_t width, uint32_t height) {
    rtscenes::EarthImage& e = rtscenes::earthOverride();
    if (!rgb_top_down) { e = rtscenes::EarthImage(); return RT_OK; }
    if (width == 0 || height == 0) return RT_ERR_INVALID;
    e.width = width; e.height = height;
    e.rgb = std::make_shared<std::vector<uint8_t>>(rgb_top_down, rgb_top_down + (size_t)width * height * 3);
    return RT_OK;
}

int rth_encode_image(int format, const uint8_t* rgb, uint32_t width, uint32_t height, uint8_t* out, uint64_t capacity, uint64_t* needed) {
    if (!rgb || !needed || width == 0 || height == 0) return RT_ERR_INVALID;
    std::vector<uint8_t> bytes;
    if (format == RTH_IMAGE_PPM_P3) {
        const std::string text = rth::encode_ppm_p3(rgb, width, height);
        bytes.assign(text.begin(), text.end());
    } else if (format == RTH_IMAGE_PNG_RGBA8) {
        bytes = rth::encode_png_rgba8(rgb, width, height);
    } else {
        return RT_ERR_UNSUPPORTED;
    }
    *needed = bytes.size();
    if (!out) return RT_OK;
    if (capacity < bytes.size()) return RT_ERR_INVALID;
    std::memcpy(out, bytes.data(), bytes.size());
    return RT_OK;
}

}  // extern "C"
