/* rt_b200_host.h — C entry points of the host-side scene-description mirror (ray-tracer_b200/host/),
 * exported by the same shared library as rt_b200.h.  They let a non-C++ caller (the ctypes harness in
 * ray-tracer_b200/__init__.py, the tests, bench.py) build the reference's example scenes through the
 * mirrored API (Sprite::builder, BoundingVolumeHierarchyNode::new, Cube::new, ConstantMedium::new,
 * PerspectiveCamera::new — examples/book-one.rs:103-205, examples/cornell-box.rs:31-150,
 * examples/main.rs:156-330), flatten them and hand them to rt_scene_create.  None of these touches
 * the GPU except rth_scene_upload. */
#ifndef RT_B200_HOST_H
#define RT_B200_HOST_H

#include "rt_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct RthFlatScene RthFlatScene;

/* PerspectiveCamera::new arguments (src/camera.rs:25-33); fov in radians. */
typedef struct RthCameraArgs {
    double eye[3], center[3], up[3], fov, aspect, focus_distance, lens_radius;
} RthCameraArgs;

/* Build + flatten an example scene: "book-one", "cornell-box", "book-two", "stress:<count>".
 * Scene randomness (sphere positions, materials, BVH split axes) comes from `seed`. */
int rth_example_scene(const char* name, uint64_t seed, RthFlatScene** out);
void rth_flat_destroy(RthFlatScene* flat);
/* The example's camera (examples/book-one.rs:35-47 etc.) for a given aspect ratio. */
int rth_flat_camera_args(const RthFlatScene* flat, double aspect, RthCameraArgs* out);
/* Flat tables (pointers stay valid until rth_flat_destroy). */
int rth_flat_desc(const RthFlatScene* flat, RtSceneDesc* out);
/* rt_scene_create on the flattened tables. */
int rth_scene_upload(RtContext* ctx, const RthFlatScene* flat, RtScene** out);
/* PerspectiveCamera::new -> device camera constants (f64 math of src/camera.rs:34-45, rounded to f32). */
int rth_camera(const RthCameraArgs* args, RtCamera* out);


/* Encoders of the example drivers' output formats, fed with the top-down 8-bit image of
 * rt_resolve_u8.  RTH_IMAGE_PPM_P3: "P3\n<w> <h>\n255\n" + one "r g b\n" line per pixel, rows top to bottom
 * (examples/book-one.rs:28-30,90-100).  RTH_IMAGE_PNG_RGBA8: 8-bit RGBA PNG, alpha 255, row 0 = top
 * (examples/main.rs:105-135).  Two-call pattern: *needed receives the encoded size; bytes are written when
 * out != NULL and capacity >= *needed (else RT_ERR_INVALID with *needed set). */
typedef enum RthImageFormat { RTH_IMAGE_PPM_P3 = 0, RTH_IMAGE_PNG_RGBA8 = 1 } RthImageFormat;
int rth_encode_image(int format, const uint8_t* rgb_top_down, uint32_t width, uint32_t height, uint8_t* out, uint64_t capacity,
                     uint64_t* needed);


/* Texture ingestion (examples/main.rs:266 `image::open("./earthmap.jpg")`): decode a PNG (8-bit, colour types
 * 0/2/3/4/6, non-interlaced) or PPM (P3/P6, maxval <= 255) byte stream to top-down 8-bit RGB — what `as_rgb8()` gives the texture
 * closure of main.rs:267-280.  Same two-call pattern as rth_encode_image.  RT_ERR_UNSUPPORTED: 16-bit / interlaced / JPEG-free build. */
int rth_decode_image(const uint8_t* bytes, uint64_t n, uint32_t* width, uint32_t* height, uint8_t* out_rgb_top_down, uint64_t capacity,
                     uint64_t* needed);
/* Install (or with NULL remove) the image that the "book-two" example maps onto its earth sphere instead of the synthetic
 * stand-in; takes effect for scenes built afterwards (rth_example_scene). */
int rth_set_earth_image(const uint8_t* rgb_top_down, uint32_t width, uint32_t height);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_HOST_H */
