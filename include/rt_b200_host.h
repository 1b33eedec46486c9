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

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_HOST_H */
