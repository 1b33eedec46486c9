//! `src/gpu.rs` — the B200 (sm_100a) wavefront path-tracing backend behind aiifabbf/ray-tracer's own API.
//!
//! What stays as it is: `Sprite::builder`, the `Hit` / `Bound` traits, `BoundingVolumeHierarchyNode::new`,
//! `PerspectiveCamera::new`, `render::color`.  What this module adds:
//!
//! * `Flatten` — a supertrait of `Hit` (src/ray.rs:85-87 becomes `pub trait Hit: Send + Sync + Flatten`): the hook that walks a scene
//!   graph of `dyn Hit` / `dyn Bound<AxisAlignedBoundingBox>` trait objects (which cannot be down-cast) in the reference's
//!   left-first DFS order and emits flat device primitives with every enclosing `Sprite` / `TransformedGeometry` transform
//!   pre-composed (src/sprite.rs:87-139, src/geometry.rs:210-246).  Implemented here for every `Hit` type of the crate.
//! * `Describe` / `DescribeTexture` — the same for `Material` (src/material.rs:13-21) and `Texture` (src/material.rs:196-198).
//! * `Gpu` (context), `World` (an uploaded scene), `render_image` — replaces the `for y / for x / for _ in 0..spp` nest of the
//!   example drivers (examples/book-one.rs:52-88) — `trace_rays` (batched `world.hit(&ray)`, src/optimize.rs:469-498) and
//!   `color` (batched `render::color`, src/render.rs:5-29); `MultiGpu` / `MultiWorld` for several GPUs of one process.
//!
//! The `extern "C"` block below is the C ABI of `include/rt_b200.h` / `include/rt_b200_host.h`, name for name and argument for
//! argument (tests/test_rust_binding.py in the backend's repository diffs the two).  There is no CPU fallback: without a
//! CUDA device `Gpu::new` returns `Err`.
//!
//! Patch to the rest of the crate (rust/reference.patch): `pub mod gpu;` in src/lib.rs; the supertraits on `Hit`, `Material`,
//! `Texture`; five getters for fields this module must read but that are private to their modules (`DiffuseLight::emission`,
//! `Isotropic::albedo`, `CheckerTexture::black` / `white`, `Vec<..>` needs none).  Camera constants are recomputed from the
//! public getters (src/camera.rs:61-87) because `lowerLeft` / `horizontal` / `vertical` have none (src/camera.rs:18-20).
//!
//! No Rust toolchain exists in the image the backend was developed in: this file is written against the reference's sources
//! (file:line cited per item) and against the header, and has not been compiled by its authors.
#![allow(non_snake_case, non_camel_case_types, dead_code)]

use crate::camera::PerspectiveCamera;
use crate::geometry::{Rectangle, Sphere, TransformedGeometry};
use crate::mat4::Mat4;
use crate::material::{CheckerTexture, Dielectric, DiffuseLight, ImageTexture, Isotropic, Lambertian, Material, Metal, SolidColor, Texture};
use crate::optimize::{AxisAlignedBoundingBox, Bound, BoundingVolumeHierarchyNode};
use crate::ray::{Hit, Ray};
use crate::sprite::Sprite;
use crate::vec3::Vec3;
use crate::volume::ConstantMedium;

use std::collections::HashMap;
use std::ffi::CStr;
use std::os::raw::{c_char, c_int, c_void};
use std::sync::Arc;

// ---------------------------------------------------------------------------------------------------------------------
// C ABI (include/rt_b200.h, ABI version 2).  Every struct is #[repr(C)] plain-old-data.
// ---------------------------------------------------------------------------------------------------------------------
pub mod ffi {
    use super::*;

    pub const RT_ABI_VERSION: u32 = 2;
    pub const RT_OK: c_int = 0;
    pub const RT_ERR_INVALID: c_int = -1;
    pub const RT_ERR_UNSUPPORTED: c_int = -2;
    pub const RT_ERR_CUDA: c_int = -3;
    pub const RT_ERR_NO_DEVICE: c_int = -4;
    pub const RT_ERR_OOM: c_int = -5;
    pub const RT_ERR_NCCL: c_int = -6;

    pub const RT_PRIM_SPHERE: u32 = 0;
    pub const RT_PRIM_RECT: u32 = 1;
    pub const RT_PRIM_MEDIUM_SPHERE: u32 = 2;
    pub const RT_PRIM_MEDIUM: u32 = 3;

    pub const RT_MAT_LAMBERTIAN: u32 = 0;
    pub const RT_MAT_METAL: u32 = 1;
    pub const RT_MAT_DIELECTRIC: u32 = 2;
    pub const RT_MAT_DIFFUSE_LIGHT: u32 = 3;
    pub const RT_MAT_ISOTROPIC: u32 = 4;

    pub const RT_TEX_SOLID: u32 = 0;
    pub const RT_TEX_CHECKER: u32 = 1;
    pub const RT_TEX_IMAGE: u32 = 2;

    pub const RT_NO_PRIM: u32 = 0xFFFF_FFFF;

    // RtRenderFlags / RtContextFlags / RtSceneFlags (all-zero = defaults; the flags exist for A/B measurements and tests)
    pub const RT_RENDER_STAGE_TIMING: u32 = 1;
    pub const RT_RENDER_COUNT_WORK: u32 = 2;
    pub const RT_RENDER_SAMPLE_MAJOR: u32 = 4;
    pub const RT_RENDER_EXTEND_PER_THREAD: u32 = 8;
    pub const RT_RENDER_EXTEND_PERSISTENT: u32 = 16;
    pub const RT_RENDER_TIME_EVERY_ITERATION: u32 = 32;
    pub const RT_CTX_NO_ACCUMULATE_OVERLAP: u32 = 1;
    pub const RT_CTX_STAGE_TIMING: u32 = 2;
    pub const RT_CTX_SHADE_ONE_RAY: u32 = 4;
    pub const RT_CTX_NO_FUSED_RECTS: u32 = 8;
    pub const RT_SCENE_NO_FRAME_RECTS: u32 = 1;
    pub const RT_SCENE_NO_STAGED_PRIMS: u32 = 2;
    pub const RT_SCENE_NO_SMALL_LIST: u32 = 4;
    pub const RT_SCENE_NO_FRAME_BOXES: u32 = 8;
    pub const RT_SCENE_NO_BOX_PRIMS: u32 = 16;
    pub const RT_SCENE_NO_WHOLE_TREE: u32 = 32;
    // rt_render_noalloc: limits of the static device scratch
    pub const RT_NOALLOC_MAX_PIXELS: u32 = 65536;
    pub const RT_NOALLOC_QUEUE_RAYS: u32 = 32768;
    pub const RT_NOALLOC_SCRATCH_BYTES: u32 = 9 << 20;

    #[repr(C)]
    #[derive(Clone, Copy, Debug, Default)]
    pub struct RtPrim {
        pub kind: u32,
        pub material: i32,
        pub order: u32,
        pub aux: u32,
        pub param: [f32; 4],
        pub local_from_world: [f32; 12],
        pub world_from_local: [f32; 12],
    } // 128 B

    #[repr(C)]
    #[derive(Clone, Copy, Debug, Default)]
    pub struct RtMaterial {
        pub kind: u32,
        pub texture: i32,
        pub fuzz: f32,
        pub ior: f32,
    } // 16 B

    #[repr(C)]
    #[derive(Clone, Copy, Debug, Default)]
    pub struct RtTexture {
        pub kind: u32,
        pub a: i32,
        pub b: i32,
        pub width: u32,
        pub height: u32,
        pub reserved: u32,
        pub texel_offset: u64,
        pub rgb: [f32; 4],
    } // 48 B

    #[repr(C)]
    #[derive(Clone, Copy, Debug)]
    pub struct RtSceneDesc {
        pub abi_version: u32,
        pub n_prims: u32,
        pub n_materials: u32,
        pub n_textures: u32,
        pub prims: *const RtPrim,
        pub materials: *const RtMaterial,
        pub textures: *const RtTexture,
        pub texels: *const f32,
        pub n_texels: u64,
        pub boundary_prims: *const RtPrim,
        pub n_boundary_prims: u32,
        pub reserved: u32,
    }

    #[repr(C)]
    #[derive(Clone, Copy, Debug, Default)]
    pub struct RtCamera {
        pub eye: [f32; 3],
        pub lower_left: [f32; 3],
        pub horizontal: [f32; 3],
        pub vertical: [f32; 3],
        pub lens_radius: f32,
        pub reserved: [f32; 3],
    } // 64 B

    #[repr(C)]
    #[derive(Clone, Copy, Debug, Default)]
    pub struct RtRenderParams {
        pub width: u32,
        pub height: u32,
        pub spp: u32,
        pub max_depth: u32,
        pub seed: u64,
        pub row_begin: u32,
        pub row_stride: u32,
        pub queue_capacity: u32,
        pub flags: u32,
    }

    #[repr(C)]
    #[derive(Clone, Copy, Debug, Default)]
    pub struct RtStats {
        pub paths: u64,
        pub rays: u64,
        pub iterations: u64,
        pub launches: u64,
        pub ms_total: f32,
        pub ms_generate: f32,
        pub ms_extend: f32,
        pub ms_shade: f32,
        pub ms_accumulate: f32,
        pub bvh_nodes: u32,
        pub bvh_always: u32,
        pub box_tests: u64,
        pub prim_tests: u64,
        pub timed_iterations: u64,
    }

    #[repr(C)]
    #[derive(Clone, Copy, Debug, Default)]
    pub struct RtRayIn {
        pub origin: [f32; 3],
        pub direction: [f32; 3],
    } // 24 B

    #[repr(C)]
    #[derive(Clone, Copy, Debug, Default)]
    pub struct RtHitOut {
        pub hit: i32,
        pub prim: i32,
        pub t: f32,
        pub p: [f32; 3],
        pub n: [f32; 3],
        pub uv: [f32; 2],
        pub material: i32,
    } // 48 B

    #[repr(C)]
    #[derive(Clone, Copy, Debug, Default)]
    pub struct RtContextOptions {
        pub lanes: u32,
        pub queue_capacity: u32,
        pub flags: u32,
        pub gen_blocks: u32,
        pub acc_blocks: u32,
        pub reserved: [u32; 3],
    } // 32 B

    #[repr(C)]
    #[derive(Clone, Copy, Debug, Default)]
    pub struct RtSceneOptions {
        pub bvh_builder: u32,
        pub bvh_max_leaf: u32,
        pub bvh_treelet_nodes: u32,
        pub staged_nodes: u32,
        pub flags: u32,
        pub bvh_threads: u32,
        pub reserved: [u32; 2],
    } // 32 B

    #[repr(C)]
    #[derive(Clone, Copy, Debug, Default)]
    pub struct RtScenePlan {
        pub n_box_prims: u32,
        pub n_leaf_items: u32,
        pub n_nodes: u32,
        pub bvh_depth: u32,
        pub n_always: u32,
        pub whole_tree_threads: u32,
        pub whole_tree_smem: u64,
    } // 32 B

    #[repr(C)]
    pub struct RtContext {
        _private: [u8; 0],
    }
    #[repr(C)]
    pub struct RtScene {
        _private: [u8; 0],
    }
    #[repr(C)]
    pub struct RtMulti {
        _private: [u8; 0],
    }
    #[repr(C)]
    pub struct RtMultiScene {
        _private: [u8; 0],
    }

    extern "C" {
        pub fn rt_init(device_id: c_int, out: *mut *mut RtContext) -> c_int;
        pub fn rt_init_ex(device_id: c_int, options: *const RtContextOptions, out: *mut *mut RtContext) -> c_int;
        pub fn rt_shutdown(ctx: *mut RtContext);
        pub fn rt_options_from_env(ctx_options: *mut RtContextOptions, scene_options: *mut RtSceneOptions);
        pub fn rt_scene_create(ctx: *mut RtContext, desc: *const RtSceneDesc, out: *mut *mut RtScene) -> c_int;
        pub fn rt_scene_create_ex(ctx: *mut RtContext, desc: *const RtSceneDesc, options: *const RtSceneOptions, out: *mut *mut RtScene) -> c_int;
        pub fn rt_scene_destroy(scene: *mut RtScene);
        pub fn rt_render(scene: *mut RtScene, cam: *const RtCamera, params: *const RtRenderParams, out_sum_rgb: *mut f32, out_sumsq_rgb: *mut f32,
                         stats: *mut RtStats) -> c_int;
        pub fn rt_render_noalloc(scene: *mut RtScene, cam: *const RtCamera, params: *const RtRenderParams, out_sum_rgb: *mut f32, out_sumsq_rgb: *mut f32,
                                 stats: *mut RtStats) -> c_int;
        pub fn rt_render_device(scene: *mut RtScene, cam: *const RtCamera, params: *const RtRenderParams, d_sum_rgb: *mut f32, d_sumsq_rgb: *mut f32,
                                cuda_stream: *mut c_void, stats: *mut RtStats) -> c_int;
        pub fn rt_color_rays(scene: *mut RtScene, rays: *const RtRayIn, n: u64, spp: u32, max_depth: u32, seed: u64, out_sum_rgb: *mut f32,
                             out_sumsq_rgb: *mut f32, stats: *mut RtStats) -> c_int;
        pub fn rt_trace(scene: *mut RtScene, rays: *const RtRayIn, n: u64, seed: u64, out: *mut RtHitOut) -> c_int;
        pub fn rt_trace_from(scene: *mut RtScene, rays: *const RtRayIn, self_order: *const u32, n: u64, seed: u64, out: *mut RtHitOut) -> c_int;
        pub fn rt_resolve_u8(ctx: *mut RtContext, sum_rgb: *const f32, width: u32, height: u32, spp: u32, out_rgb_top_down: *mut u8) -> c_int;
        pub fn rt_deinterleave_rows(ctx: *mut RtContext, d_gathered: *const f32, d_full: *mut f32, width: u32, height: u32, n_shards: u32,
                                    cuda_stream: *mut c_void) -> c_int;
        pub fn rt_init_multi(n_devices: c_int, device_ids: *const c_int, options: *const RtContextOptions, out: *mut *mut RtMulti) -> c_int;
        pub fn rt_shutdown_multi(multi: *mut RtMulti);
        pub fn rt_multi_device_count(multi: *const RtMulti) -> c_int;
        pub fn rt_multi_context(multi: *mut RtMulti, index: c_int) -> *mut RtContext;
        pub fn rt_multi_scene_create(multi: *mut RtMulti, desc: *const RtSceneDesc, options: *const RtSceneOptions, out: *mut *mut RtMultiScene) -> c_int;
        pub fn rt_multi_scene_destroy(scene: *mut RtMultiScene);
        pub fn rt_render_multi(scene: *mut RtMultiScene, cam: *const RtCamera, params: *const RtRenderParams, out_sum_rgb: *mut f32,
                               out_sumsq_rgb: *mut f32, stats: *mut RtStats) -> c_int;
        pub fn rt_scene_bvh_info(scene: *const RtScene, n_nodes: *mut u32, n_always: *mut u32, max_depth: *mut u32) -> c_int;
        pub fn rt_scene_frame_rect_info(scene: *const RtScene, n_frames: *mut u32, n_boxes: *mut u32, n_loose: *mut u32) -> c_int;
        pub fn rt_scene_extend_info(scene: *const RtScene, n_box_prims: *mut u32, n_leaf_items: *mut u32, whole_tree_threads: *mut u32) -> c_int;
        pub fn rt_scene_plan(desc: *const RtSceneDesc, options: *const RtSceneOptions, out: *mut RtScenePlan) -> c_int;
        pub fn rt_scene_device_bytes(scene: *const RtScene) -> u64;
        pub fn rt_last_error() -> *const c_char;
        pub fn rt_abi_version() -> u32;
        // include/rt_b200_host.h: the encoders / decoder of the example drivers' image formats (the scene-building entry points of that
        // header serve callers WITHOUT a scene graph of their own — the ctypes harness — and are not needed from Rust)
        pub fn rth_encode_image(format: c_int, rgb_top_down: *const u8, width: u32, height: u32, out: *mut u8, capacity: u64, needed: *mut u64) -> c_int;
        pub fn rth_decode_image(bytes: *const u8, n: u64, width: *mut u32, height: *mut u32, out_rgb_top_down: *mut u8, capacity: u64,
                                needed: *mut u64) -> c_int;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Errors
// ---------------------------------------------------------------------------------------------------------------------
#[derive(Debug, Clone)]
pub enum GpuError {
    /// a scene node / material / texture without a device implementation (the reference's own types all have one)
    Unsupported(String),
    /// the C ABI returned a negative status; the string is rt_last_error()
    Backend(i32, String),
}

impl std::fmt::Display for GpuError {
    fn fmt(&self, f: &mut std::fmt::Formatter) -> std::fmt::Result {
        match self {
            GpuError::Unsupported(what) => write!(f, "unsupported scene node: {}", what),
            GpuError::Backend(code, msg) => write!(f, "rt_b200 error {}: {}", code, msg),
        }
    }
}
impl std::error::Error for GpuError {}

fn check(rc: c_int) -> Result<(), GpuError> {
    if rc == ffi::RT_OK {
        return Ok(());
    }
    let msg = unsafe {
        let p = ffi::rt_last_error();
        if p.is_null() { String::new() } else { CStr::from_ptr(p).to_string_lossy().into_owned() }
    };
    Err(GpuError::Backend(rc, msg))
}

// ---------------------------------------------------------------------------------------------------------------------
// Flatten state
// ---------------------------------------------------------------------------------------------------------------------

/// What is handed down the graph: the composed transform and the material of the OUTERMOST enclosing Sprite
/// (Sprite::hit overwrites the inner record's material unconditionally, None included: src/sprite.rs:124).
#[derive(Clone, Copy)]
pub struct FlattenState<'a> {
    pub world_from_local: Mat4,
    pub material: Option<&'a dyn Material>,
    pub in_sprite: bool,
}

impl<'a> FlattenState<'a> {
    pub fn root() -> Self {
        Self { world_from_local: Mat4::identity(), material: None, in_sprite: false }
    }
}

/// The flat tables (RtSceneDesc of include/rt_b200.h) under construction.
#[derive(Default)]
pub struct FlattenCtx {
    pub prims: Vec<ffi::RtPrim>,
    pub boundary_prims: Vec<ffi::RtPrim>,
    pub materials: Vec<ffi::RtMaterial>,
    pub textures: Vec<ffi::RtTexture>,
    pub texels: Vec<f32>, // RGBA f32
    /// DFS ranks handed out (>= prims.len(): leaves under a singular transform keep their rank but emit nothing)
    pub leaves_visited: u32,
    /// flattening the boundary of a ConstantMedium: leaves go to boundary_prims and take no rank
    in_boundary: bool,
    material_ids: HashMap<usize, i32>,
    texture_ids: HashMap<usize, i32>,
    /// grid on which ImageTexture closures are baked (examples/main.rs:267-280 looks up the nearest texel of a W x H image:
    /// baking at the image's own size reproduces that lookup exactly)
    pub bake_width: u32,
    pub bake_height: u32,
}

/// Row-major affine 3x4 (x' = m[0] x + m[1] y + m[2] z + m[3], ...) from the reference's column-major Mat4
/// (`a[index.1 * 4 + index.0]`, src/mat4.rs:390-396).
fn affine(m: &Mat4) -> [f32; 12] {
    let mut out = [0.0f32; 12];
    for r in 0..3 {
        for c in 0..4 {
            out[4 * r + c] = m[(r, c)] as f32;
        }
    }
    out
}

impl FlattenCtx {
    pub fn new() -> Self {
        Self { bake_width: 1024, bake_height: 512, ..Default::default() }
    }

    fn intern_texture(&mut self, t: &dyn Texture) -> Result<i32, GpuError> {
        let key = t as *const dyn Texture as *const u8 as usize;
        if let Some(id) = self.texture_ids.get(&key) {
            return Ok(*id);
        }
        let rec = t.describe_texture(self)?;
        let id = self.textures.len() as i32;
        self.textures.push(rec);
        self.texture_ids.insert(key, id);
        Ok(id)
    }

    fn intern_material(&mut self, m: Option<&dyn Material>) -> Result<i32, GpuError> {
        let m = match m {
            Some(m) => m,
            None => return Ok(-1), // hit without material returns black (src/render.rs:18-20)
        };
        let key = m as *const dyn Material as *const u8 as usize;
        if let Some(id) = self.material_ids.get(&key) {
            return Ok(*id);
        }
        let rec = m.describe(self)?;
        let id = self.materials.len() as i32;
        self.materials.push(rec);
        self.material_ids.insert(key, id);
        Ok(id)
    }

    /// One leaf primitive.  `p0`, `p1`: sphere radius / rectangle width, height / medium boundary radius, density.
    pub fn leaf(&mut self, kind: u32, st: &FlattenState, p0: f64, p1: f64) -> Result<(), GpuError> {
        let inv = st.world_from_local.inversed();
        if self.in_boundary {
            if let Some(inv) = inv {
                let mut p = ffi::RtPrim::default();
                p.kind = kind;
                p.material = -1;
                p.param[0] = p0 as f32;
                p.param[1] = p1 as f32;
                p.local_from_world = affine(&inv);
                p.world_from_local = affine(&st.world_from_local);
                self.boundary_prims.push(p);
            }
            return Ok(());
        }
        let order = self.leaves_visited;
        self.leaves_visited += 1;
        let inv = match inv {
            Some(inv) => inv,
            None => return Ok(()), // det == 0: the sprite is never hit (src/sprite.rs:131-134); the rank stays reserved
        };
        let mut p = ffi::RtPrim::default();
        p.kind = kind;
        p.material = self.intern_material(st.material)?;
        p.order = order;
        p.param[0] = p0 as f32;
        p.param[1] = p1 as f32;
        p.local_from_world = affine(&inv);
        p.world_from_local = affine(&st.world_from_local);
        self.prims.push(p);
        Ok(())
    }

    /// ConstantMedium over a boundary that is not a plain Sphere (src/volume.rs:40-43 takes any Hit).
    pub fn medium_over(&mut self, boundary: &dyn Flatten, st: &FlattenState, density: f64) -> Result<(), GpuError> {
        if self.in_boundary {
            return Err(GpuError::Unsupported("ConstantMedium inside the boundary of a ConstantMedium".into()));
        }
        let order = self.leaves_visited;
        self.leaves_visited += 1;
        let first = self.boundary_prims.len() as u32;
        self.in_boundary = true;
        let res = boundary.flatten(self, st);
        self.in_boundary = false;
        res?;
        let count = self.boundary_prims.len() as u32 - first;
        if count == 0 {
            return Ok(());
        }
        if count >= (1 << 24) {
            return Err(GpuError::Unsupported("ConstantMedium boundary with 2^24 or more surfaces".into()));
        }
        let mut p = ffi::RtPrim::default();
        p.kind = ffi::RT_PRIM_MEDIUM;
        p.material = self.intern_material(st.material)?;
        p.order = order;
        p.aux = first;
        p.param[1] = density as f32;
        p.param[2] = count as f32;
        p.local_from_world = affine(&Mat4::identity());
        p.world_from_local = affine(&Mat4::identity());
        self.prims.push(p);
        Ok(())
    }

    pub fn desc(&self) -> ffi::RtSceneDesc {
        ffi::RtSceneDesc {
            abi_version: ffi::RT_ABI_VERSION,
            n_prims: self.prims.len() as u32,
            n_materials: self.materials.len() as u32,
            n_textures: self.textures.len() as u32,
            prims: self.prims.as_ptr(),
            materials: self.materials.as_ptr(),
            textures: self.textures.as_ptr(),
            texels: self.texels.as_ptr(),
            n_texels: (self.texels.len() / 4) as u64,
            boundary_prims: self.boundary_prims.as_ptr(),
            n_boundary_prims: self.boundary_prims.len() as u32,
            reserved: 0,
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// The flatten hook.  `pub trait Hit: Send + Sync + Flatten` (src/ray.rs:85-87 patched) makes it reachable through every
// `dyn Hit` and — because `Bound<T>: Hit` (src/optimize.rs:88-95) — through every `dyn Bound<AxisAlignedBoundingBox>`.
// The default body reports "unsupported", so a downstream `impl Flatten for MyShape {}` keeps user types compiling.
// ---------------------------------------------------------------------------------------------------------------------
pub trait Flatten {
    fn flatten(&self, _ctx: &mut FlattenCtx, _st: &FlattenState) -> Result<(), GpuError> {
        Err(GpuError::Unsupported(std::any::type_name::<Self>().to_string()))
    }
    /// `Some(radius)` for a plain Sphere: ConstantMedium<Sphere> keeps its analytic two-root device form.
    fn as_sphere_radius(&self) -> Option<f64> {
        None
    }
}

/// Sphere::hit (src/geometry.rs:42-74)
impl Flatten for Sphere {
    fn flatten(&self, ctx: &mut FlattenCtx, st: &FlattenState) -> Result<(), GpuError> {
        ctx.leaf(ffi::RT_PRIM_SPHERE, st, self.radius(), 0.0)
    }
    fn as_sphere_radius(&self) -> Option<f64> {
        Some(self.radius())
    }
}

/// Rectangle::hit (src/geometry.rs:152-181)
impl Flatten for Rectangle {
    fn flatten(&self, ctx: &mut FlattenCtx, st: &FlattenState) -> Result<(), GpuError> {
        ctx.leaf(ffi::RT_PRIM_RECT, st, self.width(), self.height())
    }
}

/// A box is only ever a cull volume (HitRecord::default() has t = +inf, src/ray.rs:89-99): it contributes no primitive.
impl Flatten for AxisAlignedBoundingBox {
    fn flatten(&self, _ctx: &mut FlattenCtx, _st: &FlattenState) -> Result<(), GpuError> {
        Ok(())
    }
}

/// TransformedGeometry::hit (src/geometry.rs:210-246): ray into object space with M^-1, point and normal back with M; the inner
/// record's material is dropped (src/geometry.rs:234), which the state already expresses (only Sprites set one).
impl<T> Flatten for TransformedGeometry<T>
where
    T: Hit,
{
    fn flatten(&self, ctx: &mut FlattenCtx, st: &FlattenState) -> Result<(), GpuError> {
        let mut inner = *st;
        inner.world_from_local = st.world_from_local.multiplied(self.transform().origin());
        self.geometry().flatten(ctx, &inner)
    }
}

/// Sprite::hit (src/sprite.rs:87-139).  The outermost Sprite's material wins (src/sprite.rs:124), None included.
impl<T, U> Flatten for Sprite<T, U>
where
    T: Hit,
    U: Material + 'static,
{
    fn flatten(&self, ctx: &mut FlattenCtx, st: &FlattenState) -> Result<(), GpuError> {
        let geometry = match self.geometry() {
            Some(g) => g,
            None => return Ok(()), // a Sprite without geometry is never hit (src/sprite.rs:136-138)
        };
        let mut inner = *st;
        inner.world_from_local = st.world_from_local.multiplied(self.transform().origin());
        if !st.in_sprite {
            inner.material = self.material().as_ref().map(|m| m.as_ref() as &dyn Material);
        }
        inner.in_sprite = true;
        geometry.flatten(ctx, &inner)
    }
}

/// ConstantMedium::hit (src/volume.rs:46-100).
impl<T> Flatten for ConstantMedium<T>
where
    T: Hit,
{
    fn flatten(&self, ctx: &mut FlattenCtx, st: &FlattenState) -> Result<(), GpuError> {
        match self.boundary().as_sphere_radius() {
            Some(radius) => ctx.leaf(ffi::RT_PRIM_MEDIUM_SPHERE, st, radius, self.density()),
            None => ctx.medium_over(self.boundary().as_ref(), st, self.density()),
        }
    }
}

/// BoundingVolumeHierarchyNode::hit (src/optimize.rs:469-498): left before right, strict `<` — the DFS order that breaks
/// closest-hit ties is the order of emission here.  The device rebuilds its own tree over the flattened primitives.
impl<T> Flatten for BoundingVolumeHierarchyNode<T>
where
    T: Bound<T>,
{
    fn flatten(&self, ctx: &mut FlattenCtx, st: &FlattenState) -> Result<(), GpuError> {
        if let Some(left) = self.left() {
            left.flatten(ctx, st)?;
        }
        if let Some(right) = self.right() {
            right.flatten(ctx, st)?;
        }
        Ok(())
    }
}

/// The four linear lists (src/geometry.rs:76-116, src/optimize.rs:256-334): closest hit in list order with strict `<`.
impl Flatten for Vec<Box<dyn Hit>> {
    fn flatten(&self, ctx: &mut FlattenCtx, st: &FlattenState) -> Result<(), GpuError> {
        self.iter().try_for_each(|v| v.flatten(ctx, st))
    }
}
impl Flatten for Vec<Arc<dyn Hit>> {
    fn flatten(&self, ctx: &mut FlattenCtx, st: &FlattenState) -> Result<(), GpuError> {
        self.iter().try_for_each(|v| v.flatten(ctx, st))
    }
}
impl Flatten for Vec<Box<dyn Bound<AxisAlignedBoundingBox>>> {
    fn flatten(&self, ctx: &mut FlattenCtx, st: &FlattenState) -> Result<(), GpuError> {
        self.iter().try_for_each(|v| v.flatten(ctx, st))
    }
}
impl Flatten for Vec<Arc<dyn Bound<AxisAlignedBoundingBox>>> {
    fn flatten(&self, ctx: &mut FlattenCtx, st: &FlattenState) -> Result<(), GpuError> {
        self.iter().try_for_each(|v| v.flatten(ctx, st))
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Materials and textures.  `pub trait Material: Send + Sync + Debug + Describe` (src/material.rs:13 patched),
// `pub trait Texture: Send + Sync + Debug + DescribeTexture` (src/material.rs:196 patched).
// ---------------------------------------------------------------------------------------------------------------------
pub trait Describe {
    fn describe(&self, _ctx: &mut FlattenCtx) -> Result<ffi::RtMaterial, GpuError> {
        Err(GpuError::Unsupported(std::any::type_name::<Self>().to_string()))
    }
}
pub trait DescribeTexture {
    fn describe_texture(&self, _ctx: &mut FlattenCtx) -> Result<ffi::RtTexture, GpuError> {
        Err(GpuError::Unsupported(std::any::type_name::<Self>().to_string()))
    }
}

/// Lambertian::scatter (src/material.rs:60-70)
impl Describe for Lambertian {
    fn describe(&self, ctx: &mut FlattenCtx) -> Result<ffi::RtMaterial, GpuError> {
        Ok(ffi::RtMaterial { kind: ffi::RT_MAT_LAMBERTIAN, texture: ctx.intern_texture(self.albedo().as_ref())?, fuzz: 0.0, ior: 1.0 })
    }
}
/// Metal::scatter (src/material.rs:98-119)
impl Describe for Metal {
    fn describe(&self, ctx: &mut FlattenCtx) -> Result<ffi::RtMaterial, GpuError> {
        Ok(ffi::RtMaterial { kind: ffi::RT_MAT_METAL, texture: ctx.intern_texture(self.albedo().as_ref())?, fuzz: self.fuzziness() as f32, ior: 1.0 })
    }
}
/// Dielectric::scatter (src/material.rs:140-193): attenuation (1, 1, 1), no texture
impl Describe for Dielectric {
    fn describe(&self, _ctx: &mut FlattenCtx) -> Result<ffi::RtMaterial, GpuError> {
        Ok(ffi::RtMaterial { kind: ffi::RT_MAT_DIELECTRIC, texture: -1, fuzz: 0.0, ior: self.refractive() as f32 })
    }
}
/// DiffuseLight (src/material.rs:274-298); `emission()` is one of the getters of rust/reference.patch
impl Describe for DiffuseLight {
    fn describe(&self, ctx: &mut FlattenCtx) -> Result<ffi::RtMaterial, GpuError> {
        Ok(ffi::RtMaterial { kind: ffi::RT_MAT_DIFFUSE_LIGHT, texture: ctx.intern_texture(self.emission().as_ref())?, fuzz: 0.0, ior: 1.0 })
    }
}
/// Isotropic::scatter (src/material.rs:317-326); `albedo()` is one of the getters of rust/reference.patch
impl Describe for Isotropic {
    fn describe(&self, ctx: &mut FlattenCtx) -> Result<ffi::RtMaterial, GpuError> {
        Ok(ffi::RtMaterial { kind: ffi::RT_MAT_ISOTROPIC, texture: ctx.intern_texture(self.albedo().as_ref())?, fuzz: 0.0, ior: 1.0 })
    }
}

/// SolidColor (src/material.rs:200-215): `value()` is the colour whatever uv and point are — no getter needed.
impl DescribeTexture for SolidColor {
    fn describe_texture(&self, _ctx: &mut FlattenCtx) -> Result<ffi::RtTexture, GpuError> {
        let c = self.value(&(0.0, 0.0), &Vec3::new(0.0, 0.0, 0.0));
        let mut t = ffi::RtTexture::default();
        t.kind = ffi::RT_TEX_SOLID;
        t.a = -1;
        t.b = -1;
        t.rgb = [c.r() as f32, c.g() as f32, c.b() as f32, 0.0];
        Ok(t)
    }
}
/// CheckerTexture (src/material.rs:217-245): `sin(20 pi u) sin(20 pi v) > 0` -> black child, else white; children may be checkers.
impl DescribeTexture for CheckerTexture {
    fn describe_texture(&self, ctx: &mut FlattenCtx) -> Result<ffi::RtTexture, GpuError> {
        let mut t = ffi::RtTexture::default();
        t.kind = ffi::RT_TEX_CHECKER;
        t.a = ctx.intern_texture(self.black().as_ref())?;
        t.b = ctx.intern_texture(self.white().as_ref())?;
        Ok(t)
    }
}
/// ImageTexture<F> (src/material.rs:247-265): an arbitrary closure of uv cannot run on the device; it is evaluated at the texel
/// centres of a bake_width x bake_height grid (row 0 = v near 1).  The device looks up px = floor(u W), py = floor((1 - v) H),
/// clamped — with the grid equal to the source image this IS the example's nearest lookup (examples/main.rs:271-274).
impl<F> DescribeTexture for ImageTexture<F>
where
    F: (Fn(&(f64, f64)) -> Vec3) + Send + Sync,
{
    fn describe_texture(&self, ctx: &mut FlattenCtx) -> Result<ffi::RtTexture, GpuError> {
        let (w, h) = (ctx.bake_width, ctx.bake_height);
        let mut t = ffi::RtTexture::default();
        t.kind = ffi::RT_TEX_IMAGE;
        t.a = -1;
        t.b = -1;
        t.width = w;
        t.height = h;
        t.texel_offset = (ctx.texels.len() / 4) as u64;
        ctx.texels.reserve((w as usize) * (h as usize) * 4);
        let origin = Vec3::new(0.0, 0.0, 0.0);
        for py in 0..h {
            for px in 0..w {
                let u = (px as f64 + 0.5) / w as f64;
                let v = 1.0 - (py as f64 + 0.5) / h as f64;
                let c = self.value(&(u, v), &origin);
                ctx.texels.extend_from_slice(&[c.r() as f32, c.g() as f32, c.b() as f32, 1.0]);
            }
        }
        Ok(t)
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Camera: lowerLeft / horizontal / vertical are private without getters (src/camera.rs:18-20), so they are recomputed from
// the public ones exactly as PerspectiveCamera::new derives them in f64 (src/camera.rs:34-45) — `u = up x w` is NOT
// normalised there and is not here — then rounded to f32.
// ---------------------------------------------------------------------------------------------------------------------
pub fn camera_constants(camera: &PerspectiveCamera) -> ffi::RtCamera {
    let eye = *camera.eye();
    let center = *camera.center();
    let up = camera.up().normalized(); // already normalised by new(); idempotent
    let height = (camera.fov() / 2.0).tan() * 2.0;
    let width = camera.aspect() * height;
    let w = (eye - center).normalized();
    let u = up.cross(&w);
    let v = w.cross(&u);
    let horizontal = u * width * camera.focusDistance();
    let vertical = v * height * camera.focusDistance();
    let lower_left = eye - horizontal / 2.0 - vertical / 2.0 - w * camera.focusDistance();
    let f3 = |a: Vec3| [a.x() as f32, a.y() as f32, a.z() as f32];
    ffi::RtCamera {
        eye: f3(eye),
        lower_left: f3(lower_left),
        horizontal: f3(horizontal),
        vertical: f3(vertical),
        lens_radius: camera.lensRadius() as f32,
        reserved: [0.0; 3],
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Gpu (RtContext), World (RtScene)
// ---------------------------------------------------------------------------------------------------------------------

/// One CUDA device.  Used from one thread at a time; several `Gpu`s may render concurrently from several threads.
pub struct Gpu {
    ctx: *mut ffi::RtContext,
}
unsafe impl Send for Gpu {}

impl Gpu {
    pub fn new(device: i32) -> Result<Self, GpuError> {
        let mut ctx = std::ptr::null_mut();
        check(unsafe { ffi::rt_init(device as c_int, &mut ctx) })?;
        Ok(Self { ctx })
    }

    pub fn with_options(device: i32, options: &ffi::RtContextOptions) -> Result<Self, GpuError> {
        let mut ctx = std::ptr::null_mut();
        check(unsafe { ffi::rt_init_ex(device as c_int, options, &mut ctx) })?;
        Ok(Self { ctx })
    }

    /// mean -> sqrt -> * 255 -> min(255) -> truncation, rows top to bottom: the bytes examples/book-one.rs:90-100 prints.
    /// `means` is what `render_image` returns (row-major, y up).
    pub fn resolve_u8(&self, means: &[Vec3], width: usize, height: usize) -> Result<Vec<u8>, GpuError> {
        assert_eq!(means.len(), width * height);
        let sums: Vec<f32> = means.iter().flat_map(|p| vec![p.r() as f32, p.g() as f32, p.b() as f32]).collect();
        let mut out = vec![0u8; width * height * 3];
        check(unsafe { ffi::rt_resolve_u8(self.ctx, sums.as_ptr(), width as u32, height as u32, 1, out.as_mut_ptr()) })?;
        Ok(out)
    }
}

impl Drop for Gpu {
    fn drop(&mut self) {
        unsafe { ffi::rt_shutdown(self.ctx) }
    }
}

/// Flatten any world the reference can render: `&dyn Hit`, `&dyn Bound<AxisAlignedBoundingBox>`, a `BoundingVolumeHierarchyNode`, a `Vec`.
pub fn flatten_world(world: &dyn Flatten) -> Result<FlattenCtx, GpuError> {
    let mut ctx = FlattenCtx::new();
    world.flatten(&mut ctx, &FlattenState::root())?;
    Ok(ctx)
}

/// A flattened, uploaded scene: device BVH + primitive / material / texture tables.  Borrows the `Gpu` it lives on.
pub struct World<'g> {
    scene: *mut ffi::RtScene,
    _gpu: &'g Gpu,
}

impl<'g> World<'g> {
    pub fn new(gpu: &'g Gpu, world: &dyn Flatten) -> Result<Self, GpuError> {
        Self::from_flat(gpu, &flatten_world(world)?)
    }

    pub fn from_flat(gpu: &'g Gpu, flat: &FlattenCtx) -> Result<Self, GpuError> {
        let desc = flat.desc();
        let mut scene = std::ptr::null_mut();
        check(unsafe { ffi::rt_scene_create(gpu.ctx, &desc, &mut scene) })?;
        Ok(Self { scene, _gpu: gpu })
    }

    /// (BVH nodes, primitives in the always-tested list, tree depth)
    pub fn bvh_info(&self) -> (u32, u32, u32) {
        let (mut n, mut a, mut d) = (0u32, 0u32, 0u32);
        unsafe { ffi::rt_scene_bvh_info(self.scene, &mut n, &mut a, &mut d) };
        (n, a, d)
    }
}

impl<'g> Drop for World<'g> {
    fn drop(&mut self) {
        unsafe { ffi::rt_scene_destroy(self.scene) }
    }
}

#[derive(Clone, Copy, Debug)]
pub struct RenderOptions {
    pub seed: u64,
    /// this call renders rows y = row_begin + k * row_stride — the reference's thread partition (examples/book-one.rs:63)
    pub row_begin: u32,
    pub row_stride: u32,
    pub queue_capacity: u32,
    pub flags: u32,
}

impl Default for RenderOptions {
    fn default() -> Self {
        Self { seed: 1, row_begin: 0, row_stride: 1, queue_capacity: 0, flags: 0 }
    }
}

fn to_ray_in(ray: &Ray) -> ffi::RtRayIn {
    let (o, d) = (ray.origin(), ray.direction());
    ffi::RtRayIn { origin: [o.x() as f32, o.y() as f32, o.z() as f32], direction: [d.x() as f32, d.y() as f32, d.z() as f32] }
}

/// Replaces
/// ```text
/// for y in (0..height).rev() { for x in 0..width { for _ in 0..spp {
///     let u = (x as f64 + U) / width as f64; let v = (y as f64 + U) / height as f64;
///     pixel += color(&camera.ray(u, v), world, maxDepth); } pixel /= spp; buffer[y][x] = pixel; } }
/// ```
/// (examples/book-one.rs:52-88).  Returns per-pixel MEANS, row-major with y up like `buffer[y][x]` (book-one.rs:53,86):
/// `result[k * width + x]` is row `y = row_begin + k * row_stride`.
pub fn render_image(world: &World, camera: &PerspectiveCamera, width: usize, height: usize, spp: usize, maxDepth: usize,
                    options: &RenderOptions) -> Result<(Vec<Vec3>, ffi::RtStats), GpuError> {
    let cam = camera_constants(camera);
    let prm = ffi::RtRenderParams {
        width: width as u32,
        height: height as u32,
        spp: spp as u32,
        max_depth: maxDepth as u32,
        seed: options.seed,
        row_begin: options.row_begin,
        row_stride: options.row_stride,
        queue_capacity: options.queue_capacity,
        flags: options.flags,
    };
    if options.row_stride == 0 || options.row_begin as usize >= height {
        return Err(GpuError::Backend(ffi::RT_ERR_INVALID, "bad row_begin / row_stride".into()));
    }
    let rows = (height - options.row_begin as usize + options.row_stride as usize - 1) / options.row_stride as usize;
    let mut sums = vec![0.0f32; rows * width * 3];
    let mut stats = ffi::RtStats::default();
    check(unsafe { ffi::rt_render(world.scene, &cam, &prm, sums.as_mut_ptr(), std::ptr::null_mut(), &mut stats) })?;
    let inv = 1.0 / spp as f64; // pixel /= subPixelSampleCount (book-one.rs:76)
    let out = sums.chunks_exact(3).map(|c| Vec3::new(c[0] as f64 * inv, c[1] as f64 * inv, c[2] as f64 * inv)).collect();
    Ok((out, stats))
}

/// What `world.hit(&ray)` returns (src/ray.rs:36-43), plus the DFS rank of the leaf that was hit.
#[derive(Clone, Copy, Debug)]
pub struct TraceHit {
    pub prim: u32,
    pub t: f64,
    pub intersection: Vec3,
    pub normal: Vec3,
    pub uv: (f64, f64),
    /// index into the flattened material table, -1 = none
    pub material: i32,
}

fn to_trace_hit(h: &ffi::RtHitOut) -> Option<TraceHit> {
    if h.hit == 0 {
        return None;
    }
    Some(TraceHit {
        prim: h.prim as u32,
        t: h.t as f64,
        intersection: Vec3::new(h.p[0] as f64, h.p[1] as f64, h.p[2] as f64),
        normal: Vec3::new(h.n[0] as f64, h.n[1] as f64, h.n[2] as f64),
        uv: (h.uv[0] as f64, h.uv[1] as f64),
        material: h.material,
    })
}

/// Batched `world.hit(&ray)` (src/optimize.rs:469-498).  ConstantMedium free-flight draws are keyed by (ray index, leaf, seed).
pub fn trace_rays(world: &World, rays: &[Ray], seed: u64) -> Result<Vec<Option<TraceHit>>, GpuError> {
    let input: Vec<ffi::RtRayIn> = rays.iter().map(to_ray_in).collect();
    let mut out = vec![ffi::RtHitOut::default(); rays.len()];
    check(unsafe { ffi::rt_trace(world.scene, input.as_ptr(), input.len() as u64, seed, out.as_mut_ptr()) })?;
    Ok(out.iter().map(to_trace_hit).collect())
}

/// The same for scattered rays that start ON a primitive: `from[i]` = `TraceHit::prim` of the hit ray i leaves from (None: free ray).
pub fn trace_rays_from(world: &World, rays: &[Ray], from: &[Option<u32>], seed: u64) -> Result<Vec<Option<TraceHit>>, GpuError> {
    assert_eq!(rays.len(), from.len());
    let input: Vec<ffi::RtRayIn> = rays.iter().map(to_ray_in).collect();
    let orders: Vec<u32> = from.iter().map(|o| o.unwrap_or(ffi::RT_NO_PRIM)).collect();
    let mut out = vec![ffi::RtHitOut::default(); rays.len()];
    check(unsafe { ffi::rt_trace_from(world.scene, input.as_ptr(), orders.as_ptr(), input.len() as u64, seed, out.as_mut_ptr()) })?;
    Ok(out.iter().map(to_trace_hit).collect())
}

/// Batched `render::color(&ray, world, maxDepth)` (src/render.rs:5-29): the mean of `spp` independent evaluations per ray.  The
/// direction is used as given (color() does not normalise it).
pub fn color(world: &World, rays: &[Ray], maxDepth: usize, spp: usize, seed: u64) -> Result<Vec<Vec3>, GpuError> {
    let input: Vec<ffi::RtRayIn> = rays.iter().map(to_ray_in).collect();
    let mut sums = vec![0.0f32; rays.len() * 3];
    check(unsafe {
        ffi::rt_color_rays(world.scene, input.as_ptr(), input.len() as u64, spp as u32, maxDepth as u32, seed, sums.as_mut_ptr(), std::ptr::null_mut(),
                           std::ptr::null_mut())
    })?;
    let inv = 1.0 / spp as f64;
    Ok(sums.chunks_exact(3).map(|c| Vec3::new(c[0] as f64 * inv, c[1] as f64 * inv, c[2] as f64 * inv)).collect())
}

// ---------------------------------------------------------------------------------------------------------------------
// Several GPUs of one process: the reference's `for i in 0..cpuCount { thread::spawn(.. if y % cpuCount != i { continue } ..) }`
// (examples/book-one.rs:56-66) with devices instead of cores; the mpsc funnel (book-one.rs:83-88) is one NCCL gather.
// ---------------------------------------------------------------------------------------------------------------------
pub struct MultiGpu {
    multi: *mut ffi::RtMulti,
}
unsafe impl Send for MultiGpu {}

impl MultiGpu {
    pub fn new(devices: &[i32]) -> Result<Self, GpuError> {
        let ids: Vec<c_int> = devices.iter().map(|d| *d as c_int).collect();
        let mut multi = std::ptr::null_mut();
        check(unsafe { ffi::rt_init_multi(ids.len() as c_int, ids.as_ptr(), std::ptr::null(), &mut multi) })?;
        Ok(Self { multi })
    }
    pub fn device_count(&self) -> usize {
        unsafe { ffi::rt_multi_device_count(self.multi) as usize }
    }
}
impl Drop for MultiGpu {
    fn drop(&mut self) {
        unsafe { ffi::rt_shutdown_multi(self.multi) }
    }
}

pub struct MultiWorld<'g> {
    scene: *mut ffi::RtMultiScene,
    _gpus: &'g MultiGpu,
}
impl<'g> MultiWorld<'g> {
    pub fn new(gpus: &'g MultiGpu, world: &dyn Flatten) -> Result<Self, GpuError> {
        let flat = flatten_world(world)?;
        let desc = flat.desc();
        let mut scene = std::ptr::null_mut();
        check(unsafe { ffi::rt_multi_scene_create(gpus.multi, &desc, std::ptr::null(), &mut scene) })?;
        Ok(Self { scene, _gpus: gpus })
    }

    /// `render_image` over every device of the `MultiGpu`: device i renders the rows y % n == i, the image is reassembled on device 0.
    pub fn render_image(&self, camera: &PerspectiveCamera, width: usize, height: usize, spp: usize, maxDepth: usize, seed: u64)
                        -> Result<(Vec<Vec3>, ffi::RtStats), GpuError> {
        let cam = camera_constants(camera);
        let prm = ffi::RtRenderParams {
            width: width as u32, height: height as u32, spp: spp as u32, max_depth: maxDepth as u32, seed,
            row_begin: 0, row_stride: 1, queue_capacity: 0, flags: 0,
        };
        let mut sums = vec![0.0f32; width * height * 3];
        let mut stats = ffi::RtStats::default();
        check(unsafe { ffi::rt_render_multi(self.scene, &cam, &prm, sums.as_mut_ptr(), std::ptr::null_mut(), &mut stats) })?;
        let inv = 1.0 / spp as f64;
        Ok((sums.chunks_exact(3).map(|c| Vec3::new(c[0] as f64 * inv, c[1] as f64 * inv, c[2] as f64 * inv)).collect(), stats))
    }
}
impl<'g> Drop for MultiWorld<'g> {
    fn drop(&mut self) {
        unsafe { ffi::rt_multi_scene_destroy(self.scene) }
    }
}

/// Encode the bytes of `Gpu::resolve_u8` as the example drivers do: format 0 = PPM P3 text (examples/book-one.rs:28-30,90-100),
/// format 1 = 8-bit RGBA PNG, alpha 255, top row first (examples/main.rs:105-135).
pub fn encode_image(format: i32, rgb_top_down: &[u8], width: usize, height: usize) -> Result<Vec<u8>, GpuError> {
    assert_eq!(rgb_top_down.len(), width * height * 3);
    let mut needed = 0u64;
    check(unsafe { ffi::rth_encode_image(format as c_int, rgb_top_down.as_ptr(), width as u32, height as u32, std::ptr::null_mut(), 0, &mut needed) })?;
    let mut out = vec![0u8; needed as usize];
    check(unsafe { ffi::rth_encode_image(format as c_int, rgb_top_down.as_ptr(), width as u32, height as u32, out.as_mut_ptr(), needed, &mut needed) })?;
    Ok(out)
}
