// examples/book-one-gpu.rs — examples/book-one.rs with its pixel loop replaced by ONE call.
//
// `randomScene()` and the camera are the example's own (examples/book-one.rs:35-47,103-205: include that file's function here or
// `mod` it); what changes is main(): the `for i in 0..cpuCount { thread::spawn(..) }` nest and the mpsc funnel
// (book-one.rs:52-88) become `gpu::render_image`, and the `println!` loop (book-one.rs:90-100) becomes `Gpu::resolve_u8` +
// `gpu::encode_image` — same "P3\n<w> <h>\n255\n" header, same `(c.sqrt() * 255.0).min(255.0) as usize` bytes, rows top to bottom.
use ray_tracer::camera::PerspectiveCamera;
use ray_tracer::gpu::{self, Gpu, RenderOptions, World};
use ray_tracer::optimize::{AxisAlignedBoundingBox, Bound, BoundingVolumeHierarchyNode};
use ray_tracer::vec3::Vec3;

use std::io::Write;
use std::sync::Arc;

mod book_one_scene; // fn randomScene() -> Vec<Arc<dyn Bound<AxisAlignedBoundingBox>>>, verbatim from examples/book-one.rs:103-205

fn main() -> Result<(), Box<dyn std::error::Error>> {
    let (width, height) = (1600usize, 800usize);
    let scene: Vec<Arc<dyn Bound<AxisAlignedBoundingBox>>> = book_one_scene::randomScene();
    let world = BoundingVolumeHierarchyNode::new(scene).unwrap(); // the reference's own tree: its DFS order is the tie-break contract

    let camera = PerspectiveCamera::new(
        Vec3::new(13.0, 2.0, 3.0),
        Vec3::new(0.0, 0.0, 0.0),
        Vec3::new(0.0, 1.0, 0.0),
        (20.0 as f64).to_radians(),
        width as f64 / height as f64,
        10.0,
        0.05,
    );

    let device = Gpu::new(0)?; // no CPU fallback: Err without an sm_100 device
    let device_world = World::new(&device, &world)?;
    let (pixels, stats) = gpu::render_image(&device_world, &camera, width, height, 100, 100, &RenderOptions::default())?;
    eprintln!("{} paths, {} rays, {:.1} ms on the device", stats.paths, stats.rays, stats.ms_total);

    let bytes = device.resolve_u8(&pixels, width, height)?;
    std::io::stdout().write_all(&gpu::encode_image(0, &bytes, width, height)?)?; // PPM P3, byte for byte what book-one.rs prints
    Ok(())
}
