//! FFI binding of include/ottomarcher.h (B200 / sm_100a renderer for ottomarcher) and the safe wrapper that replaces
//! main.rs:196-214.
//! NOT COMPILED IN THIS IMAGE (no rustc/cargo).  What keeps it honest instead:
//!  * `ffi.rs` (constants, #[repr(C)] structs, the extern "C" block) is GENERATED from the header by tools/gen_rust_ffi.py;
//!  * tests/test_rust_ffi.py fails when it is stale, when a symbol / argument count differs from the header or from the
//!    built library's exports, and when a struct's size or a field's offset differs from the ctypes mirrors that every GPU
//!    test binds (raytracingoneweekend_b200/_abi.py).
#![allow(non_camel_case_types, non_upper_case_globals, dead_code)]
mod ffi;
pub use ffi::*;
/// `From<&Sphere> for om_sphere` and friends, written against the reference's own types; compiled inside the ottomarcher crate
/// (`--features ottomarcher`), where those types live.
#[cfg(feature = "ottomarcher")]
pub mod convert;

impl Default for om_pixel {
    fn default() -> Self { om_pixel { sum: [0.0; 3], n: 0, avg_depth: 0.0, bad_avgs: 0, bloom: 0, color: [0; 3], _pad: [0; 5] } }
}
impl Default for om_render_params {
    fn default() -> Self {
        om_render_params { width: 0, height: 0, spp_max: 0, max_depth: 50, tmin: 0.001, tmax: 100.0, seed: 0, flags: 0, sample_base: 0, sample_count: 0,
                           part_rank: 0, part_world: 1, device: -1 }
    }
}
/// A user type opts in to the device by describing itself as data: `impl DeviceMarched for MyShape` next to `impl Marched for MyShape`
/// (marched_objects: Vec<Arc<dyn Marched + Send + Sync>>, hits.rs:40).
pub trait DeviceMarched {
    fn describe(&self) -> (om_dyn_marched, Vec<u32>);
}

/// Error text of the last failing call on this thread.
pub fn last_error() -> String {
    unsafe { std::ffi::CStr::from_ptr(om_last_error()).to_string_lossy().into_owned() }
}
fn ck(s: om_status) -> Result<(), String> { if s == OM_OK { Ok(()) } else { Err(last_error()) } }

/// The lists of a `HittableList` (hits.rs:37-43), already converted element by element
/// (`From<&Sphere> for om_sphere` etc. live in the ottomarcher crate, next to the types).
#[derive(Default)]
pub struct SceneLists {
    pub spheres: Vec<om_sphere>, pub cubes: Vec<om_cube>, pub triangles: Vec<om_barycentric>,
    pub infinite_planes: Vec<om_infinite_plane>, pub parallelograms: Vec<om_barycentric>,
    pub marched_spheres: Vec<om_marched_sphere>, pub marched_boxes: Vec<om_marched_box>, pub marched_tori: Vec<om_marched_torus>,
    /// marched_objects: Vec<Arc<dyn Marched + Send + Sync>> (hits.rs:40) whose types implement DeviceMarched
    pub marched_objects: Vec<(om_dyn_marched, Vec<u32>)>,
}

/// Replaces main.rs:196-214 (spawn of the render threads): renders `spp` samples per pixel
/// (a maximum when OM_RENDER_ADAPTIVE is set) in passes of `pass_spp`, calling `progress(done,total)`
/// between passes (the log thread's counter contract, main.rs:151-168).
pub fn render_gpu(cam: &om_camera, world: &SceneLists, max_depth: u32, tmin: f32, tmax: f32, spp: u32, width: u32, height: u32,
                  seed: u64, flags: u32, pass_spp: u32, mut progress: impl FnMut(u64, u64)) -> Result<Vec<om_pixel>, String> {
    unsafe {
        let mut s: *mut om_scene = std::ptr::null_mut();
        ck(om_scene_create(&mut s))?;
        let res = (|| {
            ck(om_scene_add_spheres(s, world.spheres.as_ptr(), world.spheres.len()))?;
            ck(om_scene_add_cubes(s, world.cubes.as_ptr(), world.cubes.len()))?;
            ck(om_scene_add_triangles(s, world.triangles.as_ptr(), world.triangles.len()))?;
            ck(om_scene_add_infinite_planes(s, world.infinite_planes.as_ptr(), world.infinite_planes.len()))?;
            ck(om_scene_add_parallelograms(s, world.parallelograms.as_ptr(), world.parallelograms.len()))?;
            ck(om_scene_add_marched_spheres(s, world.marched_spheres.as_ptr(), world.marched_spheres.len()))?;
            ck(om_scene_add_marched_boxes(s, world.marched_boxes.as_ptr(), world.marched_boxes.len()))?;
            ck(om_scene_add_marched_tori(s, world.marched_tori.as_ptr(), world.marched_tori.len()))?;
            for (obj, code) in &world.marched_objects { ck(om_scene_add_dyn_marched(s, obj, code.as_ptr(), code.len()))?; }
            ck(om_scene_freeze(s, cam, 0))?;
            let p = om_render_params { width, height, spp_max: spp, max_depth, tmin, tmax, seed, flags, device: -1, ..Default::default() };
            let mut r: *mut om_render = std::ptr::null_mut();
            ck(om_render_create(&mut r, s, cam, &p))?;
            let out = (|| {
                let total = width as u64 * height as u64 * spp as u64;
                let mut done = 0u64;
                let mut passes = 0u32;
                while done < total && passes * pass_spp.max(1) < spp {
                    ck(om_render_launch(r, pass_spp))?;
                    ck(om_render_samples_done(r, &mut done))?;
                    progress(done, total);
                    passes += 1;
                }
                let mut px = vec![om_pixel::default(); (width * height) as usize];
                ck(om_render_read_pixels(r, px.as_mut_ptr(), px.len()))?;
                Ok(px)
            })();
            om_render_destroy(r);
            out
        })();
        om_scene_destroy(s);
        res
    }
}
