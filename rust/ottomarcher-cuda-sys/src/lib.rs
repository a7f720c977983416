//! FFI mirror of include/ottomarcher.h (B200 / sm_100a renderer for ottomarcher).
//! NOT COMPILED IN THIS IMAGE (no rustc/cargo): written to the header, checked by eye and by the
//! ctypes mirror's sizeof/offset tests (tests/test_host_abi.py), which bind the same structs.
#![allow(non_camel_case_types, non_upper_case_globals, dead_code)]
use std::os::raw::{c_char, c_void};

pub type om_status = u32;
pub const OM_OK: om_status = 0;
pub const OM_ERR_INVALID: om_status = 1;
pub const OM_ERR_CUDA: om_status = 2;
pub const OM_ERR_UNSUPPORTED: om_status = 3;
pub const OM_ERR_OOM: om_status = 4;
pub const OM_ERR_NO_DEVICE: om_status = 5;

pub const OM_RENDER_ADAPTIVE: u32 = 1 << 0;
pub const OM_RENDER_FAST_MATH: u32 = 1 << 1;
pub const OM_RENDER_BRUTE_FORCE: u32 = 1 << 2;
pub const OM_RENDER_NO_SMEM_SCENE: u32 = 1 << 4;
pub const OM_RENDER_COUNTERS: u32 = 1 << 5;
pub const OM_RENDER_WAVEFRONT: u32 = 1 << 3;
pub const OM_RENDER_PERSISTENT: u32 = 1 << 6;
pub const OM_RENDER_PHASE_ALIGN: u32 = 1 << 7;
pub const OM_RENDER_NO_PHASE_ALIGN: u32 = 1 << 8;
pub const OM_FREEZE_NO_ALWAYS_LIST: u32 = 1 << 0;
pub const OM_FREEZE_RAY_PADDING: u32 = 1 << 1;
pub const OM_FREEZE_STATIC_PADDING: u32 = 1 << 2;
pub const OM_FREEZE_NO_MARCHED_BVH: u32 = 1 << 3;
pub const OM_FREEZE_GPU_BVH: u32 = 1 << 4;

/// materials.rs:11-23
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct om_material { pub albedo: [f32; 3], pub fuzz: f32, pub ior: f32, pub mat_type: u32 }
/// traced.rs:13-19 / 229-235 (Sphere and Cube share the layout)
#[repr(C)] #[derive(Clone, Copy)]
pub struct om_sphere { pub m_local_to_world: [f32; 16], pub m_world_to_local: [f32; 16], pub material: om_material }
pub type om_cube = om_sphere;
/// traced.rs:77-82
#[repr(C)] #[derive(Clone, Copy)]
pub struct om_infinite_plane { pub center: [f32; 3], pub normal: [f32; 3], pub material: om_material }
/// traced.rs:118-131
#[repr(C)] #[derive(Clone, Copy)]
pub struct om_barycentric {
    pub origin: [f32; 3], pub u: [f32; 3], pub u_length: f32, pub v: [f32; 3], pub v_length: f32,
    pub uxv: [f32; 3], pub uxvxu: [f32; 3], pub base_inv: [f32; 9], pub v_in_base: [f32; 3], pub material: om_material,
}
/// marched.rs:49-54, 78-83, 104-113
#[repr(C)] #[derive(Clone, Copy)]
pub struct om_marched_sphere { pub center: [f32; 3], pub radius: f32, pub material: om_material }
#[repr(C)] #[derive(Clone, Copy)]
pub struct om_marched_box { pub center: [f32; 3], pub sizes: [f32; 3], pub material: om_material }
#[repr(C)] #[derive(Clone, Copy)]
pub struct om_marched_torus {
    pub m_local_to_world_translate_rotate: [f32; 16], pub m_world_to_local_translate_rotate: [f32; 16],
    pub m_local_to_world_scale: [f32; 4], pub m_world_to_local_scale: [f32; 4], pub sizes: [f32; 3], pub material: om_material,
}
/// `Arc<dyn Marched + Send + Sync>` as data (marched.rs:8-45): to_local / to_world matrices, to_world_f factor, world bounds
/// (or +-INF); local_sdf travels beside it as a postfix f32 program (OM_SDF_* opcodes, see ottomarcher.h).
#[repr(C)] #[derive(Clone, Copy)]
pub struct om_dyn_marched {
    pub m_world_to_local: [f32; 16], pub m_local_to_world: [f32; 16], pub world_f: f32,
    pub bounds_lo: [f32; 3], pub bounds_hi: [f32; 3], pub material: om_material,
}
pub const OM_SDF_X: u32 = 1; pub const OM_SDF_Y: u32 = 2; pub const OM_SDF_Z: u32 = 3; pub const OM_SDF_CONST: u32 = 4;
pub const OM_SDF_ADD: u32 = 5; pub const OM_SDF_SUB: u32 = 6; pub const OM_SDF_MUL: u32 = 7; pub const OM_SDF_DIV: u32 = 8;
pub const OM_SDF_MIN: u32 = 9; pub const OM_SDF_MAX: u32 = 10; pub const OM_SDF_NEG: u32 = 11; pub const OM_SDF_ABS: u32 = 12;
pub const OM_SDF_SQRT: u32 = 13; pub const OM_SDF_LEN2: u32 = 14; pub const OM_SDF_LEN3: u32 = 15;
pub const OM_SDF_STORE: u32 = 16; pub const OM_SDF_LOAD: u32 = 17;
pub const fn om_sdf_insn(op: u32, arg: u32) -> u32 { op | (arg << 8) }
/// A user type opts in by describing itself as data; `impl DeviceMarched for MyShape` next to `impl Marched for MyShape`.
pub trait DeviceMarched {
    fn describe(&self) -> (om_dyn_marched, Vec<u32>);
}
/// camera.rs:9-29
#[repr(C)] #[derive(Clone, Copy)]
pub struct om_camera {
    pub origin: [f32; 3], pub horizontal: [f32; 3], pub vertical: [f32; 3], pub lower_left_corner: [f32; 3],
    pub u_of_plane: [f32; 3], pub v_of_plane: [f32; 3], pub w_of_plane: [f32; 3],
    pub lens_radius: f32, pub aspect_ratio: f32, pub focus_dist: f32, pub viewport_width: f32, pub viewport_height: f32,
}
/// render_thread.rs:9-17 (Stats), 40 bytes
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct om_pixel { pub sum: [f32; 3], pub n: u32, pub avg_depth: f32, pub bad_avgs: u32, pub bloom: u64, pub color: [u8; 3], pub _pad: [u8; 5] }
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct om_render_params {
    pub width: u32, pub height: u32, pub spp_max: u32, pub max_depth: u32, pub tmin: f32, pub tmax: f32, pub seed: u64,
    pub flags: u32, pub sample_base: u32, pub sample_count: u32, pub part_rank: u32, pub part_world: u32, pub device: i32,
}
#[repr(C)] pub struct om_scene { _p: [u8; 0] }
#[repr(C)] pub struct om_render { _p: [u8; 0] }

extern "C" {
    pub fn om_scene_create(out: *mut *mut om_scene) -> om_status;
    pub fn om_scene_clear(s: *mut om_scene) -> om_status;
    pub fn om_scene_destroy(s: *mut om_scene) -> om_status;
    pub fn om_scene_add_spheres(s: *mut om_scene, a: *const om_sphere, n: usize) -> om_status;
    pub fn om_scene_add_cubes(s: *mut om_scene, a: *const om_cube, n: usize) -> om_status;
    pub fn om_scene_add_triangles(s: *mut om_scene, a: *const om_barycentric, n: usize) -> om_status;
    pub fn om_scene_add_infinite_planes(s: *mut om_scene, a: *const om_infinite_plane, n: usize) -> om_status;
    pub fn om_scene_add_parallelograms(s: *mut om_scene, a: *const om_barycentric, n: usize) -> om_status;
    pub fn om_scene_add_marched_spheres(s: *mut om_scene, a: *const om_marched_sphere, n: usize) -> om_status;
    pub fn om_scene_add_marched_boxes(s: *mut om_scene, a: *const om_marched_box, n: usize) -> om_status;
    pub fn om_scene_add_marched_tori(s: *mut om_scene, a: *const om_marched_torus, n: usize) -> om_status;
    pub fn om_scene_add_dyn_marched(s: *mut om_scene, obj: *const om_dyn_marched, code: *const u32, n_words: usize) -> om_status;
    pub fn om_scene_dyn_marched_count(s: *const om_scene, count: *mut u64) -> om_status;
    pub fn om_scene_get_dyn_marched(s: *const om_scene, index: u64, obj: *mut om_dyn_marched, code: *mut u32, n_words: *mut usize) -> om_status;
    pub fn om_scene_add_dyn_object(s: *mut om_scene, trait_object: *const c_void) -> om_status;
    pub fn om_trim_device_cache() -> om_status;
    pub fn om_scene_freeze(s: *mut om_scene, cam: *const om_camera, flags: u32) -> om_status;
    pub fn om_render_create(out: *mut *mut om_render, frozen: *const om_scene, cam: *const om_camera, p: *const om_render_params) -> om_status;
    pub fn om_render_destroy(r: *mut om_render) -> om_status;
    pub fn om_render_set_stream(r: *mut om_render, cuda_stream: *mut c_void) -> om_status;
    pub fn om_render_launch(r: *mut om_render, spp_this_pass: u32) -> om_status;
    pub fn om_render_sync(r: *mut om_render) -> om_status;
    pub fn om_render_samples_done(r: *mut om_render, credited: *mut u64) -> om_status;
    pub fn om_render_samples_traced(r: *mut om_render, traced: *mut u64) -> om_status;
    pub fn om_render_read_pixels(r: *mut om_render, host: *mut om_pixel, n: usize) -> om_status;
    pub fn om_render_read_rgb(r: *mut om_render, host_rgb: *mut u8, nbytes: usize) -> om_status;
    pub fn om_render_read_view(r: *mut om_render, mode: u32, host_rgb: *mut u8, nbytes: usize) -> om_status;
    /// no counterpart in the reference; fixed device scratch, at most OM_VIEW_NOALLOC_MAX_PIXELS (2 097 152) pixels
    pub fn om_render_read_view_noalloc(r: *mut om_render, mode: u32, host_rgb: *mut u8, nbytes: usize) -> om_status;
    pub fn om_write_bmp(path: *const c_char, rgb: *const u8, width: u32, height: u32) -> om_status;
    pub fn om_render_reset(r: *mut om_render) -> om_status;
    pub fn om_render_primary(r: *mut om_render, mode: u32, sample_index: u32, ids: *mut u64, t: *mut f32, n: usize) -> om_status;
    pub fn om_render_set_chunk_owners(r: *mut om_render, owners: *const u8, n_chunks: usize, my_id: u32) -> om_status;
    pub fn om_render_export_state(r: *mut om_render, dev_state: *mut u32, only_owned: u32) -> om_status;
    pub fn om_render_import_state(r: *mut om_render, dev_state: *const u32) -> om_status;
    pub fn om_render_live_per_chunk(r: *mut om_render, host_counts: *mut u32, n_chunks: usize) -> om_status;
    pub fn om_last_error() -> *const c_char;
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
