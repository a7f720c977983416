//! Conversions from the reference's own types (octaviogarcia/RaytracingOneWeekend, src/*.rs) to the C ABI's PODs: every field is
//! copied one to one, matrices row by row (`Mat4x4.e[row].e[col]`, mat4x4.rs:8-10 — the ABI is row-major too).
//! NOT COMPILED IN THIS IMAGE.  Built with `--features ottomarcher` from inside the ottomarcher crate (these `use crate::…`
//! paths are the reference's modules).  Two of the reference's types keep their fields private, so their conversions cannot be
//! written from outside: the snippets to paste next to them are at the end of this file.
use crate::materials::{Material, MaterialType};
use crate::math::mat3x3::Mat3x3;
use crate::math::mat4x4::Mat4x4;
use crate::math::vec3::Vec3;
use crate::math::vec4::Vec4;
use crate::marched::{MarchedBox, MarchedSphere, MarchedTorus};
use crate::render_thread::{Pixel, Stats};
use crate::traced::{Barycentric, Cube, InfinitePlane, Sphere};
use crate::utils::BloomFilter;
use super::*;

fn v3(v: &Vec3) -> [f32; 3] { v.e }
fn v4(v: &Vec4) -> [f32; 4] { v.e }
/// Mat4x4 { e: [Vec4; 4] } -> 16 floats, row-major
pub fn rows(m: &Mat4x4) -> [f32; 16] {
    let mut o = [0f32; 16];
    for r in 0..4 { for c in 0..4 { o[4 * r + c] = m.e[r].e[c]; } }
    o
}
fn rows3(m: &Mat3x3) -> [f32; 9] {
    let mut o = [0f32; 9];
    for r in 0..3 { for c in 0..3 { o[3 * r + c] = m.e[r].e[c]; } }
    o
}

impl From<&Material> for om_material {   // materials.rs:18-23
    fn from(m: &Material) -> Self {
        let mat_type = match m.mat_type { MaterialType::LAMBERTIAN => OM_MAT_LAMBERTIAN, MaterialType::METAL => OM_MAT_METAL, MaterialType::DIELECTRIC => OM_MAT_DIELECTRIC };
        om_material { albedo: v3(&m.albedo), fuzz: m.fuzz, ior: m.ior, mat_type: mat_type }
    }
}
impl From<&Sphere> for om_sphere {       // traced.rs:13-19 (the field is spelled m_word_to_local there)
    fn from(s: &Sphere) -> Self {
        om_sphere { m_local_to_world: rows(&s.m_local_to_world), m_world_to_local: rows(&s.m_word_to_local), material: (&s.material).into() }
    }
}
impl From<&Cube> for om_cube {           // traced.rs:229-235
    fn from(s: &Cube) -> Self {
        om_cube { m_local_to_world: rows(&s.m_local_to_world), m_world_to_local: rows(&s.m_word_to_local), material: (&s.material).into() }
    }
}
impl From<&InfinitePlane> for om_infinite_plane {   // traced.rs:77-82
    fn from(p: &InfinitePlane) -> Self { om_infinite_plane { center: v3(&p.center), normal: v3(&p.normal), material: (&p.material).into() } }
}
impl<const BT: usize> From<&Barycentric<BT>> for om_barycentric {   // traced.rs:118-131; Parallelogram = 0 goes to om_scene_add_parallelograms, Triangle = 1 to _triangles
    fn from(b: &Barycentric<BT>) -> Self {
        om_barycentric { origin: v3(&b.origin), u: v3(&b.u), u_length: b.u_length, v: v3(&b.v), v_length: b.v_length, uxv: v3(&b.uxv), uxvxu: v3(&b.uxvxu),
                         base_inv: rows3(&b.base_inv), v_in_base: v3(&b.v_in_base), material: (&b.material).into() }
    }
}
impl From<&MarchedSphere> for om_marched_sphere {   // marched.rs:49-54
    fn from(s: &MarchedSphere) -> Self { om_marched_sphere { center: v3(&s.center), radius: s.radius, material: (&s.material).into() } }
}
impl From<&MarchedBox> for om_marched_box {         // marched.rs:78-83
    fn from(s: &MarchedBox) -> Self { om_marched_box { center: v3(&s.center), sizes: v3(&s.sizes), material: (&s.material).into() } }
}
impl From<&MarchedTorus> for om_marched_torus {     // marched.rs:104-113
    fn from(s: &MarchedTorus) -> Self {
        om_marched_torus { m_local_to_world_translate_rotate: rows(&s.m_local_to_world_translate_rotate),
                           m_world_to_local_translate_rotate: rows(&s.m_world_to_local_translate_rotate),
                           m_local_to_world_scale: v4(&s.m_local_to_world_scale), m_world_to_local_scale: v4(&s.m_world_to_local_scale),
                           sizes: v3(&s.sizes), material: (&s.material).into() }
    }
}
/// om_pixel -> Pixel (render_thread.rs:9-51): what draw_to_sdl reads (main.rs:228-232,377,386,402,432)
impl From<&om_pixel> for Pixel {
    fn from(p: &om_pixel) -> Self {
        Pixel { c: Vec3 { e: [0.0; 3] },
                stats: Stats { n: p.n, sum: Vec3 { e: p.sum }, color: (p.color[0], p.color[1], p.color[2]), bad_avgs: p.bad_avgs, avg_depth: p.avg_depth,
                               bloom_filter: BloomFilter { state: p.bloom } } }
    }
}

// ---- to paste into the reference, next to the private fields ---------------------------------------------------------------------
// camera.rs, inside `impl Camera` (camera.rs:9-29: origin .. lens_radius are private):
//     pub fn to_om(&self) -> ottomarcher_cuda_sys::om_camera {
//         ottomarcher_cuda_sys::om_camera { origin: self.origin.e, horizontal: self.horizontal.e, vertical: self.vertical.e,
//             lower_left_corner: self.lower_left_corner.e, u_of_plane: self.u_of_plane.e, v_of_plane: self.v_of_plane.e, w_of_plane: self.w_of_plane.e,
//             lens_radius: self.lens_radius, aspect_ratio: self.aspect_ratio, focus_dist: self.focus_dist,
//             viewport_width: self.viewport_width, viewport_height: self.viewport_height }
//     }
// hits.rs, inside `impl HittableList` of the `hittable_list!` macro (hits.rs:37-43: the eight Vec<T> are private):
//     pub fn to_om(&self) -> ottomarcher_cuda_sys::SceneLists {
//         ottomarcher_cuda_sys::SceneLists {
//             spheres: self.spheres.iter().map(Into::into).collect(), cubes: self.cubes.iter().map(Into::into).collect(),
//             triangles: self.triangles.iter().map(Into::into).collect(), infinite_planes: self.infinite_planes.iter().map(Into::into).collect(),
//             parallelograms: self.parallelograms.iter().map(Into::into).collect(), marched_spheres: self.marched_spheres.iter().map(Into::into).collect(),
//             marched_boxes: self.marched_boxes.iter().map(Into::into).collect(), marched_tori: self.marched_torus.iter().map(Into::into).collect(),
//             marched_objects: Vec::new(),   // Arc<dyn Marched> that also implement DeviceMarched: push their describe()
//         }
//     }
