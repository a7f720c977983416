// Host-side mirror of the reference's construction code: the parts of the reference that stay
// on the CPU in front of render_thread::render (matrix algebra, object constructors, camera,
// scene builders, bloom hash table, jitter table).  Plain row-major float[16] matrices.
#pragma once
#include <stdint.h>

#include <vector>

#include "../../include/ottomarcher.h"

namespace omh {

// ---- Philox4x32-10 (same contract as the device generator, see om_device.cuh) ----------------
void philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
struct Rng {
    uint32_t key[2], a, b, stream, draw, buf[4];
    Rng(uint64_t seed, uint32_t a_, uint32_t b_, uint32_t stream_);
    uint32_t next_u32();
    float rand();                         // utils.rs:25
    float rand_range(float mn, float mx); // utils.rs:26
};
enum { STREAM_PATH = 0, STREAM_JITTER = 1, STREAM_SCENE = 2 };

// ---- mat4x4.rs / mat3x3.rs -------------------------------------------------------------------
void mat4_identity(float o[16]);
void mat4_translate(float o[16], float x, float y, float z);   // mat4x4.rs:88-93
void mat4_scale(float o[16], float x, float y, float z);       // :94-99
void mat4_rotate_x(float o[16], float f);                      // :100-107
void mat4_rotate_y(float o[16], float f);                      // :108-115
void mat4_rotate_z(float o[16], float f);                      // :116-123
void mat4_mul(float o[16], const float a[16], const float b[16]);          // dot_mat :66-72
void mat4_fast_homogenous_inverse(float o[16], const float m[16]);         // :59-64
float mat3_determinant(const float m[9]);                                  // mat3x3.rs:52-68 (Kahan)
void mat3_inverse(float o[9], const float m[9]);                           // mat3x3.rs:69-81

// ---- constructors ----------------------------------------------------------------------------
void camera_new(om_camera* out, const float lookfrom[3], const float lookat[3], const float vup[3], float vfov_deg,
                float aspect, float aperture, float focus_dist);
om_material lambertian(float r, float g, float b);
om_material metal(float r, float g, float b, float fuzz);
om_material dielectric(float ior);
void sphere_new(om_sphere* out, const float m[16], const om_material& mat);
void sphere_new_with_radius(om_sphere* out, const float o[3], float r, const om_material& mat);
void cube_new(om_cube* out, const float m[16], const om_material& mat);
void cube_new_with_length(om_cube* out, const float o[3], float len, const om_material& mat);
void infinite_plane_new(om_infinite_plane* out, const float c[3], const float n[3], const om_material& mat);
void barycentric_new3points(om_barycentric* out, const float o[3], const float up[3], const float vp[3], const om_material& mat);
void marched_torus_new(om_marched_torus* out, const float m[16], const float sizes[3], const om_material& mat);

// ---- utils.rs:45-108 -------------------------------------------------------------------------
uint64_t scramble(uint64_t id);
uint64_t bloom_get_hash(uint64_t id);
void jitter_table(uint32_t spp, uint64_t seed, std::vector<uint8_t>& out);  // render_thread.rs:164-174

// ---- scene lists + generators ----------------------------------------------------------------
struct SceneLists {  // HittableList, hits.rs:37-43
    std::vector<om_sphere> spheres;
    std::vector<om_cube> cubes;
    std::vector<om_barycentric> triangles;
    std::vector<om_infinite_plane> infinite_planes;
    std::vector<om_barycentric> parallelograms;
    std::vector<om_marched_sphere> marched_spheres;
    std::vector<om_marched_box> marched_boxes;
    std::vector<om_marched_torus> marched_tori;
    // marched_objects: Vec<Arc<dyn Marched>> as data (om_dyn_marched); object i's program = dyn_code[dyn_code_first[i] .. dyn_code_first[i + 1])
    std::vector<om_dyn_marched> dyn_marched;
    std::vector<uint32_t> dyn_code, dyn_code_first;
    // traced_objects: Vec<Arc<dyn Traced>> as data (om_dyn_traced: quadrics in a local frame), scanned after the parallelograms
    std::vector<om_dyn_traced> dyn_traced;
};
// "" when the description is usable (see om_scene_add_dyn_traced), else what is wrong with it
const char* check_dyn_traced(const om_dyn_traced& o);
// "" when the program is well-formed (see om_scene_add_dyn_marched), else what is wrong with it
const char* check_sdf_program(const uint32_t* code, size_t n_words);
bool generate(SceneLists& s, uint32_t kind, uint64_t seed, uint64_t n_param);

}  // namespace omh
