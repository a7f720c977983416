// ORACLE — TEST INFRASTRUCTURE ONLY.  C entry points (ctypes) over the CPU restatement of the
// reference in ref_*.hpp.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs load this library.  PARITY UNPINNED (see ref_math.hpp).
//
// The POD scene/camera/pixel structs are the ones of include/ottomarcher.h so that the very
// same bytes can be handed to the oracle and to the CUDA library ("the scene crosses the
// boundary as data").
#include <cstring>

#include "../include/ottomarcher.h"
#include "ref_render.hpp"
#include "ref_view.hpp"

using namespace ref;

namespace {
Vec3 v3(const float* p) { return Vec3(p[0], p[1], p[2]); }
void put3(float* p, const Vec3& v) { p[0] = v.e[0]; p[1] = v.e[1]; p[2] = v.e[2]; }
Material mat_in(const om_material& m) { Material o; o.albedo = v3(m.albedo); o.fuzz = m.fuzz; o.ior = m.ior; o.mat_type = m.mat_type; return o; }
om_material mat_out(const Material& m) { om_material o; put3(o.albedo, m.albedo); o.fuzz = m.fuzz; o.ior = m.ior; o.mat_type = m.mat_type; return o; }

Sphere sphere_in(const om_sphere& s) { Sphere o; o.m_local_to_world = Mat4x4::from_rows16(s.m_local_to_world); o.m_word_to_local = Mat4x4::from_rows16(s.m_world_to_local); o.material = mat_in(s.material); return o; }
om_sphere sphere_out(const Sphere& s) { om_sphere o; s.m_local_to_world.to_rows16(o.m_local_to_world); s.m_word_to_local.to_rows16(o.m_world_to_local); o.material = mat_out(s.material); return o; }
Cube cube_in(const om_cube& s) { Cube o; o.m_local_to_world = Mat4x4::from_rows16(s.m_local_to_world); o.m_word_to_local = Mat4x4::from_rows16(s.m_world_to_local); o.material = mat_in(s.material); return o; }
om_cube cube_out(const Cube& s) { om_cube o; s.m_local_to_world.to_rows16(o.m_local_to_world); s.m_word_to_local.to_rows16(o.m_world_to_local); o.material = mat_out(s.material); return o; }
InfinitePlane plane_in(const om_infinite_plane& s) { InfinitePlane o; o.center = v3(s.center); o.normal = v3(s.normal); o.material = mat_in(s.material); return o; }
Barycentric bary_in(const om_barycentric& s) {
    Barycentric o;
    o.origin = v3(s.origin); o.u = v3(s.u); o.u_length = s.u_length; o.v = v3(s.v); o.v_length = s.v_length;
    o.uxv = v3(s.uxv); o.uxvxu = v3(s.uxvxu);
    for (int r = 0; r < 3; r++) o.base_inv.e[r] = Vec3(s.base_inv[3 * r], s.base_inv[3 * r + 1], s.base_inv[3 * r + 2]);
    o.v_in_base = v3(s.v_in_base); o.material = mat_in(s.material);
    return o;
}
om_barycentric bary_out(const Barycentric& s) {
    om_barycentric o;
    put3(o.origin, s.origin); put3(o.u, s.u); o.u_length = s.u_length; put3(o.v, s.v); o.v_length = s.v_length;
    put3(o.uxv, s.uxv); put3(o.uxvxu, s.uxvxu);
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) o.base_inv[3 * r + c] = s.base_inv.e[r].e[c];
    put3(o.v_in_base, s.v_in_base); o.material = mat_out(s.material);
    return o;
}
MarchedSphere msphere_in(const om_marched_sphere& s) { MarchedSphere o; o.center = v3(s.center); o.radius = s.radius; o.material = mat_in(s.material); return o; }
MarchedBox mbox_in(const om_marched_box& s) { MarchedBox o; o.center = v3(s.center); o.sizes = v3(s.sizes); o.material = mat_in(s.material); return o; }
MarchedTorus mtorus_in(const om_marched_torus& s) {
    MarchedTorus o;
    o.m_local_to_world_translate_rotate = Mat4x4::from_rows16(s.m_local_to_world_translate_rotate);
    o.m_world_to_local_translate_rotate = Mat4x4::from_rows16(s.m_world_to_local_translate_rotate);
    o.m_local_to_world_scale = Vec4(s.m_local_to_world_scale[0], s.m_local_to_world_scale[1], s.m_local_to_world_scale[2], s.m_local_to_world_scale[3]);
    o.m_world_to_local_scale = Vec4(s.m_world_to_local_scale[0], s.m_world_to_local_scale[1], s.m_world_to_local_scale[2], s.m_world_to_local_scale[3]);
    o.sizes = v3(s.sizes); o.material = mat_in(s.material);
    return o;
}
om_marched_torus mtorus_out(const MarchedTorus& s) {
    om_marched_torus o;
    s.m_local_to_world_translate_rotate.to_rows16(o.m_local_to_world_translate_rotate);
    s.m_world_to_local_translate_rotate.to_rows16(o.m_world_to_local_translate_rotate);
    for (int i = 0; i < 4; i++) { o.m_local_to_world_scale[i] = s.m_local_to_world_scale.e[i]; o.m_world_to_local_scale[i] = s.m_world_to_local_scale.e[i]; }
    put3(o.sizes, s.sizes); o.material = mat_out(s.material);
    return o;
}
Camera cam_in(const om_camera& c) {
    Camera o;
    o.origin = v3(c.origin); o.horizontal = v3(c.horizontal); o.vertical = v3(c.vertical); o.lower_left_corner = v3(c.lower_left_corner);
    o.u_of_plane = v3(c.u_of_plane); o.v_of_plane = v3(c.v_of_plane); o.w_of_plane = v3(c.w_of_plane);
    o.lens_radius = c.lens_radius; o.aspect_ratio = c.aspect_ratio; o.focus_dist = c.focus_dist;
    o.viewport_width = c.viewport_width; o.viewport_height = c.viewport_height;
    return o;
}
om_camera cam_out(const Camera& c) {
    om_camera o;
    put3(o.origin, c.origin); put3(o.horizontal, c.horizontal); put3(o.vertical, c.vertical); put3(o.lower_left_corner, c.lower_left_corner);
    put3(o.u_of_plane, c.u_of_plane); put3(o.v_of_plane, c.v_of_plane); put3(o.w_of_plane, c.w_of_plane);
    o.lens_radius = c.lens_radius; o.aspect_ratio = c.aspect_ratio; o.focus_dist = c.focus_dist;
    o.viewport_width = c.viewport_width; o.viewport_height = c.viewport_height;
    return o;
}
Stats stats_in(const om_pixel& p) {
    Stats s; s.n = p.n; s.sum = v3(p.sum); s.color[0] = p.color[0]; s.color[1] = p.color[1]; s.color[2] = p.color[2];
    s.bad_avgs = p.bad_avgs; s.avg_depth = p.avg_depth; s.bloom = p.bloom;
    return s;
}
om_pixel stats_out(const Stats& s) {
    om_pixel p; std::memset(&p, 0, sizeof p);
    put3(p.sum, s.sum); p.n = s.n; p.avg_depth = s.avg_depth; p.bad_avgs = s.bad_avgs; p.bloom = s.bloom;
    p.color[0] = s.color[0]; p.color[1] = s.color[1]; p.color[2] = s.color[2];
    return p;
}
RenderParams params_in(const om_render_params& p) {
    RenderParams r;
    r.width = p.width; r.height = p.height; r.spp_max = p.spp_max; r.max_depth = p.max_depth; r.tmin = p.tmin; r.tmax = p.tmax;
    r.seed = p.seed; r.adaptive = (p.flags & OM_RENDER_ADAPTIVE) != 0;
    r.sample_base = p.sample_base;
    r.sample_count = p.sample_count ? p.sample_count : (p.spp_max - p.sample_base);
    r.part_rank = p.part_rank; r.part_world = p.part_world ? p.part_world : 1;
    return r;
}
}  // namespace

struct omo_scene { World w; };

extern "C" {

#define OMO_API __attribute__((visibility("default")))

OMO_API omo_scene* omo_scene_create() { return new omo_scene(); }
OMO_API void omo_scene_destroy(omo_scene* s) { delete s; }
OMO_API void omo_scene_add_spheres(omo_scene* s, const om_sphere* a, size_t n) { for (size_t i = 0; i < n; i++) s->w.spheres.push_back(sphere_in(a[i])); }
OMO_API void omo_scene_add_cubes(omo_scene* s, const om_cube* a, size_t n) { for (size_t i = 0; i < n; i++) s->w.cubes.push_back(cube_in(a[i])); }
OMO_API void omo_scene_add_triangles(omo_scene* s, const om_barycentric* a, size_t n) { for (size_t i = 0; i < n; i++) s->w.triangles.push_back(bary_in(a[i])); }
OMO_API void omo_scene_add_infinite_planes(omo_scene* s, const om_infinite_plane* a, size_t n) { for (size_t i = 0; i < n; i++) s->w.infinite_planes.push_back(plane_in(a[i])); }
OMO_API void omo_scene_add_parallelograms(omo_scene* s, const om_barycentric* a, size_t n) { for (size_t i = 0; i < n; i++) s->w.parallelograms.push_back(bary_in(a[i])); }
OMO_API void omo_scene_add_marched_spheres(omo_scene* s, const om_marched_sphere* a, size_t n) { for (size_t i = 0; i < n; i++) s->w.marched_spheres.push_back(msphere_in(a[i])); }
OMO_API void omo_scene_add_marched_boxes(omo_scene* s, const om_marched_box* a, size_t n) { for (size_t i = 0; i < n; i++) s->w.marched_boxes.push_back(mbox_in(a[i])); }
OMO_API void omo_scene_add_marched_tori(omo_scene* s, const om_marched_torus* a, size_t n) { for (size_t i = 0; i < n; i++) s->w.marched_torus.push_back(mtorus_in(a[i])); }
OMO_API void omo_scene_add_dyn_marched(omo_scene* s, const om_dyn_marched* a, const uint32_t* code, size_t n_words) {
    DynMarched o;
    o.m_world_to_local = Mat4x4::from_rows16(a->m_world_to_local); o.m_local_to_world = Mat4x4::from_rows16(a->m_local_to_world);
    o.world_f = a->world_f; o.code.assign(code, code + n_words); o.material = mat_in(a->material);
    s->w.marched_objects.push_back(o);
}
OMO_API void omo_scene_add_dyn_traced(omo_scene* s, const om_dyn_traced* a, size_t n) {
    for (size_t i = 0; i < n; i++) {
        DynTraced o;
        o.m_local_to_world = Mat4x4::from_rows16(a[i].m_local_to_world);
        o.m_word_to_local = Mat4x4::from_rows16(a[i].m_world_to_local);
        for (int k = 0; k < 10; k++) o.q[k] = a[i].q[k];
        o.clip_lo = v3(a[i].clip_lo); o.clip_hi = v3(a[i].clip_hi);
        o.material = mat_in(a[i].material);
        s->w.traced_objects.push_back(o);
    }
}
OMO_API void omo_scene_counts(const omo_scene* s, uint64_t c[8]) {
    c[0] = s->w.spheres.size(); c[1] = s->w.cubes.size(); c[2] = s->w.triangles.size(); c[3] = s->w.infinite_planes.size();
    c[4] = s->w.parallelograms.size(); c[5] = s->w.marched_spheres.size(); c[6] = s->w.marched_boxes.size(); c[7] = s->w.marched_torus.size();
}
// the oracle's own restatement of main.rs:37-110 — used to check the product's generators
OMO_API void omo_scene_random(omo_scene* s, uint64_t seed) { random_scene(s->w, seed); }
OMO_API void omo_scene_basic(omo_scene* s) { basic_scene(s->w); }
OMO_API void omo_scene_get_spheres(const omo_scene* s, om_sphere* out) { for (size_t i = 0; i < s->w.spheres.size(); i++) out[i] = sphere_out(s->w.spheres[i]); }
OMO_API void omo_scene_get_cubes(const omo_scene* s, om_cube* out) { for (size_t i = 0; i < s->w.cubes.size(); i++) out[i] = cube_out(s->w.cubes[i]); }
OMO_API void omo_scene_get_triangles(const omo_scene* s, om_barycentric* out) { for (size_t i = 0; i < s->w.triangles.size(); i++) out[i] = bary_out(s->w.triangles[i]); }
OMO_API void omo_scene_get_parallelograms(const omo_scene* s, om_barycentric* out) { for (size_t i = 0; i < s->w.parallelograms.size(); i++) out[i] = bary_out(s->w.parallelograms[i]); }
OMO_API void omo_scene_get_marched_tori(const omo_scene* s, om_marched_torus* out) { for (size_t i = 0; i < s->w.marched_torus.size(); i++) out[i] = mtorus_out(s->w.marched_torus[i]); }

// render(): `pixels` is in/out (Vec<Pixel>), so passes can be chained like the reference's outer loop.
// counters (nullable): segments, traced_tests, sdf_evals, normal_evals, march_steps
OMO_API int omo_render(const omo_scene* s, const om_camera* cam, const om_render_params* params, om_pixel* pixels, double* sumsq,
                       uint32_t spp_pass, uint32_t nthreads, uint64_t* credited, uint64_t* traced, uint64_t* counters) {
    RenderParams p = params_in(*params);
    if (p.width < 2 || p.height < 2 || p.spp_max == 0) return 1;
    size_t npx = (size_t)p.width * p.height;
    std::vector<Stats> st(npx);
    for (size_t i = 0; i < npx; i++) st[i] = stats_in(pixels[i]);
    if (spp_pass == 0) spp_pass = p.sample_count;
    World::Counters cn;
    uint64_t cr = 0, tr = 0;
    render(cam_in(*cam), s->w, p, st.data(), sumsq, spp_pass, nthreads, &cr, &tr, counters ? &cn : nullptr);
    for (size_t i = 0; i < npx; i++) pixels[i] = stats_out(st[i]);
    if (credited) *credited += cr;
    if (traced) *traced += tr;
    if (counters) { counters[0] += cn.segments; counters[1] += cn.traced_tests; counters[2] += cn.sdf_evals; counters[3] += cn.normal_evals; counters[4] += cn.march_steps; }
    return 0;
}

// primary rays: mode 0 = pixel centre, no lens; mode 1 = the ray of sample `sample_index`
OMO_API void omo_primary_rays(const om_camera* cam, const om_render_params* params, uint32_t mode, uint32_t sample_index, float* origins, float* dirs) {
    RenderParams p = params_in(*params);
    Camera c = cam_in(*cam);
    std::vector<uint8_t> jit = make_jitter_table(p.spp_max, p.seed);
    uint32_t npx = p.width * p.height;
    for (uint32_t i = 0; i < npx; i++) {
        Ray r;
        if (mode == 0) r = center_ray(c, p, i);
        else { Rng g(p.seed, i, sample_index, STREAM_PATH); r = primary_ray(c, p, jit, i, sample_index, g); }
        put3(origins + 3 * (size_t)i, r.orig); put3(dirs + 3 * (size_t)i, r.dir);
    }
}
OMO_API void omo_primary(const omo_scene* s, const om_camera* cam, const om_render_params* params, uint32_t mode, uint32_t sample_index, uint64_t* ids, float* t) {
    RenderParams p = params_in(*params);
    Camera c = cam_in(*cam);
    std::vector<uint8_t> jit = make_jitter_table(p.spp_max, p.seed);
    uint32_t npx = p.width * p.height;
    for (uint32_t i = 0; i < npx; i++) {
        Ray r;
        if (mode == 0) r = center_ray(c, p, i);
        else { Rng g(p.seed, i, sample_index, STREAM_PATH); r = primary_ray(c, p, jit, i, sample_index, g); }
        HitRecord hr;
        if (s->w.hit(r, p.tmin, p.tmax, hr)) { ids[i] = hr.obj_id; t[i] = hr.t; } else { ids[i] = 0; t[i] = INF; }
    }
}
// FrozenHittableList::hit for arbitrary rays; dirs are normalised like Ray::new
OMO_API void omo_trace_rays(const omo_scene* s, const float* origins, const float* dirs, size_t n, float tmin, float tmax,
                            uint64_t* ids, float* t, float* points, float* normals) {
    for (size_t i = 0; i < n; i++) {
        Ray r = Ray::make(v3(origins + 3 * i), v3(dirs + 3 * i));
        HitRecord hr;
        bool h = s->w.hit(r, tmin, tmax, hr);
        if (ids) ids[i] = h ? hr.obj_id : 0;
        if (t) t[i] = h ? hr.t : INF;
        if (points) put3(points + 3 * i, h ? hr.point : Vec3());
        if (normals) put3(normals + 3 * i, h ? hr.normal : Vec3());
    }
}
// one full sample (ray_color) for given (pixel, sample): colour, depth, id — for fine-grained parity
OMO_API void omo_sample(const omo_scene* s, const om_camera* cam, const om_render_params* params, uint32_t pixel, uint32_t sample, float color[3], float* depth, uint64_t* id) {
    RenderParams p = params_in(*params);
    Camera c = cam_in(*cam);
    std::vector<uint8_t> jit = make_jitter_table(p.spp_max, p.seed);
    Rng g(p.seed, pixel, sample, STREAM_PATH);
    Ray r = primary_ray(c, p, jit, pixel, sample, g);
    SampleResult sr = ray_color(r, s->w, p.max_depth, p.tmin, p.tmax, g);
    put3(color, sr.color); *depth = sr.depth; *id = sr.obj_id;
}

// ---- constructors (restated) -------------------------------------------------------------------
OMO_API void omo_camera_new(om_camera* out, const float* lookfrom, const float* lookat, const float* vup, float vfov, float aspect, float aperture, float focus_dist) {
    *out = cam_out(Camera::make(v3(lookfrom), v3(lookat), v3(vup), vfov, aspect, aperture, focus_dist));
}
OMO_API void omo_sphere_new(om_sphere* out, const float* m, const om_material* mat) { *out = sphere_out(Sphere::make(Mat4x4::from_rows16(m), mat_in(*mat))); }
OMO_API void omo_sphere_new_with_radius(om_sphere* out, const float* o, float r, const om_material* mat) { *out = sphere_out(Sphere::new_with_radius(v3(o), r, mat_in(*mat))); }
OMO_API void omo_cube_new(om_cube* out, const float* m, const om_material* mat) { *out = cube_out(Cube::make(Mat4x4::from_rows16(m), mat_in(*mat))); }
OMO_API void omo_cube_new_with_length(om_cube* out, const float* o, float len, const om_material* mat) { *out = cube_out(Cube::new_with_length(v3(o), len, mat_in(*mat))); }
OMO_API void omo_infinite_plane_new(om_infinite_plane* out, const float* c, const float* n, const om_material* mat) {
    InfinitePlane p = InfinitePlane::make(v3(c), v3(n), mat_in(*mat));
    put3(out->center, p.center); put3(out->normal, p.normal); out->material = mat_out(p.material);
}
OMO_API void omo_barycentric_new3points(om_barycentric* out, const float* o, const float* u, const float* v, const om_material* mat) {
    *out = bary_out(Barycentric::new3points(v3(o), v3(u), v3(v), mat_in(*mat)));
}
OMO_API void omo_marched_torus_new(om_marched_torus* out, const float* m, const float* sizes, const om_material* mat) {
    *out = mtorus_out(MarchedTorus::make(Mat4x4::from_rows16(m), v3(sizes), mat_in(*mat)));
}
OMO_API void omo_mat4_translate(float* out, float x, float y, float z) { Mat4x4::new_translate(Vec3(x, y, z)).to_rows16(out); }
OMO_API void omo_mat4_scale(float* out, float x, float y, float z) { Mat4x4::new_scale(Vec3(x, y, z)).to_rows16(out); }
OMO_API void omo_mat4_rotate_x(float* out, float f) { Mat4x4::new_rotate_x(f).to_rows16(out); }
OMO_API void omo_mat4_rotate_y(float* out, float f) { Mat4x4::new_rotate_y(f).to_rows16(out); }
OMO_API void omo_mat4_rotate_z(float* out, float f) { Mat4x4::new_rotate_z(f).to_rows16(out); }
OMO_API void omo_mat4_mul(float* out, const float* a, const float* b) { Mat4x4::from_rows16(a).dot_mat(Mat4x4::from_rows16(b)).to_rows16(out); }
OMO_API void omo_mat4_fast_homogenous_inverse(float* out, const float* m) { Mat4x4::from_rows16(m).fast_homogenous_inverse().to_rows16(out); }

// ---- unit hooks for known-answer tests ------------------------------------------------------------
OMO_API void omo_philox(const uint32_t* ctr, const uint32_t* key, uint32_t* out) { philox4x32_10(ctr, key, out); }
OMO_API void omo_rand_stream(uint64_t seed, uint32_t a, uint32_t b, uint32_t stream, uint32_t n, float* out) { Rng g(seed, a, b, stream); for (uint32_t i = 0; i < n; i++) out[i] = g.rand(); }
OMO_API void omo_jitter_table(uint32_t spp, uint64_t seed, uint8_t* out) { auto t = make_jitter_table(spp, seed); std::memcpy(out, t.data(), spp); }
OMO_API uint64_t omo_scramble(uint64_t id) { return scramble(id); }
OMO_API uint64_t omo_bloom_hash(uint64_t id) { return bloom_get_hash(id); }
OMO_API void omo_u64_to_color(uint64_t id, uint8_t* out) { u64_to_color(id, out); }
OMO_API void omo_tone_map(const float* rgb, uint8_t* out) { to_u8x3(normalize_color(v3(rgb)), out); }
OMO_API int omo_stats_add(om_pixel* px, const float* color, float depth, uint64_t id) { Stats s = stats_in(*px); bool d = s.add(v3(color), depth, id); *px = stats_out(s); return d ? 1 : 0; }
OMO_API void omo_reflect(const float* v, const float* n, float* out) { put3(out, reflect(v3(v), v3(n))); }
OMO_API void omo_refract(const float* uv, const float* n, float eta, float* out) { put3(out, refract(v3(uv), v3(n), eta)); }
OMO_API float omo_reflectance(float cos, float ref_idx) { return reflectance(cos, ref_idx); }
// kind 0 sphere / 1 box / 2 torus; obj points at the matching om_marched_* struct
OMO_API float omo_marched_sdf(uint32_t kind, const void* obj, const float* p) {
    if (kind == 0) return marched_sdf(msphere_in(*(const om_marched_sphere*)obj), v3(p));
    if (kind == 1) return marched_sdf(mbox_in(*(const om_marched_box*)obj), v3(p));
    return marched_sdf(mtorus_in(*(const om_marched_torus*)obj), v3(p));
}
OMO_API void omo_marched_normal(uint32_t kind, const void* obj, const float* p, float* out) {
    if (kind == 0) put3(out, msphere_in(*(const om_marched_sphere*)obj).get_outward_normal(v3(p)));
    else if (kind == 1) put3(out, mbox_in(*(const om_marched_box*)obj).get_outward_normal(v3(p)));
    else put3(out, mtorus_in(*(const om_marched_torus*)obj).get_outward_normal(v3(p)));
}
// scatter one ray off a synthetic hit record with the Philox stream (seed, a, b)
OMO_API void omo_scatter(const om_material* m, const float* in_orig, const float* in_dir, const float* point, const float* normal,
                         uint64_t seed, uint32_t a, uint32_t b, float* att, float* out_orig, float* out_dir) {
    Ray r; r.orig = v3(in_orig); r.dir = v3(in_dir);
    HitRecord hr; hr.point = v3(point); hr.normal = v3(normal); hr.material = mat_in(*m); hr.t = 0.f; hr.obj_id = 0;
    Rng g(seed, a, b, STREAM_PATH);
    ScatterResult s = scatter(hr.material, r, hr, g);
    put3(att, s.attenuation); put3(out_orig, s.ray.orig); put3(out_dir, s.ray.dir);
}
OMO_API void omo_view(int mode, const om_pixel* px, int W, int H, uint8_t* rgb) { view(mode, px, W, H, rgb); }
OMO_API uint32_t omo_hardware_threads() { unsigned n = std::thread::hardware_concurrency(); return n ? n : 1; }
OMO_API const char* omo_version() { return "om_oracle 0.1 (C++ restatement of octaviogarcia/RaytracingOneWeekend; parity unpinned)"; }

}  // extern "C"
