// ottomarcher.hpp — C++17 host side above the C ABI (ottomarcher.h), mirroring the reference's own
// interface for the render path so that a host program reads like the reference's src/main.rs:
//
//     HittableList world;                                         // hits.rs:72
//     world += Sphere::new_with_radius({0,-1000,0}, 1000, Material::new_lambertian({.5,.5,.5}));   // main.rs:41
//     FrozenHittableList frozen = world.freeze(&camera);           // hits.rs:87-89
//     std::vector<render_thread::Pixel> pixels(w * h);             // main.rs:192
//     std::atomic<uint64_t> samples_atom{0};                       // main.rs:147
//     render_thread::render(camera, frozen, max_depth, tmin, tmax, samples_per_pixel, w, h,
//                           pixels, tid, n_threads, samples_atom); // render_thread.rs:145-147 / main.rs:207-213
//
// The reference is Rust and cannot be compiled in this image; this header is the host language
// layer that CAN be compiled here (g++), used by harness/om_main.cpp and the C++ tests.  Rust
// associated functions called `new` are spelled `new_` (C++ keyword).  Where the reference panics,
// these functions throw ottomarcher::Error carrying om_last_error().  Header-only; link with
// libottomarcher_cuda.so.  There is no CPU fallback: without a CUDA device render() throws.
#ifndef OTTOMARCHER_HPP
#define OTTOMARCHER_HPP

#include <algorithm>
#include <cmath>
#include <cstring>
#include <atomic>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "ottomarcher.h"

namespace ottomarcher {

struct Error : std::runtime_error {
    om_status status;
    Error(om_status s, const std::string& what) : std::runtime_error(what), status(s) {}
};
inline void check(om_status s) {
    if (s != OM_OK) throw Error(s, om_last_error());
}

// ---- math/vec3.rs: only what scene construction needs ------------------------------------------
struct Vec3 {
    float e[3];
    Vec3() : e{0.f, 0.f, 0.f} {}
    Vec3(float x, float y, float z) : e{x, y, z} {}
    static Vec3 new_(float x, float y, float z) { return Vec3(x, y, z); }  // vec3.rs:17
    float x() const { return e[0]; }
    float y() const { return e[1]; }
    float z() const { return e[2]; }
};
using Point3 = Vec3;  // vec3.rs:13
using Color = Vec3;   // vec3.rs:14

// ---- math/mat4x4.rs: row-major 4x4, `a ^ b` is the matrix product (mat4x4.rs:167-209) ------------
struct Mat4x4 {
    float m[16];
    static Mat4x4 identity() { Mat4x4 r; om_mat4_identity(r.m); return r; }
    static Mat4x4 translate(float x, float y, float z) { Mat4x4 r; om_mat4_translate(r.m, x, y, z); return r; }  // m4x4!(TR x,y,z)
    static Mat4x4 scale(float x, float y, float z) { Mat4x4 r; om_mat4_scale(r.m, x, y, z); return r; }          // m4x4!(SC x,y,z)
    static Mat4x4 rotate_x(float f) { Mat4x4 r; om_mat4_rotate_x(r.m, f); return r; }                           // m4x4!(RX f)
    static Mat4x4 rotate_y(float f) { Mat4x4 r; om_mat4_rotate_y(r.m, f); return r; }
    static Mat4x4 rotate_z(float f) { Mat4x4 r; om_mat4_rotate_z(r.m, f); return r; }
    Mat4x4 operator^(const Mat4x4& b) const { Mat4x4 r; om_mat4_mul(r.m, m, b.m); return r; }
    Mat4x4 fast_homogenous_inverse() const { Mat4x4 r; om_mat4_fast_homogenous_inverse(r.m, m); return r; }      // mat4x4.rs:59
};

// ---- materials.rs:11-37 -------------------------------------------------------------------------------
struct Material : om_material {
    static Material new_lambertian(Color albedo) { Material m; om_material_lambertian(&m, albedo.e); return m; }
    static Material new_metal(Color albedo) { Material m; om_material_metal(&m, albedo.e, 0.f); return m; }                    // :29
    static Material new_metal_fuzz(Color albedo, float fuzz) { Material m; om_material_metal(&m, albedo.e, fuzz); return m; }  // :32 (no clamp)
    static Material new_dielectric(float ior) { Material m; om_material_dielectric(&m, ior); return m; }
};

// ---- traced.rs ----------------------------------------------------------------------------------------
struct Sphere : om_sphere {
    static Sphere new_(const Mat4x4& local_to_world, const Material& mat) { Sphere s; om_sphere_new(&s, local_to_world.m, &mat); return s; }       // traced.rs:22
    static Sphere new_with_radius(Point3 o, float r, const Material& mat) { Sphere s; om_sphere_new_with_radius(&s, o.e, r, &mat); return s; }    // traced.rs:26
};
struct Cube : om_cube {
    static Cube new_(const Mat4x4& local_to_world, const Material& mat) { Cube s; om_cube_new(&s, local_to_world.m, &mat); return s; }             // traced.rs:238
    static Cube new_with_length(Point3 o, float len, const Material& mat) { Cube s; om_cube_new_with_length(&s, o.e, len, &mat); return s; }      // traced.rs:242
};
struct InfinitePlane : om_infinite_plane {
    static InfinitePlane new_(Point3 center, Vec3 normal, const Material& mat) { InfinitePlane p; om_infinite_plane_new(&p, center.e, normal.e, &mat); return p; }  // traced.rs:86
};
// Barycentric<BT> (traced.rs:118-131): the two instantiations are distinct types, as in the reference (traced.rs:225-226)
struct Parallelogram : om_barycentric {
    static Parallelogram new3points(Point3 origin, Point3 upoint, Point3 vpoint, const Material& mat) { Parallelogram b; om_barycentric_new3points(&b, origin.e, upoint.e, vpoint.e, &mat); return b; }
};
struct Triangle : om_barycentric {
    static Triangle new3points(Point3 origin, Point3 upoint, Point3 vpoint, const Material& mat) { Triangle b; om_barycentric_new3points(&b, origin.e, upoint.e, vpoint.e, &mat); return b; }
};

// ---- marched.rs ---------------------------------------------------------------------------------------
struct MarchedSphere : om_marched_sphere {
    static MarchedSphere new_(Point3 center, float radius, const Material& mat) {  // marched.rs:56
        MarchedSphere s;
        for (int i = 0; i < 3; i++) s.center[i] = center.e[i];
        s.radius = radius; s.material = mat;
        return s;
    }
};
struct MarchedBox : om_marched_box {
    static MarchedBox new_(Point3 center, Vec3 sizes, const Material& mat) {  // marched.rs:85
        MarchedBox b;
        for (int i = 0; i < 3; i++) { b.center[i] = center.e[i]; b.sizes[i] = sizes.e[i]; }
        b.material = mat;
        return b;
    }
};
struct MarchedTorus : om_marched_torus {
    static MarchedTorus new_(const Mat4x4& local_to_world, Vec3 sizes, const Material& mat) { MarchedTorus t; om_marched_torus_new(&t, local_to_world.m, sizes.e, &mat); return t; }  // marched.rs:115
};

// A user's `impl Marched` (marched.rs:8-45) as data, for `world += Arc::new(obj) as Arc<dyn Marched + Send + Sync>`
// (hits.rs:96-100): to_local / to_world as matrices, to_world_f as a factor, local_sdf as a postfix f32 program
// (OM_SDF_* in ottomarcher.h).  sdf() and get_outward_normal() are the trait's defaults.
struct SdfProgram {
    std::vector<uint32_t> words;
    SdfProgram& op(uint32_t opcode, uint32_t arg = 0) { words.push_back(OM_SDF_INSN(opcode, arg)); return *this; }
    SdfProgram& x() { return op(OM_SDF_X); }
    SdfProgram& y() { return op(OM_SDF_Y); }
    SdfProgram& z() { return op(OM_SDF_Z); }
    SdfProgram& c(float v) { uint32_t b; std::memcpy(&b, &v, 4); words.push_back(OM_SDF_CONST); words.push_back(b); return *this; }
    SdfProgram& add() { return op(OM_SDF_ADD); }
    SdfProgram& sub() { return op(OM_SDF_SUB); }
    SdfProgram& mul() { return op(OM_SDF_MUL); }
    SdfProgram& div() { return op(OM_SDF_DIV); }
    SdfProgram& min() { return op(OM_SDF_MIN); }
    SdfProgram& max() { return op(OM_SDF_MAX); }
    SdfProgram& neg() { return op(OM_SDF_NEG); }
    SdfProgram& abs() { return op(OM_SDF_ABS); }
    SdfProgram& sqrt() { return op(OM_SDF_SQRT); }
    SdfProgram& len2() { return op(OM_SDF_LEN2); }
    SdfProgram& len3() { return op(OM_SDF_LEN3); }
    SdfProgram& store(uint32_t r) { return op(OM_SDF_STORE, r); }
    SdfProgram& load(uint32_t r) { return op(OM_SDF_LOAD, r); }
};
struct DynMarched {
    om_dyn_marched pod;
    SdfProgram local_sdf;
    // `bounds` = Bounded::build_world_bounding_box (bounding_box.rs:69-73); nullptr = the trait default (draw always)
    static DynMarched new_(const Mat4x4& world_to_local, const Mat4x4& local_to_world, float world_f, const SdfProgram& prog, const Material& mat,
                           const Point3* bounds_lo = nullptr, const Point3* bounds_hi = nullptr) {
        DynMarched d;
        std::memcpy(d.pod.m_world_to_local, world_to_local.m, sizeof d.pod.m_world_to_local);
        std::memcpy(d.pod.m_local_to_world, local_to_world.m, sizeof d.pod.m_local_to_world);
        d.pod.world_f = world_f;
        for (int i = 0; i < 3; i++) {
            d.pod.bounds_lo[i] = bounds_lo ? bounds_lo->e[i] : -INFINITY;
            d.pod.bounds_hi[i] = bounds_hi ? bounds_hi->e[i] : INFINITY;
        }
        d.pod.material = mat;
        d.local_sdf = prog;
        return d;
    }
};

// ---- camera.rs:38-59 ----------------------------------------------------------------------------------
struct Camera : om_camera {
    static Camera new_(Point3 lookfrom, Point3 lookat, Vec3 vup, float vfov_in_degrees, float aspect_ratio, float aperture, float focus_dist) {
        Camera c;
        om_camera_new(&c, lookfrom.e, lookat.e, vup.e, vfov_in_degrees, aspect_ratio, aperture, focus_dist);
        return c;
    }
};

// ---- ray.rs:5-13, hits.rs:11-17 ---------------------------------------------------------------------
struct Ray {
    Point3 orig;
    Vec3 dir;  // normalised by the library exactly like Ray::new (ray.rs:11-13)
    static Ray new_(Point3 o, Vec3 d) { return Ray{o, d}; }
};
struct HitRecord {
    Point3 point;
    Vec3 normal;
    float t;
    uint64_t obj_id;  // (OM_TYPE_* << 32) | (index in its list + 1); the reference uses the object's address (utils.rs:110)
};

class FrozenHittableList;

// ---- hits.rs:37-110 HittableList: one Vec per concrete type, filled with `world += obj` ---------------------
class HittableList {
public:
    std::vector<om_sphere> spheres;
    std::vector<om_cube> cubes;
    std::vector<om_barycentric> triangles;
    std::vector<om_infinite_plane> infinite_planes;
    std::vector<om_barycentric> parallelograms;
    std::vector<om_marched_sphere> marched_spheres;
    std::vector<om_marched_box> marched_boxes;
    std::vector<om_marched_torus> marched_tori;
    std::vector<DynMarched> marched_objects;  // Vec<Arc<dyn Marched + Send + Sync>> (hits.rs:40), as data

    static HittableList new_() { return HittableList(); }  // hits.rs:72
    void clear() {                                          // hits.rs:81
        spheres.clear(); cubes.clear(); triangles.clear(); infinite_planes.clear(); parallelograms.clear();
        marched_spheres.clear(); marched_boxes.clear(); marched_tori.clear(); marched_objects.clear();
    }
    // impl AddAssign<&T> for HittableList (hits.rs:101-110)
    HittableList& operator+=(const Sphere& o) { spheres.push_back(o); return *this; }
    HittableList& operator+=(const Cube& o) { cubes.push_back(o); return *this; }
    HittableList& operator+=(const Triangle& o) { triangles.push_back(o); return *this; }
    HittableList& operator+=(const InfinitePlane& o) { infinite_planes.push_back(o); return *this; }
    HittableList& operator+=(const Parallelogram& o) { parallelograms.push_back(o); return *this; }
    HittableList& operator+=(const MarchedSphere& o) { marched_spheres.push_back(o); return *this; }
    HittableList& operator+=(const MarchedBox& o) { marched_boxes.push_back(o); return *this; }
    HittableList& operator+=(const MarchedTorus& o) { marched_tori.push_back(o); return *this; }
    // AddAssign<Arc<dyn Marched + Send + Sync>> (hits.rs:96-100) for a Marched type described as data
    HittableList& operator+=(const DynMarched& o) { marched_objects.push_back(o); return *this; }
    // AddAssign<Arc<dyn Traced>> (hits.rs:91-95) and native trait objects: compiled host code cannot cross to the device
    HittableList& operator+=(const void* dyn_object) {
        om_scene* s = nullptr;
        check(om_scene_create(&s));
        const om_status st = om_scene_add_dyn_object(s, dyn_object);
        const std::string msg = om_last_error();
        om_scene_destroy(s);
        throw Error(st, msg);
    }
    size_t len() const {
        return spheres.size() + cubes.size() + triangles.size() + infinite_planes.size() + parallelograms.size() + marched_spheres.size() +
               marched_boxes.size() + marched_tori.size() + marched_objects.size();
    }
    // The library's seeded synthetic scenes (om_scene_generate), pulled back into host lists so that they enter the
    // renderer the same way a hand-built world does.
    static HittableList generate(uint32_t kind, uint64_t seed, uint64_t n_param = 0) {
        om_scene* s = nullptr;
        check(om_scene_create(&s));
        HittableList w;
        try {
            check(om_scene_generate(s, kind, seed, n_param));
            uint64_t n[8];
            check(om_scene_counts(s, n));
            auto pull = [&](uint32_t which, auto& vec) {
                vec.resize(size_t(n[which]));
                if (!vec.empty()) check(om_scene_get_list(s, which, vec.data(), vec.size() * sizeof(vec[0])));
            };
            pull(0, w.spheres); pull(1, w.cubes); pull(2, w.triangles); pull(3, w.infinite_planes); pull(4, w.parallelograms);
            pull(5, w.marched_spheres); pull(6, w.marched_boxes); pull(7, w.marched_tori);
        } catch (...) { om_scene_destroy(s); throw; }
        om_scene_destroy(s);
        return w;
    }
    inline FrozenHittableList freeze(const Camera* cam, uint32_t flags = 0) const;  // hits.rs:87-89
};

// ---- hits.rs:63-69 FrozenHittableList: the frozen scene (BVH in place of CameraHash), shareable between renders
class FrozenHittableList {
public:
    FrozenHittableList() = default;
    FrozenHittableList(const HittableList& hl, const Camera* cam, uint32_t flags = 0) {  // FrozenHittableList::new hits.rs:116
        check(om_scene_create(&h_));
        try {
            check(om_scene_add_spheres(h_, hl.spheres.data(), hl.spheres.size()));
            check(om_scene_add_cubes(h_, hl.cubes.data(), hl.cubes.size()));
            check(om_scene_add_triangles(h_, hl.triangles.data(), hl.triangles.size()));
            check(om_scene_add_infinite_planes(h_, hl.infinite_planes.data(), hl.infinite_planes.size()));
            check(om_scene_add_parallelograms(h_, hl.parallelograms.data(), hl.parallelograms.size()));
            check(om_scene_add_marched_spheres(h_, hl.marched_spheres.data(), hl.marched_spheres.size()));
            check(om_scene_add_marched_boxes(h_, hl.marched_boxes.data(), hl.marched_boxes.size()));
            check(om_scene_add_marched_tori(h_, hl.marched_tori.data(), hl.marched_tori.size()));
            for (const DynMarched& d : hl.marched_objects) check(om_scene_add_dyn_marched(h_, &d.pod, d.local_sdf.words.data(), d.local_sdf.words.size()));
            check(om_scene_freeze(h_, cam, flags));
        } catch (...) { om_scene_destroy(h_); h_ = nullptr; throw; }
    }
    FrozenHittableList(FrozenHittableList&& o) noexcept : h_(o.h_) { o.h_ = nullptr; }
    FrozenHittableList& operator=(FrozenHittableList&& o) noexcept { if (this != &o) { om_scene_destroy(h_); h_ = o.h_; o.h_ = nullptr; } return *this; }
    FrozenHittableList(const FrozenHittableList&) = delete;
    FrozenHittableList& operator=(const FrozenHittableList&) = delete;
    ~FrozenHittableList() { om_scene_destroy(h_); }
    const om_scene* handle() const { return h_; }

    // hits.rs:270-334 for one ray (on the device; meant for tests, a batch call is om_trace_rays). false = None.
    bool hit(const Ray& r, float t_min, float t_max, HitRecord* rec, uint32_t flags = 0, int32_t device = -1) const {
        uint64_t id = 0; float t = 0.f, p[3], n[3];
        check(om_trace_rays(h_, r.orig.e, r.dir.e, 1, t_min, t_max, flags, device, &id, &t, p, n));
        if (id == 0) return false;
        if (rec) { rec->point = Point3(p[0], p[1], p[2]); rec->normal = Vec3(n[0], n[1], n[2]); rec->t = t; rec->obj_id = id; }
        return true;
    }
    // hits.rs:187: the camera-hash accelerated first hit gives the same answer as hit() by construction
    bool first_hit(const Ray& r, float t_min, float t_max, float /*u*/, float /*v*/, HitRecord* rec) const { return hit(r, t_min, t_max, rec); }

private:
    om_scene* h_ = nullptr;
};
inline FrozenHittableList HittableList::freeze(const Camera* cam, uint32_t flags) const { return FrozenHittableList(*this, cam, flags); }

namespace render_thread {

using Pixel = om_pixel;  // render_thread.rs:43-46 with Stats (render_thread.rs:10-17) flattened; Pixel::new() == Pixel{}

// What the reference leaves to the OS (RNG seed) or does not have (flags, pass size), with defaults that reproduce it.
struct Options {
    uint64_t seed = 1;                    // Philox key in place of the OS-seeded thread_rng (utils.rs:25)
    uint32_t flags = OM_RENDER_ADAPTIVE;  // the reference always retires pixels early (render_thread.rs:31-38,194-198)
    uint32_t pass_spp = 0;                // samples per pixel per launch; 0 = all at once.  Small passes make the counter move
    int32_t device = -2;                  // -2: thread `tid` drives device `tid % om_device_count()`
};

// render_thread::render (render_thread.rs:145-202), same arguments in the same order.  Each call stands for one of
// the reference's render threads: it owns the pixels p with (p / 2730) % num_threads == tid (`assigned_thread`,
// main.rs:172-189), renders them on its device, adds what it traced to `samples_atom` after every pass (leftover
// crediting included, render_thread.rs:196-198) and stores the finished pixels into `pixels`.  Threads may run
// concurrently on the same vector: they write disjoint elements.
inline void render(const Camera& camera, const FrozenHittableList& world, uint32_t max_depth, float tmin, float tmax, uint32_t samples_per_pixel,
                   uint32_t image_width, uint32_t image_height, std::vector<Pixel>& pixels, uint32_t tid, uint32_t num_threads,
                   std::atomic<uint64_t>& samples_atom, const Options& opt = Options()) {
    const size_t npx = size_t(image_width) * image_height;
    if (pixels.size() != npx) throw Error(OM_ERR_INVALID, "render: pixels.len() != image_width*image_height");
    if (num_threads == 0 || tid >= num_threads) throw Error(OM_ERR_INVALID, "render: tid out of range");
    om_render_params p{};
    p.width = image_width; p.height = image_height; p.spp_max = samples_per_pixel; p.max_depth = max_depth;
    p.tmin = tmin; p.tmax = tmax; p.seed = opt.seed; p.flags = opt.flags;
    p.part_rank = tid; p.part_world = num_threads;
    if (opt.device == -2) {
        const int32_t n = om_device_count();
        if (n <= 0) throw Error(OM_ERR_NO_DEVICE, "render: no CUDA device (this library has no CPU path)");
        p.device = int32_t(tid % uint32_t(n));
    } else p.device = opt.device;
    om_render* r = nullptr;
    check(om_render_create(&r, world.handle(), &camera, &p));
    try {
        // total this thread has to account for = its pixels x spp (main.rs:154)
        uint64_t mine = 0;
        for (size_t c = 0; c * OM_CHUNK_PIXELS < npx; c++)
            if (c % num_threads == tid) mine += std::min<size_t>(OM_CHUNK_PIXELS, npx - c * OM_CHUNK_PIXELS);
        const uint64_t total = mine * samples_per_pixel;
        uint64_t seen = 0;
        const uint32_t pass = opt.pass_spp ? opt.pass_spp : samples_per_pixel;
        for (uint32_t done_spp = 0; done_spp < samples_per_pixel && seen < total; done_spp += pass) {
            check(om_render_launch(r, pass));
            uint64_t credited = 0;
            check(om_render_samples_done(r, &credited));
            samples_atom.fetch_add(credited - seen, std::memory_order_relaxed);  // render_thread.rs:199
            seen = credited;
        }
        std::vector<Pixel> all(npx);
        check(om_render_read_pixels(r, all.data(), npx));
        for (size_t c = 0; c * OM_CHUNK_PIXELS < npx; c++) {
            if (c % num_threads != tid) continue;
            const size_t b = c * OM_CHUNK_PIXELS, e = std::min(npx, b + OM_CHUNK_PIXELS);
            for (size_t i = b; i < e; i++) pixels[i] = all[i];
        }
    } catch (...) { om_render_destroy(r); throw; }
    om_render_destroy(r);
}

}  // namespace render_thread

// main.rs:473-476: the viewer's "save" key writes what is on screen; MODE_NORMAL shows pixel.stats.color (main.rs:374-381)
inline void save_bmp(const std::string& path, const std::vector<render_thread::Pixel>& pixels, uint32_t width, uint32_t height) {
    std::vector<uint8_t> rgb(pixels.size() * 3);
    for (size_t i = 0; i < pixels.size(); i++) { rgb[3 * i] = pixels[i].color[0]; rgb[3 * i + 1] = pixels[i].color[1]; rgb[3 * i + 2] = pixels[i].color[2]; }
    check(om_write_bmp(path.c_str(), rgb.data(), width, height));
}

}  // namespace ottomarcher
#endif
