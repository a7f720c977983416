// ORACLE — TEST INFRASTRUCTURE ONLY (see ref_math.hpp header).  PARITY UNPINNED.
//
// Follows: src/ray.rs, src/camera.rs:38-74, src/materials.rs, src/traced.rs, src/marched.rs,
//          src/hits.rs:11-17,270-365, src/math/vec3.rs:80-114 (samplers).
// Deviations from the reference, all documented in DESIGN.md:
//  * first_hit (hits.rs:187-268, the CameraHash-culled primary query) is DEFINED as hit(): the
//    grid at HEAD registers marched objects in the wrong list and overflows its 16-slot cells
//    (SURVEY.md F3), so its intended semantics — closest hit over all objects — is used.
//  * obj_id is (type << 32) | (index + 1) instead of an address (utils.rs:110).
//  * unstuck()'s inner loop (hits.rs:360-363) has no bound in the reference; here it stops after
//    UNSTUCK_MAX_STEPS steps (never reached by any test scene).
//  * random numbers come from ref::Rng (ref_rng.hpp) passed explicitly.
#pragma once
#include <cstring>
#include <vector>

#include "ref_math.hpp"
#include "ref_rng.hpp"

namespace ref {

// ---- vec3.rs:80-114 -----------------------------------------------------------------------
inline Vec3 vec3_rand(Rng& g) { float a = g.rand(), b = g.rand(), c = g.rand(); return Vec3(a, b, c); }
inline Vec3 vec3_rand_range(Rng& g, float mn, float mx) {
    float a = g.rand_range(mn, mx), b = g.rand_range(mn, mx), c = g.rand_range(mn, mx);
    return Vec3(a, b, c);
}
inline Vec3 rand_in_unit_sphere(Rng& g) {  // vec3.rs:92-97
    for (;;) {
        Vec3 p = vec3_rand_range(g, -1.f, 1.f);
        if (p.length_squared() < 1.f) return p;
    }
}
inline Vec3 rand_unit_vector(Rng& g) { return rand_in_unit_sphere(g).unit(); }  // :98-100
inline Vec3 rand_in_unit_disc(Rng& g) {                                         // :108-113
    for (;;) {
        float a = g.rand_range(-1.f, 1.f), b = g.rand_range(-1.f, 1.f);
        Vec3 p(a, b, 0.f);
        if (p.length_squared() < 1.f) return p;
    }
}

// ---- ray.rs --------------------------------------------------------------------------------
struct Ray {
    Vec3 orig, dir;
    static Ray make(const Vec3& o, const Vec3& d) { Ray r; r.orig = o; r.dir = d.unit(); return r; }  // ray.rs:11-13
    Vec3 at(float t) const { return orig + t * dir; }                                                 // :14-16
    Ray transform(const Mat4x4& m) const { Ray r; r.orig = m.dot_p3(orig); r.dir = m.dot_v3(dir); return r; }  // :17-19
};

// ---- materials.rs --------------------------------------------------------------------------
enum MaterialType { LAMBERTIAN = 0, METAL = 1, DIELECTRIC = 2 };
struct Material {
    Vec3 albedo;
    float fuzz = 0.f, ior = 0.f;
    uint32_t mat_type = LAMBERTIAN;
    static Material new_lambertian(const Vec3& a) { Material m; m.albedo = a; m.mat_type = LAMBERTIAN; return m; }
    static Material new_metal(const Vec3& a) { Material m; m.albedo = a; m.mat_type = METAL; return m; }
    static Material new_metal_fuzz(const Vec3& a, float f) { Material m; m.albedo = a; m.fuzz = f; m.mat_type = METAL; return m; }
    static Material new_dielectric(float ior) { Material m; m.ior = ior; m.mat_type = DIELECTRIC; return m; }
};

struct HitRecord {  // hits.rs:11-17
    Vec3 point, normal;
    Material material;
    float t;
    uint64_t obj_id;
};

inline Vec3 reflect(const Vec3& v, const Vec3& n) { return v - 2.f * v.dot(n) * n; }  // materials.rs:98-100
inline Vec3 refract(const Vec3& uv, const Vec3& n, float etai_over_etat) {            // :102-108
    float cos_theta = fmin_((-uv).dot(n), 1.0f);
    Vec3 r_out_perp = etai_over_etat * (uv + cos_theta * n);
    float aux = -std::sqrt(std::fabs(1.0f - r_out_perp.length_squared()));
    Vec3 r_out_parallel = aux * n;
    return r_out_perp + r_out_parallel;
}
inline float reflectance(float cos, float ref_idx) {  // :110-116 (Schlick)
    float r0 = (1.f - ref_idx) / (1.f + ref_idx);
    float r0_2 = r0 * r0;
    float cos_5 = (1.f - cos) * (1.f - cos) * (1.f - cos) * (1.f - cos) * (1.f - cos);
    return r0_2 + (1.f - r0_2) * cos_5;
}

struct ScatterResult { Vec3 attenuation; Ray ray; };

inline ScatterResult scatter(const Material& m, const Ray& r_in, const HitRecord& hr, Rng& g) {  // materials.rs:39-51
    ScatterResult out;
    if (m.mat_type == LAMBERTIAN) {  // :52-61
        Vec3 new_dir = hr.normal + rand_unit_vector(g);
        if (new_dir.near_zero()) new_dir = hr.normal;
        out.attenuation = m.albedo;
        out.ray = Ray::make(hr.point, new_dir);
    } else if (m.mat_type == METAL) {  // :62-66 (no absorb test; draws even when fuzz == 0)
        Vec3 reflected = reflect(r_in.dir, hr.normal);
        Vec3 s = rand_in_unit_sphere(g);
        out.attenuation = m.albedo;
        out.ray = Ray::make(hr.point, reflected + m.fuzz * s);
    } else {  // :68-95
        Vec3 dir_unit = r_in.dir;
        bool front_face = dir_unit.dot(hr.normal) < 0.f;
        float refraction_ratio;
        Vec3 normal;
        if (front_face) { refraction_ratio = 1.0f / m.ior; normal = hr.normal; }
        else            { refraction_ratio = m.ior;        normal = -hr.normal; }
        float cos_theta = fmin_((-dir_unit).dot(normal), 1.0f);
        float sin_theta = std::sqrt(1.0f - cos_theta * cos_theta);
        bool cannot_refract = (refraction_ratio * sin_theta) > 1.0f;
        bool reflect_by_reflectance = reflectance(cos_theta, refraction_ratio) > g.rand();
        Vec3 new_dir = (cannot_refract || reflect_by_reflectance) ? reflect(dir_unit, normal)
                                                                  : refract(dir_unit, normal, refraction_ratio);
        out.attenuation = Vec3(1.f, 1.f, 1.f);
        out.ray = Ray::make(hr.point, new_dir);
    }
    return out;
}

// ---- object ids ----------------------------------------------------------------------------
enum { T_SPHERE = 1, T_CUBE = 2, T_TRIANGLE = 3, T_INFINITE_PLANE = 4, T_PARALLELOGRAM = 5,
       T_MARCHED_SPHERE = 6, T_MARCHED_BOX = 7, T_MARCHED_TORUS = 8, T_DYN_MARCHED = 9, T_DYN_TRACED = 10 };
inline uint64_t make_id(uint32_t type, size_t index) { return ((uint64_t)type << 32) | (uint64_t)(index + 1); }

// ---- traced.rs -----------------------------------------------------------------------------
struct Sphere {  // traced.rs:13-63
    Mat4x4 m_local_to_world, m_word_to_local;
    Material material;
    static Sphere make(const Mat4x4& m, const Material& mat) { Sphere s; s.m_local_to_world = m; s.m_word_to_local = m.fast_homogenous_inverse(); s.material = mat; return s; }
    static Sphere new_with_radius(const Vec3& o, float r, const Material& mat) {
        return make(Mat4x4::new_translate(o).dot_mat(Mat4x4::new_scale(Vec3(r, r, r))), mat);
    }
    bool hit(const Ray& r, float t_min, float t_max, uint64_t id, HitRecord& out) const {
        Ray new_r = r.transform(m_word_to_local);
        float a = new_r.dir.length_squared();
        float half_b = new_r.orig.dot(new_r.dir);
        float c = new_r.orig.length_squared() - 1.f;
        float discriminant = half_b * half_b - a * c;
        if (discriminant < 0.f) return false;
        float sqrtd = std::sqrt(discriminant);
        float root = (-half_b - sqrtd) / a;
        if (root < t_min || root > t_max) {
            root = (-half_b + sqrtd) / a;
            if (root < t_min || root > t_max) return false;
        }
        Vec3 local_point = new_r.at(root);
        out.t = root;
        out.point = m_local_to_world.dot_p3(local_point);
        out.normal = m_local_to_world.dot_v3(local_point).unit();
        out.material = material;
        out.obj_id = id;
        return true;
    }
};

struct InfinitePlane {  // traced.rs:77-115
    Vec3 center, normal;
    Material material;
    static InfinitePlane make(const Vec3& c, const Vec3& n, const Material& m) { InfinitePlane p; p.center = c; p.normal = n.unit(); p.material = m; return p; }
};
// traced.rs:91-99: returns (t, normal·dir); t = INF when parallel
inline void ray_plane_intersect(const Ray& r, const Vec3& plane_normal, const Vec3& plane_center, float& t, float& div) {
    div = plane_normal.dot(r.dir);
    if (std::fabs(div) < 0.000001f) { t = INF; div = 0.f; return; }
    float num = -plane_normal.dot(r.orig - plane_center);
    t = num / div;
}
inline Vec3 normal_against_direction(const Vec3& normal, float normal_dot_dir) {  // :100-103
    return normal * std::copysign(1.f, -normal_dot_dir);
}
inline bool plane_hit(const InfinitePlane& p, const Ray& r, float t_min, float t_max, uint64_t id, HitRecord& out) {
    float root, ndd;
    ray_plane_intersect(r, p.normal, p.center, root, ndd);
    if (root == INF || root < t_min || root > t_max) return false;
    out.t = root;
    out.point = r.at(root);
    out.normal = normal_against_direction(p.normal, ndd);
    out.material = p.material;
    out.obj_id = id;
    return true;
}

struct Barycentric {  // traced.rs:118-200; bt 0 = Parallelogram, 1 = Triangle
    Vec3 origin, u;
    float u_length;
    Vec3 v;
    float v_length;
    Vec3 uxv, uxvxu;
    Mat3x3 base_inv;
    Vec3 v_in_base;
    Material material;
    static Barycentric make(const Vec3& origin, const Vec3& u, const Vec3& v, float u_length, float v_length, const Material& m) {  // :135-147
        Barycentric b;
        Vec3 u_unit = u.unit(), v_unit = v.unit();
        Vec3 uxv = u_unit.cross(v_unit).unit();
        Vec3 uxvxu = uxv.cross(u_unit).unit();
        b.base_inv = Mat3x3::new_3vec_vert(u_unit, uxv, uxvxu).transpose();
        b.v_in_base = b.base_inv.dot(v_unit);
        b.origin = origin; b.material = m; b.u = u_unit; b.u_length = u_length; b.v = v_unit; b.v_length = v_length;
        b.uxv = uxv; b.uxvxu = uxvxu;
        return b;
    }
    static Barycentric new3points(const Vec3& origin, const Vec3& upoint, const Vec3& vpoint, const Material& m) {  // :148-154
        Vec3 u_rel = upoint - origin; float ul = u_rel.length();
        Vec3 v_rel = vpoint - origin; float vl = v_rel.length();
        return make(origin, u_rel, v_rel, ul, vl, m);
    }
    bool hit(int bt, const Ray& r, float t_min, float t_max, uint64_t id, HitRecord& out) const {  // :176-200
        float root, ndd;
        ray_plane_intersect(r, uxv, origin, root, ndd);
        if (root == INF || root < t_min || root > t_max) return false;
        Vec3 point = r.at(root);
        Vec3 point_from_origin = point - origin;
        Vec3 coords2d = base_inv.dot(point_from_origin);
        // calc_barycentric :155-167
        float rx = coords2d.x(), ry = coords2d.z();
        float ux = 1.f, uy = 0.f;
        float vx = v_in_base.x(), vy = v_in_base.z();
        float det = ux * vy - vx * uy;
        float lambda1 = ((rx * vy - vx * ry) / det) / u_length;
        float lambda2 = (-(rx * uy - ux * ry) / det) / v_length;
        float lambda3 = 1.f - lambda1 - lambda2;
        bool correct;
        if (bt == 0) correct = lambda1 > 0.f && lambda2 > 0.f && lambda1 < 1.f && lambda2 < 1.f;  // :173-175
        else correct = lambda1 > 0.f && lambda2 > 0.f && lambda3 > 0.f && lambda1 < 1.f && lambda2 < 1.f && lambda3 < 1.f;  // :169-171
        if (!correct) return false;
        out.t = root; out.point = point; out.normal = normal_against_direction(uxv, ndd);
        out.material = material; out.obj_id = id;
        return true;
    }
};

struct Cube {  // traced.rs:229-299
    Mat4x4 m_local_to_world, m_word_to_local;
    Material material;
    static Cube make(const Mat4x4& m, const Material& mat) { Cube c; c.m_local_to_world = m; c.m_word_to_local = m.fast_homogenous_inverse(); c.material = mat; return c; }
    static Cube new_with_length(const Vec3& o, float len, const Material& mat) {
        return make(Mat4x4::new_translate(o).dot_mat(Mat4x4::new_scale(Vec3(len, len, len))), mat);
    }
    bool hit(const Ray& r, float t_min, float t_max, uint64_t id, HitRecord& out) const {
        Ray new_r = r.transform(m_word_to_local);
        float smallest_t = INF;
        int idx = -1;
        for (int i = 0; i < 3; i++) {
            if (std::fabs(new_r.dir[i]) < 0.00001f) continue;
            float t1 = (0.5f - new_r.orig[i]) / new_r.dir[i];
            float t2 = (-0.5f - new_r.orig[i]) / new_r.dir[i];
            float t;
            if (t1 >= 0.f && t2 >= 0.f) t = fmin_(t1, t2); else t = fmax_(t1, t2);
            if (t > smallest_t || t > t_max || t < t_min) continue;
            float f = new_r.at(t).abs().max_val();
            bool is_solution = std::fabs(f - 0.5f) <= 0.00001f;
            if (!is_solution) continue;
            smallest_t = t;
            idx = i;
        }
        if (idx < 0) return false;
        Vec3 local_point = new_r.at(smallest_t);
        const Vec3 axes[3] = {Vec3(1, 0, 0), Vec3(0, 1, 0), Vec3(0, 0, 1)};
        Vec3 local_outward_normal = axes[idx] * std::copysign(1.f, new_r.at(smallest_t)[idx]);
        out.t = smallest_t;
        out.point = m_local_to_world.dot_p3(local_point);
        out.normal = m_local_to_world.dot_v3(local_outward_normal);  // not normalised (reference quirk)
        out.material = material;
        out.obj_id = id;
        return true;
    }
};

// ---- marched.rs ----------------------------------------------------------------------------
// The trait's default methods (marched.rs:14-44) are written once as templates over the three
// concrete types; each type supplies to_local / to_world / to_world_f / local_sdf.
template <class M> inline float marched_sdf(const M& m, const Vec3& p) {  // marched.rs:14-18
    Vec3 pl = m.to_local(Vec4::new_p3(p)).xyz();
    return m.to_world_f(m.local_sdf(pl));
}
template <class M> inline Vec3 marched_outward_local_normal(const M& m, const Vec3& p) {  // marched.rs:25-44
    const float eps = 0.0000001f;
    Vec3 ex(eps, 0.f, 0.f), ey(0.f, eps, 0.f), ez(0.f, 0.f, eps);
    float x = m.local_sdf(p + ex) - m.local_sdf(p - ex);
    float y = m.local_sdf(p + ey) - m.local_sdf(p - ey);
    float z = m.local_sdf(p + ez) - m.local_sdf(p - ez);
    Vec3 normal = Vec3(x, y, z).unit();
    Ray test_ray = Ray::make(Vec3(), normal);
    Vec3 start = test_ray.at(0.f);
    float start_val = m.local_sdf(start);
    Vec3 end = test_ray.at(1.f);
    float end_val = m.local_sdf(end);
    float sign = (end_val > start_val) ? 1.f : -1.f;
    return normal * sign;
}
template <class M> inline Vec3 marched_outward_normal_default(const M& m, const Vec3& p) {  // marched.rs:19-24
    Vec3 pl = m.to_local(Vec4::new_p3(p)).xyz();
    Vec4 n = Vec4::new_v3(marched_outward_local_normal(m, pl));
    Vec4 world_n = m.to_world(n);
    return world_n.xyz().unit();
}

struct MarchedSphere {  // marched.rs:49-76
    Vec3 center; float radius; Material material;
    float local_sdf(const Vec3& p) const { return p.length() - radius; }
    Vec4 to_local(const Vec4& p) const { return p - p.w() * Vec4::new_v3(center); }
    Vec4 to_world(const Vec4& p) const { return p + p.w() * Vec4::new_v3(center); }
    float to_world_f(float f) const { return f; }
    Vec3 get_outward_normal(const Vec3& p) const { return (p - center).unit(); }  // override :60-63
};
struct MarchedBox {  // marched.rs:78-102
    Vec3 center, sizes; Material material;
    float local_sdf(const Vec3& p) const {
        Vec3 q = p.abs() - sizes;
        return q.max(Vec3()).length() + fmin_(fmax_(q.x(), fmax_(q.y(), q.z())), 0.f);
    }
    Vec4 to_local(const Vec4& p) const { return p - p.w() * Vec4::new_v3(center); }
    Vec4 to_world(const Vec4& p) const { return p + p.w() * Vec4::new_v3(center); }
    float to_world_f(float f) const { return f; }
    Vec3 get_outward_normal(const Vec3& p) const { return marched_outward_normal_default(*this, p); }
};
struct MarchedTorus {  // marched.rs:104-151
    Mat4x4 m_local_to_world_translate_rotate, m_world_to_local_translate_rotate;
    Vec4 m_local_to_world_scale, m_world_to_local_scale;
    Vec3 sizes; Material material;
    static MarchedTorus make(const Mat4x4& m_local_to_world, const Vec3& local_sizes, const Material& mat) {  // :115-130
        Mat4x4 t, r, s;
        m_local_to_world.decompose_into_translate_rotate_scale(t, r, s);
        Mat4x4 translate_rotate = t.dot_mat(r);
        Vec4 scale = s.diag();
        Vec4 scale_inv(1.f / scale.x(), 1.f / scale.y(), 1.f / scale.z(), 1.f / scale.w());
        MarchedTorus o;
        o.m_local_to_world_translate_rotate = translate_rotate;
        o.m_world_to_local_translate_rotate = translate_rotate.fast_homogenous_inverse();
        o.m_local_to_world_scale = scale; o.m_world_to_local_scale = scale_inv;
        o.sizes = local_sizes; o.material = mat;
        return o;
    }
    float local_sdf(const Vec3& p) const {  // :134-138
        Vec3 q(Vec3(p.x(), p.z(), 0.f).length() - sizes.x(), p.y(), 0.f);
        return q.length() - sizes.y();
    }
    Vec4 to_local(const Vec4& p) const { return m_world_to_local_translate_rotate.dot(p * m_world_to_local_scale); }  // :142-144
    Vec4 to_world(const Vec4& p) const { return m_local_to_world_translate_rotate.dot(p) * m_local_to_world_scale; }  // :145-147
    float to_world_f(float f) const { return f * m_local_to_world_scale.xyz().min_val(); }                            // :148-150
    Vec3 get_outward_normal(const Vec3& p) const { return marched_outward_normal_default(*this, p); }
};

// A user's `impl Marched` handed over as data (include/ottomarcher.h, om_dyn_marched): stands for an element of
// `marched_objects: Vec<Arc<dyn Marched + Send + Sync>>` (hits.rs:40).  The type supplies to_local / to_world /
// to_world_f / local_sdf; sdf and the normal are the trait's defaults (marched.rs:14-44).  local_sdf is a postfix
// program, one f32 operation per instruction in program order (opcodes: OM_SDF_* in the header).
struct DynMarched {
    Mat4x4 m_world_to_local, m_local_to_world;
    float world_f;
    std::vector<uint32_t> code;
    Material material;
    float local_sdf(const Vec3& p) const {
        float st[8], reg[4] = {0.f, 0.f, 0.f, 0.f};
        int sp = 0;
        for (size_t pc = 0; pc < code.size(); pc++) {
            const uint32_t op = code[pc] & 0xFFu, arg = code[pc] >> 8;
            switch (op) {
                case 1: st[sp++] = p.x(); break;
                case 2: st[sp++] = p.y(); break;
                case 3: st[sp++] = p.z(); break;
                case 4: { pc++; float f; std::memcpy(&f, &code[pc], 4); st[sp++] = f; break; }
                case 5: sp--; st[sp - 1] = st[sp - 1] + st[sp]; break;
                case 6: sp--; st[sp - 1] = st[sp - 1] - st[sp]; break;
                case 7: sp--; st[sp - 1] = st[sp - 1] * st[sp]; break;
                case 8: sp--; st[sp - 1] = st[sp - 1] / st[sp]; break;
                case 9: sp--; st[sp - 1] = fmin_(st[sp - 1], st[sp]); break;
                case 10: sp--; st[sp - 1] = fmax_(st[sp - 1], st[sp]); break;
                case 11: st[sp - 1] = -st[sp - 1]; break;
                case 12: st[sp - 1] = std::fabs(st[sp - 1]); break;
                case 13: st[sp - 1] = std::sqrt(st[sp - 1]); break;
                case 14: sp--; st[sp - 1] = Vec3(st[sp - 1], st[sp], 0.f).length(); break;
                case 15: sp -= 2; st[sp - 1] = Vec3(st[sp - 1], st[sp], st[sp + 1]).length(); break;
                case 16: sp--; reg[arg & 3u] = st[sp]; break;
                default: st[sp++] = reg[arg & 3u]; break;
            }
        }
        return st[0];
    }
    Vec4 to_local(const Vec4& p) const { return m_world_to_local.dot(p); }
    Vec4 to_world(const Vec4& p) const { return m_local_to_world.dot(p); }
    float to_world_f(float f) const { return f * world_f; }
    Vec3 get_outward_normal(const Vec3& p) const { return marched_outward_normal_default(*this, p); }
};

// ---- hits.rs -------------------------------------------------------------------------------
constexpr float HIT_SIZE = 0.001f;  // hits.rs:113
constexpr int UNSTUCK_MAX_STEPS = 1 << 20;

// Arc<dyn Traced> described as data (include/ottomarcher.h, om_dyn_traced): a quadric q(p) = A x^2 + B y^2 + C z^2 + D xy + E xz + F yz
// + G x + H y + I z + J in the object's frame, optionally cut by a box of that frame.  The reference has no code for this (the trait is
// the user's to implement, traced.rs:33-35); the semantics are the header's, stated here a second time in plain C++ in the shape of
// Sphere::hit (traced.rs:38-63), which it equals bit for bit for the unit sphere.
struct DynTraced {
    Mat4x4 m_local_to_world, m_word_to_local;
    float q[10];
    Vec3 clip_lo, clip_hi;
    Material material;
    bool inside_clip(const Vec3& p) const {
        return !(p.x() < clip_lo.x() || p.x() > clip_hi.x()) && !(p.y() < clip_lo.y() || p.y() > clip_hi.y()) && !(p.z() < clip_lo.z() || p.z() > clip_hi.z());
    }
    bool hit(const Ray& r, float t_min, float t_max, uint64_t id, HitRecord& out) const {
        const float A = q[0], B = q[1], C = q[2], D = q[3], E = q[4], F = q[5], G = q[6], H = q[7], I = q[8], J = q[9];
        Ray new_r = r.transform(m_word_to_local);
        const Vec3 o = new_r.orig, d = new_r.dir;
        float a = A * d.x() * d.x() + B * d.y() * d.y() + C * d.z() * d.z() + D * d.x() * d.y() + E * d.x() * d.z() + F * d.y() * d.z();
        float half_b = A * o.x() * d.x() + B * o.y() * d.y() + C * o.z() * d.z()
                       + 0.5f * (D * (o.x() * d.y() + o.y() * d.x()) + E * (o.x() * d.z() + o.z() * d.x()) + F * (o.y() * d.z() + o.z() * d.y()))
                       + 0.5f * (G * d.x() + H * d.y() + I * d.z());
        float c = A * o.x() * o.x() + B * o.y() * o.y() + C * o.z() * o.z() + D * o.x() * o.y() + E * o.x() * o.z() + F * o.y() * o.z() + G * o.x() + H * o.y() + I * o.z() + J;
        auto ok = [&](float root) { return !(root < t_min || root > t_max) && inside_clip(new_r.at(root)); };
        float root;
        if (a == 0.f) {  // the quadric is linear along this ray: one root
            if (half_b == 0.f) return false;
            root = -c / (2.f * half_b);
            if (!ok(root)) return false;
        } else {
            float discriminant = half_b * half_b - a * c;
            if (discriminant < 0.f) return false;
            float sqrtd = std::sqrt(discriminant);
            const float s = a < 0.f ? -1.f : 1.f;  // nearer root first, also where the form is negative along the ray
            root = (-half_b - s * sqrtd) / a;
            if (!ok(root)) {
                root = (-half_b + s * sqrtd) / a;
                if (!ok(root)) return false;
            }
        }
        Vec3 p = new_r.at(root);
        Vec3 half_grad(A * p.x() + 0.5f * (D * p.y() + E * p.z() + G), B * p.y() + 0.5f * (D * p.x() + F * p.z() + H), C * p.z() + 0.5f * (E * p.x() + F * p.y() + I));
        out.t = root;
        out.point = m_local_to_world.dot_p3(p);
        out.normal = m_local_to_world.dot_v3(half_grad).unit();
        out.material = material;
        out.obj_id = id;
        return true;
    }
};

struct World {  // HittableList / FrozenHittableList, hits.rs:37-69 (no dyn lists, no CameraHash)
    std::vector<Sphere> spheres;
    std::vector<Cube> cubes;
    std::vector<Barycentric> triangles;
    std::vector<InfinitePlane> infinite_planes;
    std::vector<Barycentric> parallelograms;
    std::vector<MarchedSphere> marched_spheres;
    std::vector<MarchedBox> marched_boxes;
    std::vector<MarchedTorus> marched_torus;
    std::vector<DynMarched> marched_objects;  // Vec<Arc<dyn Marched>>, scanned after the typed lists (hits.rs:312-320,353-359)
    std::vector<DynTraced> traced_objects;    // Vec<Arc<dyn Traced>>, scanned after the typed traced lists (hits.rs:280-285)

    // counters for the flop accounting of SURVEY.md §8(d); filled when non-null
    struct Counters { uint64_t segments = 0, traced_tests = 0, sdf_evals = 0, normal_evals = 0, march_steps = 0; };

    bool has_marched() const { return !marched_spheres.empty() || !marched_boxes.empty() || !marched_torus.empty() || !marched_objects.empty(); }
    size_t n_marched() const { return marched_spheres.size() + marched_boxes.size() + marched_torus.size() + marched_objects.size(); }

    // hits.rs:336-365
    float unstuck(float t, const Ray& r, Counters* cn) const {
        const float MIN_STEP_SIZE = HIT_SIZE / 2.f;
        float new_t = t;
        float d = INF;
        Vec3 p = r.at(t);
        int kind = -1; size_t which = 0;
        for (size_t i = 0; i < marched_spheres.size(); i++) { float nd = std::fabs(marched_sdf(marched_spheres[i], p)); if (nd < d) { d = nd; kind = 0; which = i; } }
        for (size_t i = 0; i < marched_boxes.size(); i++)   { float nd = std::fabs(marched_sdf(marched_boxes[i], p));   if (nd < d) { d = nd; kind = 1; which = i; } }
        for (size_t i = 0; i < marched_torus.size(); i++)   { float nd = std::fabs(marched_sdf(marched_torus[i], p));   if (nd < d) { d = nd; kind = 2; which = i; } }
        for (size_t i = 0; i < marched_objects.size(); i++) { float nd = std::fabs(marched_sdf(marched_objects[i], p)); if (nd < d) { d = nd; kind = 3; which = i; } }
        if (cn) cn->sdf_evals += n_marched();
        float aux = d;
        if (kind < 0) return INF;
        int steps = 0;
        while (aux < HIT_SIZE && steps < UNSTUCK_MAX_STEPS) {
            new_t += MIN_STEP_SIZE;
            Vec3 q = r.at(new_t);
            aux = std::fabs(kind == 0 ? marched_sdf(marched_spheres[which], q) : kind == 1 ? marched_sdf(marched_boxes[which], q) : kind == 2 ? marched_sdf(marched_torus[which], q) : marched_sdf(marched_objects[which], q));
            steps++;
            if (cn) cn->sdf_evals++;
        }
        return new_t;
    }

    // hits.rs:270-334.  Also stands for first_hit (see header).
    bool hit(const Ray& r, float t_min, float t_max, HitRecord& rec, Counters* cn = nullptr) const {
        float closest_so_far = t_max;
        bool have = false;
        HitRecord hr;
        for (size_t i = 0; i < spheres.size(); i++)         if (spheres[i].hit(r, t_min, closest_so_far, make_id(T_SPHERE, i), hr))                    { closest_so_far = hr.t; rec = hr; have = true; }
        for (size_t i = 0; i < cubes.size(); i++)           if (cubes[i].hit(r, t_min, closest_so_far, make_id(T_CUBE, i), hr))                        { closest_so_far = hr.t; rec = hr; have = true; }
        for (size_t i = 0; i < triangles.size(); i++)       if (triangles[i].hit(1, r, t_min, closest_so_far, make_id(T_TRIANGLE, i), hr))             { closest_so_far = hr.t; rec = hr; have = true; }
        for (size_t i = 0; i < infinite_planes.size(); i++) if (plane_hit(infinite_planes[i], r, t_min, closest_so_far, make_id(T_INFINITE_PLANE, i), hr)) { closest_so_far = hr.t; rec = hr; have = true; }
        for (size_t i = 0; i < parallelograms.size(); i++)  if (parallelograms[i].hit(0, r, t_min, closest_so_far, make_id(T_PARALLELOGRAM, i), hr))   { closest_so_far = hr.t; rec = hr; have = true; }
        for (size_t i = 0; i < traced_objects.size(); i++)  if (traced_objects[i].hit(r, t_min, closest_so_far, make_id(T_DYN_TRACED, i), hr))                { closest_so_far = hr.t; rec = hr; have = true; }
        if (cn) { cn->segments++; cn->traced_tests += spheres.size() + cubes.size() + triangles.size() + infinite_planes.size() + parallelograms.size() + traced_objects.size(); }

        float t = unstuck(t_min, r, cn);
        if (std::isinf(t)) return have;
        int max_march_iter = 1024;
        while (t < t_max && t < closest_so_far && max_march_iter > 0) {
            max_march_iter -= 1;
            Vec3 point = r.at(t);
            float distance = INF;
            Vec3 normal;
            uint64_t id = 0;
            const Material* material = nullptr;
            for (size_t i = 0; i < marched_spheres.size(); i++) {
                const MarchedSphere& o = marched_spheres[i];
                float d = std::fabs(marched_sdf(o, point));
                if (d < distance) { distance = d; normal = o.get_outward_normal(point); material = &o.material; id = make_id(T_MARCHED_SPHERE, i); if (cn) cn->normal_evals++; }
            }
            for (size_t i = 0; i < marched_boxes.size(); i++) {
                const MarchedBox& o = marched_boxes[i];
                float d = std::fabs(marched_sdf(o, point));
                if (d < distance) { distance = d; normal = o.get_outward_normal(point); material = &o.material; id = make_id(T_MARCHED_BOX, i); if (cn) cn->normal_evals++; }
            }
            for (size_t i = 0; i < marched_torus.size(); i++) {
                const MarchedTorus& o = marched_torus[i];
                float d = std::fabs(marched_sdf(o, point));
                if (d < distance) { distance = d; normal = o.get_outward_normal(point); material = &o.material; id = make_id(T_MARCHED_TORUS, i); if (cn) cn->normal_evals++; }
            }
            for (size_t i = 0; i < marched_objects.size(); i++) {
                const DynMarched& o = marched_objects[i];
                float d = std::fabs(marched_sdf(o, point));
                if (d < distance) { distance = d; normal = o.get_outward_normal(point); material = &o.material; id = make_id(T_DYN_MARCHED, i); if (cn) cn->normal_evals++; }
            }
            if (cn) { cn->march_steps++; cn->sdf_evals += n_marched(); }
            if (!material) return have;
            if (distance < HIT_SIZE) {
                rec.t = t; rec.point = point; rec.normal = normal; rec.material = *material; rec.obj_id = id;
                have = true;
                break;
            } else {
                t += distance;
            }
        }
        return have;
    }
};

// ---- camera.rs -----------------------------------------------------------------------------
struct Camera {
    Vec3 origin, horizontal, vertical, lower_left_corner, u_of_plane, v_of_plane, w_of_plane;
    float lens_radius, aspect_ratio, focus_dist, viewport_width, viewport_height;
    static Camera make(const Vec3& lookfrom, const Vec3& lookat, const Vec3& vup, float vfov_in_degrees, float aspect_ratio, float aperture, float focus_dist) {  // camera.rs:38-59
        float vfov = degrees_to_radians(vfov_in_degrees);
        float height = std::tan(vfov / 2.0f) * focus_dist;
        float viewport_height = 2.0f * height;
        float viewport_width = viewport_height * aspect_ratio;
        Vec3 w = (lookfrom - lookat).unit();
        Vec3 u = vup.cross(w).unit();
        Vec3 v = w.cross(u).unit();
        Vec3 o = lookfrom;
        Vec3 h = viewport_width * u;
        Vec3 vv = viewport_height * v;
        Vec3 llc = o - h / 2.f - vv / 2.f - focus_dist * w;
        Camera c;
        c.origin = o; c.horizontal = h; c.vertical = vv; c.lower_left_corner = llc;
        c.u_of_plane = u; c.v_of_plane = v; c.w_of_plane = w; c.lens_radius = aperture / 2.0f;
        c.aspect_ratio = aspect_ratio; c.focus_dist = focus_dist; c.viewport_width = viewport_width; c.viewport_height = viewport_height;
        return c;
    }
    Mat4x4 uv_to_dir() const {  // camera.rs:67-74
        return Mat4x4::new_4vec_vert(Vec4::new_v3(horizontal), Vec4::new_v3(vertical), Vec4(), Vec4::new_p3(lower_left_corner - origin));
    }
    // camera.rs:60-65 with the lens sample made explicit
    Ray get_ray_with_lens(float u, float v, const Vec3& unit_disc) const {
        Vec3 rand_in_lens = lens_radius * unit_disc;
        Vec3 offset = u_of_plane * rand_in_lens.x() + v_of_plane * rand_in_lens.y();
        Vec3 direction = uv_to_dir().dot(Vec4(u, v, 0.f, 1.f)).xyz();
        return Ray::make(origin + offset, (direction - offset).unit());
    }
    Ray get_ray(float u, float v, Rng& g) const { return get_ray_with_lens(u, v, rand_in_unit_disc(g)); }
};

}  // namespace ref
