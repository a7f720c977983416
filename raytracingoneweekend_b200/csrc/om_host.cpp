// Host-side mirror of the reference's construction code (see om_host.h).  Compiled with
// -ffp-contract=off: these results are scene DATA that must be identical no matter who builds
// them (the Rust host, this mirror, or the test oracle).
#include "om_host.h"

#include <cmath>
#include <math.h>
#include <string.h>

#include "om_math.h"

namespace omh {
using om::V3;
using om::v3;

// ---------------------------------------------------------------------------------------------
void philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c[4] = {ctr[0], ctr[1], ctr[2], ctr[3]};
    uint32_t k[2] = {key[0], key[1]};
    for (int round = 0; round < 10; ++round) {
        const uint64_t lo = 0xD2511F53ull * c[0];
        const uint64_t hi = 0xCD9E8D57ull * c[2];
        const uint32_t t0 = uint32_t(hi >> 32) ^ c[1] ^ k[0];
        const uint32_t t2 = uint32_t(lo >> 32) ^ c[3] ^ k[1];
        c[1] = uint32_t(hi);
        c[3] = uint32_t(lo);
        c[0] = t0;
        c[2] = t2;
        k[0] += 0x9E3779B9u;
        k[1] += 0xBB67AE85u;
    }
    memcpy(out, c, sizeof c);
}
Rng::Rng(uint64_t seed, uint32_t a_, uint32_t b_, uint32_t stream_) : a(a_), b(b_), stream(stream_), draw(0) {
    key[0] = uint32_t(seed);
    key[1] = uint32_t(seed >> 32);
}
uint32_t Rng::next_u32() {
    const uint32_t lane = draw & 3u;
    if (lane == 0) {
        const uint32_t ctr[4] = {draw >> 2, a, b, stream};
        philox4x32_10(ctr, key, buf);
    }
    ++draw;
    return buf[lane];
}
float Rng::rand() { return float(next_u32() >> 8) * 5.9604644775390625e-8f; }  // 2^-24
float Rng::rand_range(float mn, float mx) { return rand() * (mx - mn) + mn; }

// ---------------------------------------------------------------------------------------------
static inline float dot4(const float* r, const float* c) { return r[0] * c[0] + r[1] * c[1] + r[2] * c[2] + r[3] * c[3]; }

void mat4_identity(float o[16]) {
    for (int i = 0; i < 16; i++) o[i] = (i % 5 == 0) ? 1.f : 0.f;
}
void mat4_translate(float o[16], float x, float y, float z) { mat4_identity(o); o[3] = x; o[7] = y; o[11] = z; }
void mat4_scale(float o[16], float x, float y, float z) { mat4_identity(o); o[0] = x; o[5] = y; o[10] = z; }
void mat4_rotate_x(float o[16], float f) { const float c = cosf(f), s = sinf(f); mat4_identity(o); o[5] = c; o[6] = s; o[9] = -s; o[10] = c; }
void mat4_rotate_y(float o[16], float f) { const float c = cosf(f), s = sinf(f); mat4_identity(o); o[0] = c; o[2] = -s; o[8] = s; o[10] = c; }
void mat4_rotate_z(float o[16], float f) { const float c = cosf(f), s = sinf(f); mat4_identity(o); o[0] = c; o[1] = -s; o[4] = s; o[5] = c; }
void mat4_mul(float o[16], const float a[16], const float b[16]) {
    float bt[16], r[16];
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) bt[4 * j + i] = b[4 * i + j];
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) r[4 * i + j] = dot4(a + 4 * i, bt + 4 * j);
    memcpy(o, r, sizeof r);
}
float mat3_determinant(const float m[9]) {
    const float term[6] = {m[0] * m[4] * m[8], m[1] * m[5] * m[6], m[2] * m[3] * m[7], -m[0] * m[5] * m[7], -m[1] * m[3] * m[8], -m[2] * m[4] * m[6]};
    volatile float sum = 0.f, c = 0.f;  // Kahan; volatile pins the evaluation order
    for (int i = 0; i < 6; i++) {
        volatile float y = term[i] - c;
        volatile float t = sum + y;
        c = (t - sum) - y;
        sum = t;
    }
    return sum;
}
void mat3_inverse(float o[9], const float m[9]) {
    const float inv_det = 1.f / mat3_determinant(m);
    const float adj[9] = {m[4] * m[8] - m[5] * m[7], m[2] * m[7] - m[1] * m[8], m[1] * m[5] - m[2] * m[4],
                          m[5] * m[6] - m[3] * m[8], m[0] * m[8] - m[2] * m[6], m[2] * m[3] - m[0] * m[5],
                          m[3] * m[7] - m[4] * m[6], m[1] * m[6] - m[0] * m[7], m[0] * m[4] - m[1] * m[3]};
    for (int i = 0; i < 9; i++) o[i] = adj[i] * inv_det;
}
void mat4_fast_homogenous_inverse(float o[16], const float m[16]) {
    const float lin[9] = {m[0], m[1], m[2], m[4], m[5], m[6], m[8], m[9], m[10]};
    float li[9];
    mat3_inverse(li, lin);
    const float s_inv[16] = {li[0], li[1], li[2], 0.f, li[3], li[4], li[5], 0.f, li[6], li[7], li[8], 0.f, 0.f, 0.f, 0.f, 1.f};
    float t_inv[16];
    mat4_translate(t_inv, -m[3], -m[7], -m[11]);
    mat4_mul(o, s_inv, t_inv);
}

// ---------------------------------------------------------------------------------------------
static inline V3 ld3(const float* p) { return v3(p[0], p[1], p[2]); }
static inline void st3(float* p, V3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }

void camera_new(om_camera* out, const float lookfrom[3], const float lookat[3], const float vup[3], float vfov_deg, float aspect,
                float aperture, float focus_dist) {
    const float vfov = vfov_deg * 3.1415926535897932385f / 180.0f;  // utils.rs:33-35
    const float half_h = tanf(vfov / 2.0f) * focus_dist;
    const float vp_h = 2.0f * half_h;
    const float vp_w = vp_h * aspect;
    const V3 w = om::unit(ld3(lookfrom) - ld3(lookat));
    const V3 u = om::unit(om::cross(ld3(vup), w));
    const V3 v = om::unit(om::cross(w, u));
    const V3 o = ld3(lookfrom);
    const V3 h = vp_w * u;
    const V3 vert = vp_h * v;
    const V3 llc = o - om::div(h, 2.f) - om::div(vert, 2.f) - focus_dist * w;
    st3(out->origin, o); st3(out->horizontal, h); st3(out->vertical, vert); st3(out->lower_left_corner, llc);
    st3(out->u_of_plane, u); st3(out->v_of_plane, v); st3(out->w_of_plane, w);
    out->lens_radius = aperture / 2.0f; out->aspect_ratio = aspect; out->focus_dist = focus_dist;
    out->viewport_width = vp_w; out->viewport_height = vp_h;
}

om_material lambertian(float r, float g, float b) { om_material m = {{r, g, b}, 0.f, 0.f, OM_MAT_LAMBERTIAN}; return m; }
om_material metal(float r, float g, float b, float fuzz) { om_material m = {{r, g, b}, fuzz, 0.f, OM_MAT_METAL}; return m; }
om_material dielectric(float ior) { om_material m = {{0.f, 0.f, 0.f}, 0.f, ior, OM_MAT_DIELECTRIC}; return m; }

void sphere_new(om_sphere* out, const float m[16], const om_material& mat) {
    memcpy(out->m_local_to_world, m, 64);
    mat4_fast_homogenous_inverse(out->m_world_to_local, m);
    out->material = mat;
}
void sphere_new_with_radius(om_sphere* out, const float o[3], float r, const om_material& mat) {
    float t[16], s[16], m[16];
    mat4_translate(t, o[0], o[1], o[2]);
    mat4_scale(s, r, r, r);
    mat4_mul(m, t, s);
    sphere_new(out, m, mat);
}
void cube_new(om_cube* out, const float m[16], const om_material& mat) {
    memcpy(out->m_local_to_world, m, 64);
    mat4_fast_homogenous_inverse(out->m_world_to_local, m);
    out->material = mat;
}
void cube_new_with_length(om_cube* out, const float o[3], float len, const om_material& mat) {
    float t[16], s[16], m[16];
    mat4_translate(t, o[0], o[1], o[2]);
    mat4_scale(s, len, len, len);
    mat4_mul(m, t, s);
    cube_new(out, m, mat);
}
void infinite_plane_new(om_infinite_plane* out, const float c[3], const float n[3], const om_material& mat) {
    st3(out->center, ld3(c));
    st3(out->normal, om::unit(ld3(n)));
    out->material = mat;
}
// traced.rs:135-154
void barycentric_new3points(om_barycentric* out, const float o[3], const float up[3], const float vp[3], const om_material& mat) {
    const V3 origin = ld3(o);
    const V3 u_rel = ld3(up) - origin, v_rel = ld3(vp) - origin;
    const float ul = om::length(u_rel), vl = om::length(v_rel);
    const V3 uu = om::unit(u_rel), vu = om::unit(v_rel);
    const V3 uxv = om::unit(om::cross(uu, vu));
    const V3 uxvxu = om::unit(om::cross(uxv, uu));
    // new_3vec_vert(u,uxv,uxvxu).transpose(): rows are u, uxv, uxvxu
    const float bi[9] = {uu.x, uu.y, uu.z, uxv.x, uxv.y, uxv.z, uxvxu.x, uxvxu.y, uxvxu.z};
    const V3 vib = v3(om::dot(uu, vu), om::dot(uxv, vu), om::dot(uxvxu, vu));
    st3(out->origin, origin); st3(out->u, uu); out->u_length = ul; st3(out->v, vu); out->v_length = vl;
    st3(out->uxv, uxv); st3(out->uxvxu, uxvxu);
    memcpy(out->base_inv, bi, sizeof bi);
    st3(out->v_in_base, vib);
    out->material = mat;
}
// marched.rs:115-130 with mat4x4.rs:125-150
void marched_torus_new(om_marched_torus* out, const float m[16], const float sizes[3], const om_material& mat) {
    const V3 a = v3(m[0], m[4], m[8]), b = v3(m[1], m[5], m[9]), c = v3(m[2], m[6], m[10]);
    const float al = om::length(a), bl = om::length(b), cl = om::length(c);
    const V3 ra = om::div(a, al), rb = om::div(b, bl), rc = om::div(c, cl);
    float t[16], r[16], tr[16];
    mat4_translate(t, m[3], m[7], m[11]);
    const float rot[16] = {ra.x, rb.x, rc.x, 0.f, ra.y, rb.y, rc.y, 0.f, ra.z, rb.z, rc.z, 0.f, 0.f, 0.f, 0.f, 1.f};
    memcpy(r, rot, sizeof r);
    mat4_mul(tr, t, r);
    memcpy(out->m_local_to_world_translate_rotate, tr, 64);
    mat4_fast_homogenous_inverse(out->m_world_to_local_translate_rotate, tr);
    const float sc[4] = {al, bl, cl, 1.f};
    for (int i = 0; i < 4; i++) { out->m_local_to_world_scale[i] = sc[i]; out->m_world_to_local_scale[i] = 1.f / sc[i]; }
    st3(out->sizes, ld3(sizes));
    out->material = mat;
}

// ---------------------------------------------------------------------------------------------
uint64_t scramble(uint64_t id) {
    uint64_t lo = id & 0xFFFFFFFFull, hi = id >> 32;
    lo ^= lo << 13; lo ^= lo >> 7; lo ^= lo << 17;
    hi ^= hi << 13; hi ^= hi >> 17; hi ^= hi << 5;
    return (hi << 32) ^ lo ^ (lo * hi);
}
uint64_t bloom_get_hash(uint64_t id) {
    if (id == 0) return 0;  // `bit` is 0: nothing set (utils.rs:96)
    static const uint64_t K[9][3] = {{456894789ull, 348764781ull, 17287318477382145149ull}, {56456ull, 2345ull, 10520185020478678957ull},
                                     {12337ull, 7878ull, 6100366985798845493ull},          {7438554325ull, 2554ull, 2581451885731034521ull},
                                     {12345ull, 123123044ull, 2015400956511055807ull},     {6373412378ull, 12452ull, 8800267423223100703ull},
                                     {3453453ull, 7874856378ull, 7039701875810786467ull},  {999465ull, 143ull, 3008457310659543551ull},
                                     {14444ull, 111345ull, 5935720376112203207ull}};
    uint64_t mask = 0;
    for (int i = 0; i < 9; i++) mask |= 1ull << (scramble((id * K[i][0] + K[i][1]) % K[i][2]) % 64);
    return mask;
}
void jitter_table(uint32_t spp, uint64_t seed, std::vector<uint8_t>& out) {
    out.resize(spp);
    for (uint32_t s = 0; s < spp; s++) out[s] = uint8_t((((s >> 1) & 1u) << 1) | (s & 1u));  // (jitteri, jitterj)
    Rng g(seed, 0, 0, STREAM_JITTER);
    for (uint32_t i = spp; i > 1; i--) {  // Fisher-Yates from the back
        const uint32_t j = uint32_t((uint64_t(g.next_u32()) * i) >> 32);
        const uint8_t tmp = out[i - 1]; out[i - 1] = out[j]; out[j] = tmp;
    }
}

// ---------------------------------------------------------------------------------------------
// main.rs:37-100 with thread_rng replaced by Philox stream 2.  p_diffuse / p_metal are the
// material thresholds (0.8 / 0.95 in the reference); `specials` adds torus, parallelogram,
// triangle and cube.
static void random_scene(SceneLists& s, uint64_t seed, float p_diffuse, float p_metal, bool specials) {
    const float TWO_PI_STEP = 3.1415926535897932385f;
    Rng g(seed, 0, 0, STREAM_SCENE);
    om_sphere sp;
    const float ground_c[3] = {0.f, -1000.f, 0.f};
    sphere_new_with_radius(&sp, ground_c, 1000.0f, lambertian(0.5f, 0.5f, 0.5f));
    s.spheres.push_back(sp);
    for (int a = -11; a < 11; a++) {
        for (int b = -11; b < 11; b++) {
            const float cx = float(a) + 0.9f * g.rand();
            const float cz = float(b) + 0.9f * g.rand();
            const V3 center = v3(cx, 0.2f, cz);
            if (!(om::length(center - v3(4.f, 0.2f, 0.f)) > 0.9f)) continue;
            om_material mat;
            const float mat_prob = g.rand();
            if (mat_prob < p_diffuse) {
                const float r1 = g.rand(), g1 = g.rand(), b1 = g.rand();
                const float r2 = g.rand(), g2 = g.rand(), b2 = g.rand();
                mat = lambertian(r1 * r2, g1 * g2, b1 * b2);
            } else if (mat_prob < p_metal) {
                const float r1 = g.rand_range(0.5f, 1.f), g1 = g.rand_range(0.5f, 1.f), b1 = g.rand_range(0.5f, 1.f);
                const float fuzz = g.rand_range(0.f, 0.5f);
                mat = metal(r1, g1, b1, fuzz);
            } else {
                mat = dielectric(1.5f);
            }
            float m[16], t[16];
            mat4_translate(m, center.x, center.y, center.z);
            mat4_rotate_x(t, g.rand() * 2.f * TWO_PI_STEP); mat4_mul(m, m, t);
            mat4_rotate_y(t, g.rand() * 2.f * TWO_PI_STEP); mat4_mul(m, m, t);
            mat4_rotate_z(t, g.rand() * 2.f * TWO_PI_STEP); mat4_mul(m, m, t);
            const float sx = g.rand() + 1.f, sy = g.rand() + 1.f, sz = g.rand() + 1.f;
            mat4_scale(t, sx, sy, sz); mat4_mul(m, m, t);
            mat4_scale(t, 0.2f, 0.2f, 0.2f); mat4_mul(m, m, t);
            sphere_new(&sp, m, mat);
            s.spheres.push_back(sp);
        }
    }
    if (!specials) return;
    {
        float m[16], t[16];
        mat4_translate(m, 0.f, 1.f, 0.f);
        mat4_rotate_x(t, 0.6f); mat4_mul(m, m, t);
        mat4_rotate_z(t, 1.33f * 2.f * TWO_PI_STEP); mat4_mul(m, m, t);
        const float sizes[3] = {0.5f, 0.1f, 0.1f};
        om_marched_torus mt;
        marched_torus_new(&mt, m, sizes, dielectric(1.5f));
        s.marched_tori.push_back(mt);
    }
    {
        const float p1[3] = {7.f, 1.f, 0.f}, p2[3] = {6.f, 1.1f, 0.5f}, p3[3] = {6.f, 1.5f, 0.f};
        const float p1b[3] = {7.f + 0.f, 1.f + 0.5f, 0.f + 0.f};
        om_barycentric bc;
        barycentric_new3points(&bc, p1, p2, p3, metal(1.f, 0.5f, 1.f, 0.f));
        s.parallelograms.push_back(bc);
        barycentric_new3points(&bc, p1b, p2, p3, lambertian(1.f, 1.f, 0.f));
        s.triangles.push_back(bc);
    }
    {
        float m[16], t[16];
        mat4_translate(m, 4.f, 1.f, 0.f);
        mat4_rotate_x(t, g.rand() * 2.f * TWO_PI_STEP); mat4_mul(m, m, t);
        mat4_rotate_y(t, g.rand() * 2.f * TWO_PI_STEP); mat4_mul(m, m, t);
        mat4_rotate_z(t, g.rand() * 2.f * TWO_PI_STEP); mat4_mul(m, m, t);
        om_cube cb;
        cube_new(&cb, m, metal(0.7f, 0.6f, 0.5f, 0.f));
        s.cubes.push_back(cb);
    }
}

static void basic_scene(SceneLists& s) {  // main.rs:102-110
    const float c[3][3] = {{0.f, 0.f, -2.f}, {-2.f, 0.f, -2.f}, {2.f, 0.f, -2.f}};
    for (int i = 0; i < 3; i++) {
        om_sphere sp;
        sphere_new_with_radius(&sp, c[i], 1.f, lambertian(0.5f, 0.5f, 0.5f));
        s.spheres.push_back(sp);
    }
}

static om_material pick_material(Rng& g, float p_diffuse, float p_metal) {
    const float mat_prob = g.rand();
    if (mat_prob < p_diffuse) {
        const float r1 = g.rand(), g1 = g.rand(), b1 = g.rand();
        const float r2 = g.rand(), g2 = g.rand(), b2 = g.rand();
        return lambertian(r1 * r2, g1 * g2, b1 * b2);
    }
    if (mat_prob < p_metal) {
        const float r1 = g.rand_range(0.5f, 1.f), g1 = g.rand_range(0.5f, 1.f), b1 = g.rand_range(0.5f, 1.f);
        return metal(r1, g1, b1, g.rand_range(0.f, 0.5f));
    }
    return dielectric(1.5f);
}

// the reference ships no marched scene (main.rs:40,76 show the intent): a
// marched ground sphere plus an 8x8 grid alternating marched sphere / box / torus.
static void marched_scene(SceneLists& s, uint64_t seed) {
    Rng g(seed, 0, 0, STREAM_SCENE);
    om_marched_sphere ground = {{0.f, -1000.f, 0.f}, 1000.f, lambertian(0.5f, 0.5f, 0.5f)};
    s.marched_spheres.push_back(ground);
    int k = 0;
    for (int a = 0; a < 8; a++) {
        for (int b = 0; b < 8; b++, k++) {
            const float cx = -8.f + 2.f * float(a) + 0.6f * g.rand();
            const float cz = -8.f + 2.f * float(b) + 0.6f * g.rand();
            const om_material mat = pick_material(g, 0.60f, 0.85f);
            if (k % 3 == 0) {
                om_marched_sphere o = {{cx, 0.4f, cz}, 0.4f, mat};
                s.marched_spheres.push_back(o);
            } else if (k % 3 == 1) {
                om_marched_box o = {{cx, 0.35f, cz}, {0.35f, 0.35f, 0.35f}, mat};
                s.marched_boxes.push_back(o);
            } else {
                float m[16], t[16];
                mat4_translate(m, cx, 0.55f, cz);
                mat4_rotate_x(t, g.rand() * 2.f * 3.1415926535897932385f); mat4_mul(m, m, t);
                mat4_rotate_z(t, g.rand() * 2.f * 3.1415926535897932385f); mat4_mul(m, m, t);
                const float sizes[3] = {0.4f, 0.12f, 0.12f};
                om_marched_torus o;
                marched_torus_new(&o, m, sizes, mat);
                s.marched_tori.push_back(o);
            }
        }
    }
}

// n spheres, radius U[0.02,0.08], in a 200 x 4 x 200 slab on the ground sphere
static void many_spheres_scene(SceneLists& s, uint64_t seed, uint64_t n) {
    Rng g(seed, 0, 0, STREAM_SCENE);
    om_sphere sp;
    const float ground_c[3] = {0.f, -1000.f, 0.f};
    sphere_new_with_radius(&sp, ground_c, 1000.0f, lambertian(0.5f, 0.5f, 0.5f));
    s.spheres.push_back(sp);
    s.spheres.reserve(size_t(n) + 1);
    for (uint64_t i = 0; i < n; i++) {
        const float r = g.rand_range(0.02f, 0.08f);
        const float c[3] = {g.rand_range(-100.f, 100.f), g.rand_range(0.1f, 4.f), g.rand_range(-100.f, 100.f)};
        sphere_new_with_radius(&sp, c, r, pick_material(g, 0.80f, 0.95f));
        s.spheres.push_back(sp);
    }
}

bool generate(SceneLists& s, uint32_t kind, uint64_t seed, uint64_t n_param) {
    switch (kind) {
        case 0: random_scene(s, seed, 0.8f, 0.95f, true); return true;
        case 1: basic_scene(s); return true;
        case 2: random_scene(s, seed, 0.8f, 0.95f, false); return true;
        case 3: marched_scene(s, seed); return true;
        case 4: random_scene(s, seed, 0.25f, 0.40f, true); return true;
        case 5: many_spheres_scene(s, seed, n_param); return true;
        default: return false;
    }
}

// om_scene_add_dyn_marched: the program must keep the operand stack within [0, OM_SDF_STACK], end with one value,
// name existing registers and read them only after a write.  The device interpreter (om_device.cuh, sdf_program) and
// the oracle's rely on this check instead of guarding every step.
const char* check_dyn_traced(const om_dyn_traced& o) {
    for (int i = 0; i < 16; i++) if (!std::isfinite(o.m_local_to_world[i]) || !std::isfinite(o.m_world_to_local[i])) return "non-finite matrix entry";
    bool any = false;
    for (int i = 0; i < 10; i++) { if (!std::isfinite(o.q[i])) return "non-finite quadric coefficient"; if (i < 9 && o.q[i] != 0.f) any = true; }
    if (!any) return "the quadric has no term in x, y, z";
    for (int a = 0; a < 3; a++) {
        if (std::isnan(o.clip_lo[a]) || std::isnan(o.clip_hi[a]) || o.clip_lo[a] > o.clip_hi[a]) return "clip_lo > clip_hi (or NaN)";
        if (std::isnan(o.bounds_lo[a]) || std::isnan(o.bounds_hi[a]) || o.bounds_lo[a] > o.bounds_hi[a]) return "bounds_lo > bounds_hi (or NaN)";
    }
    return "";
}
const char* check_sdf_program(const uint32_t* code, size_t n_words) {
    if (n_words == 0 || n_words > OM_SDF_MAX_WORDS) return "program must have 1..OM_SDF_MAX_WORDS words";
    int depth = 0;
    bool written[OM_SDF_REGS] = {false, false, false, false};
    for (size_t pc = 0; pc < n_words; pc++) {
        const uint32_t op = code[pc] & 0xFFu, arg = code[pc] >> 8;
        int need = 0, delta = 0;
        switch (op) {
            case OM_SDF_X: case OM_SDF_Y: case OM_SDF_Z: delta = 1; break;
            case OM_SDF_CONST: delta = 1; if (++pc >= n_words) return "OM_SDF_CONST without its value word"; break;
            case OM_SDF_ADD: case OM_SDF_SUB: case OM_SDF_MUL: case OM_SDF_DIV: case OM_SDF_MIN: case OM_SDF_MAX: case OM_SDF_LEN2: need = 2; delta = -1; break;
            case OM_SDF_LEN3: need = 3; delta = -2; break;
            case OM_SDF_NEG: case OM_SDF_ABS: case OM_SDF_SQRT: need = 1; break;
            case OM_SDF_STORE: need = 1; delta = -1; if (arg >= OM_SDF_REGS) return "register out of range"; written[arg] = true; break;
            case OM_SDF_LOAD: delta = 1; if (arg >= OM_SDF_REGS) return "register out of range"; if (!written[arg]) return "register read before it is written"; break;
            default: return "unknown opcode";
        }
        if (op != OM_SDF_STORE && op != OM_SDF_LOAD && op != OM_SDF_CONST && arg != 0u) return "argument bits set on an opcode that takes none";
        if (op == OM_SDF_CONST && arg != 0u) return "argument bits set on an opcode that takes none";
        if (depth < need) return "operand stack underflow";
        depth += delta;
        if (depth > OM_SDF_STACK) return "operand stack overflow";
    }
    return depth == 1 ? "" : "program must leave exactly one value on the stack";
}

}  // namespace omh
