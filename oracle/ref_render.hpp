// ORACLE — TEST INFRASTRUCTURE ONLY (see ref_math.hpp header).  PARITY UNPINNED.
//
// Follows: src/render_thread.rs:9-40 (Stats), :105-143 (handle_hit, ray_color), :145-202 (render),
//          src/utils.rs:8-17 (lerp, normalize_color), :45-108 (scramble, u64_to_color, BloomFilter),
//          src/main.rs:37-110 (random_scene, basic_scene), :172-189 (pixel -> thread map).
#pragma once
#include <atomic>
#include <thread>
#include <vector>

#include "ref_scene.hpp"

namespace ref {

inline Vec3 lerp(float t, const Vec3& c1, const Vec3& c2) { return (1.0f - t) * c1 + t * c2; }  // utils.rs:9-11
inline Vec3 normalize_color(const Vec3& c) {                                                    // utils.rs:12-17
    return Vec3(clamp(std::sqrt(c.x()), 0.f, 0.999f), clamp(std::sqrt(c.y()), 0.f, 0.999f), clamp(std::sqrt(c.z()), 0.f, 0.999f));
}
inline void to_u8x3(const Vec3& c, uint8_t out[3]) {  // vec3.rs:73-76
    out[0] = f32_as_u8(c.e[0] * 256.0f); out[1] = f32_as_u8(c.e[1] * 256.0f); out[2] = f32_as_u8(c.e[2] * 256.0f);
}

inline uint64_t scramble(uint64_t obj_id) {  // utils.rs:45-56
    uint64_t id1 = obj_id & 0xFFFFFFFFull;
    id1 ^= id1 << 13; id1 ^= id1 >> 7; id1 ^= id1 << 17;
    uint64_t id2 = obj_id >> 32;
    id2 ^= id2 << 13; id2 ^= id2 >> 17; id2 ^= id2 << 5;
    return (id2 << 32) ^ id1 ^ (id1 * id2);
}
inline void u64_to_color(uint64_t id, uint8_t out[3]) {  // utils.rs:58-69
    uint8_t b[8];
    for (int i = 0; i < 8; i++) b[i] = (uint8_t)((id >> (8 * i)) & 0xFF);
    out[0] = b[0] ^ b[7] ^ b[3]; out[1] = b[1] ^ b[4] ^ b[5]; out[2] = b[2] ^ b[6];
}
inline uint64_t wrap_u64_madd(uint64_t x, uint64_t m, uint64_t a) { return x * m + a; }  // utils.rs:71-73
inline uint64_t bloom_get_hash(uint64_t id) {                                            // utils.rs:93-107
    uint64_t ret = 0;
    uint64_t bit = (id != 0) ? 1ull : 0ull;
    ret |= bit << (scramble(wrap_u64_madd(id, 456894789ull, 348764781ull) % 17287318477382145149ull) % 64);
    ret |= bit << (scramble(wrap_u64_madd(id, 56456ull, 2345ull) % 10520185020478678957ull) % 64);
    ret |= bit << (scramble(wrap_u64_madd(id, 12337ull, 7878ull) % 6100366985798845493ull) % 64);
    ret |= bit << (scramble(wrap_u64_madd(id, 7438554325ull, 2554ull) % 2581451885731034521ull) % 64);
    ret |= bit << (scramble(wrap_u64_madd(id, 12345ull, 123123044ull) % 2015400956511055807ull) % 64);
    ret |= bit << (scramble(wrap_u64_madd(id, 6373412378ull, 12452ull) % 8800267423223100703ull) % 64);
    ret |= bit << (scramble(wrap_u64_madd(id, 3453453ull, 7874856378ull) % 7039701875810786467ull) % 64);
    ret |= bit << (scramble(wrap_u64_madd(id, 999465ull, 143ull) % 3008457310659543551ull) % 64);
    ret |= bit << (scramble(wrap_u64_madd(id, 14444ull, 111345ull) % 5935720376112203207ull) % 64);
    return ret;
}

struct Stats {  // render_thread.rs:9-40
    uint32_t n = 0;
    Vec3 sum;
    uint8_t color[3] = {0, 0, 0};
    uint32_t bad_avgs = 0;
    float avg_depth = 0.f;
    uint64_t bloom = 0;
    bool add(const Vec3& x, float depth, uint64_t obj_id) {
        uint8_t old_color[3] = {color[0], color[1], color[2]};
        sum = sum + x;
        n += 1;
        Vec3 avg = sum / (float)n;
        to_u8x3(normalize_color(avg), color);
        bool bad_run = old_color[0] == color[0] && old_color[1] == color[1] && old_color[2] == color[2];
        bad_avgs += (uint32_t)bad_run;
        bad_avgs *= (uint32_t)bad_run;
        avg_depth = (((float)n - 1.f) * avg_depth + depth) / (float)n;
        bloom |= bloom_get_hash(obj_id);
        return bad_avgs >= 5;
    }
};

struct SampleResult { Vec3 color; float depth; uint64_t obj_id; };

// render_thread.rs:105-143.  first_hit == hit (see ref_scene.hpp header).
inline SampleResult ray_color(const Ray& r, const World& world, uint32_t depth, float tmin, float tmax, Rng& g, World::Counters* cn = nullptr) {
    Vec3 curr_color(1.f, 1.f, 1.f);
    Ray curr_ray = r;
    float depthf = INF;
    uint64_t obj_id = 0;
    for (uint32_t i = 0; i == 0 || i < depth; i++) {  // segment 0, then `for _i in 1..depth`
        HitRecord hr;
        if (world.hit(curr_ray, tmin, tmax, hr, cn)) {  // handle_hit Some(hr)
            ScatterResult rs = scatter(hr.material, curr_ray, hr, g);
            curr_color = curr_color * rs.attenuation;
            curr_ray = rs.ray;
            if (i == 0) { depthf = hr.t; obj_id = hr.obj_id; }
        } else {  // None
            float t = 0.5f * (curr_ray.dir.y() + 1.0f);
            curr_color = curr_color * lerp(t, Vec3(1.f, 1.f, 1.f), Vec3(0.5f, 0.7f, 1.0f));
            if (i == 0) return {curr_color, INF, 0};
            return {curr_color, depthf, obj_id};
        }
    }
    return {-Vec3(), depthf, obj_id};  // out of depth: -black
}

// render_thread.rs:164-174: (0,0),(0,1),(1,0),(1,1),... shuffled.  One table per render (the
// reference shuffles one per thread with thread_rng; the multiset is the same).
// code bit1 = jitter i, bit0 = jitter j.  Fisher-Yates from the back like rand's shuffle.
inline std::vector<uint8_t> make_jitter_table(uint32_t spp, uint64_t seed) {
    std::vector<uint8_t> t(spp);
    for (uint32_t s = 0; s < spp; s++) t[s] = (uint8_t)((((s / 2) & 1u) << 1) | (s & 1u));
    Rng g(seed, 0, 0, STREAM_JITTER);
    for (uint32_t i = spp; i > 1; i--) {
        uint32_t j = (uint32_t)(((uint64_t)g.next_u32() * (uint64_t)i) >> 32);
        uint8_t tmp = t[i - 1]; t[i - 1] = t[j]; t[j] = tmp;
    }
    return t;
}

struct RenderParams {
    uint32_t width, height, spp_max, max_depth;
    float tmin, tmax;
    uint64_t seed;
    bool adaptive;
    uint32_t sample_base = 0;   // spp-slice: global index of this handle's first sample
    uint32_t sample_count = 0;  // samples this handle owns per pixel (set by the caller, > 0)
    uint32_t part_rank = 0, part_world = 1;
};

// primary ray of (pixel, sample) exactly as render() builds it (render_thread.rs:183-192)
inline Ray primary_ray(const Camera& cam, const RenderParams& p, const std::vector<uint8_t>& jit, uint32_t pxl_idx, uint32_t sample, Rng& g) {
    uint32_t line = pxl_idx / p.width;
    uint32_t col = pxl_idx - p.width * line;
    float j_f = (float)line, i_f = (float)col;
    uint8_t code = jit[sample];
    float i_rand = (g.rand() + (float)((code >> 1) & 1)) / 2.f;
    float j_rand = (g.rand() + (float)(code & 1)) / 2.f;
    float u = (i_f + i_rand) / ((float)p.width - 1.f);
    float v = 1.0f - (j_f + j_rand) / ((float)p.height - 1.f);
    return cam.get_ray(u, v, g);
}
// pixel-centre ray without lens offset (deterministic; used for the primary-hit parity maps)
inline Ray center_ray(const Camera& cam, const RenderParams& p, uint32_t pxl_idx) {
    uint32_t line = pxl_idx / p.width;
    uint32_t col = pxl_idx - p.width * line;
    float u = ((float)col + 0.5f) / ((float)p.width - 1.f);
    float v = 1.0f - ((float)line + 0.5f) / ((float)p.height - 1.f);
    return cam.get_ray_with_lens(u, v, Vec3(0.f, 0.f, 0.f));
}

// main.rs:172-189: pixel -> thread, 2730-pixel chunks round-robin
inline uint32_t assigned_thread(uint32_t pxl_idx, uint32_t nthreads) { return (pxl_idx / 2730u) % nthreads; }

// One thread's share of render() (render_thread.rs:145-202) for `spp_pass` more samples.
// `pixels` is the shared Vec<Pixel>; sumsq (optional) accumulates per-channel squares for the
// variance the 3-sigma image test needs.
inline void render_thread(const Camera& cam, const World& world, const RenderParams& p, const std::vector<uint8_t>& jit,
                          Stats* pixels, double* sumsq, uint32_t spp_pass, uint32_t tid, uint32_t nthreads,
                          std::atomic<uint64_t>& samples_atom, std::atomic<uint64_t>& traced_atom, World::Counters* cn) {
    const uint32_t image_size = p.width * p.height;
    std::vector<uint32_t> list, back;  // ThreadPixels: list + backbuff_list (render_thread.rs:68-102)
    for (uint32_t pos = 0; pos < image_size; pos++) {
        if (p.part_world > 1 && (pos / 2730u) % p.part_world != p.part_rank) continue;
        if (assigned_thread(pos, nthreads) != tid) continue;
        const Stats& st = pixels[pos];
        bool retired = (p.adaptive && st.bad_avgs >= 5) || st.n >= p.sample_count;
        if (!retired) list.push_back(pos);
    }
    back.reserve(list.size());
    uint64_t traced = 0;
    for (uint32_t s = 0; s < spp_pass && !list.empty(); s++) {
        back.clear();
        for (size_t idx = 0; idx < list.size(); idx++) {
            uint32_t pxl_idx = list[idx];
            Stats& st = pixels[pxl_idx];
            uint32_t sample = p.sample_base + st.n;  // jitters[pixel.stats.n], :188-189
            Rng g(p.seed, pxl_idx, sample, STREAM_PATH);
            Ray ray = primary_ray(cam, p, jit, pxl_idx, sample, g);
            SampleResult r = ray_color(ray, world, p.max_depth, p.tmin, p.tmax, g, cn);
            bool done = st.add(r.color, r.depth, r.obj_id);
            if (sumsq) for (int c = 0; c < 3; c++) sumsq[3 * (size_t)pxl_idx + c] += (double)r.color.e[c] * (double)r.color.e[c];
            traced++;
            if (!p.adaptive) done = false;  // fixed-spp mode (not in the reference)
            uint32_t left = p.sample_count - st.n;
            uint64_t log_samples = (uint64_t)(done ? left : 0u) + 1u;  // render_thread.rs:196
            samples_atom.fetch_add(log_samples, std::memory_order_relaxed);
            if (!done && left > 0) back.push_back(pxl_idx);  // ThreadPixels::add_run :97-101
        }
        list.swap(back);  // swap_buffers :92-96
    }
    traced_atom.fetch_add(traced, std::memory_order_relaxed);
}

inline void render(const Camera& cam, const World& world, const RenderParams& p, Stats* pixels, double* sumsq,
                   uint32_t spp_pass, uint32_t nthreads, uint64_t* credited, uint64_t* traced, World::Counters* cn_total) {
    if (nthreads == 0) nthreads = 1;
    std::vector<uint8_t> jit = make_jitter_table(p.spp_max, p.seed);
    std::atomic<uint64_t> samples_atom(0), traced_atom(0);
    std::vector<World::Counters> cns(nthreads);
    std::vector<std::thread> th;
    for (uint32_t t = 0; t < nthreads; t++)
        th.emplace_back([&, t]() { render_thread(cam, world, p, jit, pixels, sumsq, spp_pass, t, nthreads, samples_atom, traced_atom, cn_total ? &cns[t] : nullptr); });
    for (auto& x : th) x.join();
    if (credited) *credited += samples_atom.load();
    if (traced) *traced += traced_atom.load();
    if (cn_total)
        for (auto& c : cns) { cn_total->segments += c.segments; cn_total->traced_tests += c.traced_tests; cn_total->sdf_evals += c.sdf_evals; cn_total->normal_evals += c.normal_evals; cn_total->march_steps += c.march_steps; }
}

// ---- main.rs:37-110 scene builders (thread_rng -> Philox stream 2) ---------------------------
inline void random_scene(World& world, uint64_t seed) {
    Rng g(seed, 0, 0, STREAM_SCENE);
    Material mat_ground = Material::new_lambertian(Vec3(0.5f, 0.5f, 0.5f));
    world.spheres.push_back(Sphere::new_with_radius(Vec3(0.f, -1000.f, 0.f), 1000.0f, mat_ground));
    for (int a = -11; a < 11; a++) {
        float af = (float)a;
        for (int b = -11; b < 11; b++) {
            float bf = (float)b;
            float cx = af + 0.9f * g.rand();
            float cz = bf + 0.9f * g.rand();
            Vec3 center(cx, 0.2f, cz);
            bool add_to_world = (center - Vec3(4.f, 0.2f, 0.f)).length() > 0.9f;
            if (add_to_world) {
                Material sphere_mat;
                float mat_prob = g.rand();
                if (mat_prob < 0.8f) {
                    Vec3 c1 = vec3_rand(g);
                    Vec3 c2 = vec3_rand(g);
                    sphere_mat = Material::new_lambertian(c1 * c2);
                } else if (mat_prob < 0.95f) {
                    Vec3 albedo = vec3_rand_range(g, 0.5f, 1.f);
                    float fuzz = g.rand_range(0.f, 0.5f);
                    sphere_mat = Material::new_metal_fuzz(albedo, fuzz);
                } else {
                    sphere_mat = Material::new_dielectric(1.5f);
                }
                Mat4x4 m = Mat4x4::new_translate(center);
                m = m.dot_mat(Mat4x4::new_rotate_x(g.rand() * 2.f * PI));
                m = m.dot_mat(Mat4x4::new_rotate_y(g.rand() * 2.f * PI));
                m = m.dot_mat(Mat4x4::new_rotate_z(g.rand() * 2.f * PI));
                float sx = g.rand() + 1.f, sy = g.rand() + 1.f, sz = g.rand() + 1.f;
                m = m.dot_mat(Mat4x4::new_scale(Vec3(sx, sy, sz)));
                m = m.dot_mat(Mat4x4::new_scale(Vec3(0.2f, 0.2f, 0.2f)));
                world.spheres.push_back(Sphere::make(m, sphere_mat));
            }
        }
    }
    {
        Material mat = Material::new_dielectric(1.5f);
        Mat4x4 l2w = Mat4x4::new_translate(Vec3(0.f, 1.f, 0.f)).dot_mat(Mat4x4::new_rotate_x(0.6f)).dot_mat(Mat4x4::new_rotate_z(1.33f * 2.f * PI));
        world.marched_torus.push_back(MarchedTorus::make(l2w, Vec3(0.5f, 0.1f, 0.1f), mat));
    }
    {
        Vec3 p1(7.f, 1.f, 0.f), p2(6.f, 1.1f, 0.5f), p3(6.f, 1.5f, 0.f);
        world.parallelograms.push_back(Barycentric::new3points(p1, p2, p3, Material::new_metal(Vec3(1.f, 0.5f, 1.f))));
        world.triangles.push_back(Barycentric::new3points(p1 + Vec3(0.f, 0.5f, 0.f), p2, p3, Material::new_lambertian(Vec3(1.f, 1.f, 0.f))));
    }
    {
        Material mat = Material::new_metal(Vec3(0.7f, 0.6f, 0.5f));
        Mat4x4 m = Mat4x4::new_translate(Vec3(4.f, 1.f, 0.f));
        m = m.dot_mat(Mat4x4::new_rotate_x(g.rand() * 2.f * PI));
        m = m.dot_mat(Mat4x4::new_rotate_y(g.rand() * 2.f * PI));
        m = m.dot_mat(Mat4x4::new_rotate_z(g.rand() * 2.f * PI));
        world.cubes.push_back(Cube::make(m, mat));
    }
}

inline void basic_scene(World& world) {  // main.rs:102-110
    Material mat_ground = Material::new_lambertian(Vec3(0.5f, 0.5f, 0.5f));
    world.spheres.push_back(Sphere::new_with_radius(Vec3(0.f, 0.f, -2.f), 1.f, mat_ground));
    world.spheres.push_back(Sphere::new_with_radius(Vec3(-2.f, 0.f, -2.f), 1.f, mat_ground));
    world.spheres.push_back(Sphere::new_with_radius(Vec3(2.f, 0.f, -2.f), 1.f, mat_ground));
}

}  // namespace ref
