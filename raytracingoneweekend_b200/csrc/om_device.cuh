// Device code of the B200 renderer: everything under render_thread::render
// (reference src/render_thread.rs:145-202) as CUDA for sm_100a.
//
// This header is compiled twice, into two namespaces:
//   om_exact  (om_kernels_exact.cu, nvcc -fmad=false): no FMA contraction, IEEE div/sqrt, the
//             reference's operation order -> bit-identical to the CPU oracle, sample by sample.
//             This is the DEFAULT build; explicit fmaf() is used only in culling code.
//   om_fast   (om_kernels_fast.cu, OM_RENDER_FAST_MATH): the same expressions with FMA contraction
//             everywhere; statistically equivalent, not bit-identical.
// OM_NS and OM_EXACT are set by the including .cu file.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ottomarcher.h"
#include "om_args.h"
#include "om_dev_types.h"
#include "om_math.h"

#ifndef OM_EXACT
#error "include from om_kernels_exact.cu / om_kernels_fast.cu"
#endif

namespace OM_NS {

using om::V3;
using om::v3;
using omd::SceneHeader;

constexpr bool kExact = (OM_EXACT != 0);
constexpr float kInf = __builtin_huge_valf();
constexpr float HIT_SIZE = 0.001f;  // hits.rs:113
constexpr int UNSTUCK_MAX_STEPS = 1 << 20;
constexpr int CTA_THREADS = 256;
#ifndef OM_RENDER_THREADS
#define OM_RENDER_THREADS 1024  // one CTA per SM: one copy of the scene image, 32 warps (64 registers) to hide the
#define OM_RENDER_MIN_CTAS 1    // instruction-cache misses of this 70 KB kernel: +23 % over 2 x 256 threads at 92 registers
#endif
constexpr int RENDER_THREADS = OM_RENDER_THREADS;  // k_render only

using omk::RenderArgs;
using omk::TraceArgs;
using namespace omk::ctr;

// ============================================================================================
// Philox4x32-10 counter RNG: key = seed, counter = (block, pixel, sample, stream).  Replaces
// rand::thread_rng (utils.rs:24-27).  Same contract as omh::Rng / the oracle's ref::Rng.
// ============================================================================================
struct Rng {
    const uint32_t* rk;  // the launch's round keys (RenderArgs::rk, kernel parameter space: they fold into the XORs as constant-bank operands)
    uint32_t a, b, draw;
    uint32_t b0, b1, b2, b3;
};
// Inlined at a handful of draw sites that the warp reaches CONVERGED (see rng_take): the first
// kernels called an out-of-line copy from every scalar draw, lane by lane, and spent 12 % of all
// issued instructions there at 3.7 active lanes.
#ifndef OM_PHILOX_INLINE
#define OM_PHILOX_INLINE 1  // 0: one out-of-line copy shared by the draw sites (experiment)
#endif
#if OM_PHILOX_INLINE
__device__ __forceinline__
#else
__device__ __noinline__
#endif
uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, const uint32_t* rk) {
    uint32_t c3 = 0u;  // STREAM_PATH
#pragma unroll
    for (int r = 0; r < 10; r++) {
        const uint32_t lo0 = 0xD2511F53u * c0, hi0 = __umulhi(0xD2511F53u, c0);
        const uint32_t lo1 = 0xCD9E8D57u * c2, hi1 = __umulhi(0xCD9E8D57u, c2);
        c0 = hi1 ^ c1 ^ rk[2 * r]; c1 = lo1; c2 = hi0 ^ c3 ^ rk[2 * r + 1]; c3 = lo0;  // key schedule precomputed by the host
    }
    return make_uint4(c0, c1, c2, c3);
}
__device__ __forceinline__ void rng_init(Rng& g, const uint32_t* rk, uint32_t pixel, uint32_t sample) {
    g.rk = rk; g.a = pixel; g.b = sample; g.draw = 0;
}
__device__ __forceinline__ float u32_to_f32(uint32_t u) { return (float)(u >> 8) * 5.9604644775390625e-8f; }  // utils.rs:25: uniform [0,1), 24 bits
__device__ __forceinline__ uint32_t sel4(uint32_t i, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    return i == 0u ? a : (i == 1u ? b : (i == 2u ? c : d));
}
__device__ __forceinline__ float sel4f(uint32_t i, float4 v) { return (i & 2u) ? ((i & 1u) ? v.w : v.z) : ((i & 1u) ? v.y : v.x); }
// The next N (<= 4) draws of the stream, i.e. words draw .. draw+N-1 of the Philox block sequence
// (word i lives in block i/4), with at most ONE block computation that every calling lane performs
// together: the cached block serves the words before the next block boundary, the new block the
// rest.  Same values in the same order as N calls of the oracle's next_f32().
template <int N>
__device__ __forceinline__ void rng_take(Rng& g, float* out) {
    const uint32_t pos = g.draw & 3u;
    const bool need = pos == 0u || pos + (uint32_t)N > 4u;
    uint4 nb = make_uint4(g.b0, g.b1, g.b2, g.b3);
    if (need) nb = philox4x32_10((g.draw + (uint32_t)N - 1u) >> 2, g.a, g.b, g.rk);
    // words pos .. pos+N-1 of the window {cached block (the new one when pos == 0), new block}: a two-level barrel shift
    // by pos (the first version picked every word through two 4-way select chains: 57 instructions for N = 3)
    static_assert(N >= 1 && N <= 3, "the window below holds words 0..6");
    const bool z = pos == 0u, s2 = (pos & 2u) != 0u, s1 = (pos & 1u) != 0u;
    uint32_t x[7] = {z ? nb.x : g.b0, z ? nb.y : g.b1, z ? nb.z : g.b2, z ? nb.w : g.b3, nb.x, nb.y, nb.z};
    uint32_t y[N + 1];
#pragma unroll
    for (int i = 0; i <= N; i++) y[i] = s2 ? x[i + 2] : x[i];
#pragma unroll
    for (int i = 0; i < N; i++) out[i] = u32_to_f32(s1 ? y[i + 1] : y[i]);
    g.b0 = nb.x; g.b1 = nb.y; g.b2 = nb.z; g.b3 = nb.w;
    g.draw += (uint32_t)N;
}
__device__ __forceinline__ float rng_f32(Rng& g) { float v; rng_take<1>(g, &v); return v; }
__device__ __forceinline__ float to_range(float u, float mn, float mx) { return u * (mx - mn) + mn; }  // utils.rs:26
__device__ __forceinline__ V3 rand_in_unit_sphere(Rng& g) {  // vec3.rs:92-97
    for (;;) {
        float u[3];
        rng_take<3>(g, u);
        const V3 p = v3(to_range(u[0], -1.f, 1.f), to_range(u[1], -1.f, 1.f), to_range(u[2], -1.f, 1.f));
        if (om::length_squared(p) < 1.f) return p;
    }
}
__device__ __forceinline__ V3 rand_in_unit_disc(Rng& g) {  // vec3.rs:108-113
    for (;;) {
        float u[2];
        rng_take<2>(g, u);
        const V3 p = v3(to_range(u[0], -1.f, 1.f), to_range(u[1], -1.f, 1.f), 0.f);
        if (om::length_squared(p) < 1.f) return p;
    }
}

// ============================================================================================
// scene access: hot image in shared memory (SMEM) or global memory; cold image in global
// ============================================================================================
extern __shared__ float4 om_smem[];

template <bool SMEM>
struct SceneView {
    const float4* hot_g;
    const float4* cold_g;
    __device__ __forceinline__ float4 hot(uint32_t u) const {
        if (SMEM) return om_smem[u];
        return __ldg(hot_g + u);
    }
    __device__ __forceinline__ float4 cold(uint32_t u) const { return __ldg(cold_g + u); }
    __device__ __forceinline__ uint32_t cold_word(uint32_t u, uint32_t i) const { return __ldg(reinterpret_cast<const uint32_t*>(cold_g + u) + i); }
    __device__ __forceinline__ uint32_t hot_word(uint32_t u, uint32_t i) const {  // 32-bit word i of the section starting at unit u
        if (SMEM) return reinterpret_cast<const uint32_t*>(om_smem + u)[i];
        return __ldg(reinterpret_cast<const uint32_t*>(hot_g + u) + i);
    }
    __device__ __forceinline__ uint32_t hot_byte(uint32_t u, uint32_t i) const {  // byte i of the section starting at unit u
        if (SMEM) return reinterpret_cast<const uint8_t*>(om_smem + u)[i];
        return __ldg(reinterpret_cast<const uint8_t*>(hot_g + u) + i);
    }
};

struct Ray {
    V3 o, d;
};
__device__ __forceinline__ V3 ray_at(const Ray& r, float t) { return r.o + t * r.d; }  // ray.rs:14-16

// Mat4x4::dot_p3 / dot_v3 over the three stored rows (mat4x4.rs:49-57)
__device__ __forceinline__ V3 xf_point(float4 r0, float4 r1, float4 r2, V3 p) {
    return v3(r0.x * p.x + r0.y * p.y + r0.z * p.z + r0.w, r1.x * p.x + r1.y * p.y + r1.z * p.z + r1.w, r2.x * p.x + r2.y * p.y + r2.z * p.z + r2.w);
}
__device__ __forceinline__ V3 xf_vec(float4 r0, float4 r1, float4 r2, V3 p) {
    if (kExact)  // the reference multiplies the translation column by w = 0 and adds it
        return v3(r0.x * p.x + r0.y * p.y + r0.z * p.z + r0.w * 0.f, r1.x * p.x + r1.y * p.y + r1.z * p.z + r1.w * 0.f, r2.x * p.x + r2.y * p.y + r2.z * p.z + r2.w * 0.f);
    return v3(r0.x * p.x + r0.y * p.y + r0.z * p.z, r1.x * p.x + r1.y * p.y + r1.z * p.z, r2.x * p.x + r2.y * p.y + r2.z * p.z);
}

// ---- traced primitives: parametric distance only (the hit record is built once, for the winner) --
// Sphere::hit traced.rs:38-56
__device__ __forceinline__ bool sphere_t(float4 r0, float4 r1, float4 r2, const Ray& r, float t_min, float t_max, float& t_out) {
    const V3 ol = xf_point(r0, r1, r2, r.o), dl = xf_vec(r0, r1, r2, r.d);
    const float a = om::length_squared(dl);
    const float half_b = om::dot(ol, dl);
    const float c = om::length_squared(ol) - 1.f;
    const float discriminant = half_b * half_b - a * c;
    if (discriminant < 0.f) return false;
    const float sqrtd = sqrtf(discriminant);
    float root = (-half_b - sqrtd) / a;
    if (root < t_min || root > t_max) {
        root = (-half_b + sqrtd) / a;
        if (root < t_min || root > t_max) return false;
    }
    t_out = root;
    return true;
}
// ray_plane_intersect traced.rs:91-99
__device__ __forceinline__ void ray_plane(const Ray& r, V3 n, V3 c, float& t, float& div) {
    div = om::dot(n, r.d);
    if (fabsf(div) < 0.000001f) { t = kInf; div = 0.f; return; }
    const float num = -om::dot(n, r.o - c);
    t = num / div;
}
__device__ __forceinline__ bool plane_t(float4 pn, float4 pc, const Ray& r, float t_min, float t_max, float& t_out) {  // traced.rs:105-110
    float root, ndd;
    ray_plane(r, v3(pn.x, pn.y, pn.z), v3(pc.x, pc.y, pc.z), root, ndd);
    if (root == kInf || root < t_min || root > t_max) return false;
    t_out = root;
    return true;
}
// Barycentric::hit_aux traced.rs:176-197;  a = (origin, u_length) b = (uxv, v_length)
// c = (base_inv row 0, v_in_base.x) d = (base_inv row 2, v_in_base.z)
__device__ __forceinline__ bool bary_t(float4 a, float4 b, float4 c, float4 d, bool triangle, const Ray& r, float t_min, float t_max, float& t_out) {
    float root, ndd;
    const V3 origin = v3(a.x, a.y, a.z);
    ray_plane(r, v3(b.x, b.y, b.z), origin, root, ndd);
    if (root == kInf || root < t_min || root > t_max) return false;
    const V3 pfo = ray_at(r, root) - origin;
    const float rx = om::dot(v3(c.x, c.y, c.z), pfo), ry = om::dot(v3(d.x, d.y, d.z), pfo);
    const float ux = 1.f, uy = 0.f, vx = c.w, vy = d.w;
    const float det = ux * vy - vx * uy;
    const float lambda1 = ((rx * vy - vx * ry) / det) / a.w;
    const float lambda2 = (-(rx * uy - ux * ry) / det) / b.w;
    const float lambda3 = 1.f - lambda1 - lambda2;
    bool ok = lambda1 > 0.f && lambda2 > 0.f && lambda1 < 1.f && lambda2 < 1.f;
    if (triangle) ok = ok && lambda3 > 0.f && lambda3 < 1.f;
    if (!ok) return false;
    t_out = root;
    return true;
}
// Cube::hit traced.rs:265-289; also returns the face axis
__device__ __forceinline__ bool cube_t(float4 r0, float4 r1, float4 r2, const Ray& r, float t_min, float t_max, float& t_out, int& axis_out) {
    Ray lr;
    lr.o = xf_point(r0, r1, r2, r.o);
    lr.d = xf_vec(r0, r1, r2, r.d);
    float smallest_t = kInf;
    int idx = -1;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const float di = om::comp(lr.d, i), oi = om::comp(lr.o, i);
        if (fabsf(di) < 0.00001f) continue;
        const float t1 = (0.5f - oi) / di;
        const float t2 = (-0.5f - oi) / di;
        const float t = (t1 >= 0.f && t2 >= 0.f) ? fminf(t1, t2) : fmaxf(t1, t2);
        if (t > smallest_t || t > t_max || t < t_min) continue;
        const float f = om::max_val(om::vabs(ray_at(lr, t)));
        if (!(fabsf(f - 0.5f) <= 0.00001f)) continue;
        smallest_t = t;
        idx = i;
    }
    if (idx < 0) return false;
    t_out = smallest_t;
    axis_out = idx;
    return true;
}

// om_dyn_traced (include/ottomarcher.h): a user's `impl Traced` described as a quadric in its own frame, optionally cut by a box of
// that frame.  Sphere::hit's arithmetic (traced.rs:38-56) generalised term by term, so that the unit sphere gives Sphere::hit's bits.
// qa = {A,B,C,D} qb = {E,F,G,H} qc = {I,J,clip_lo.x,clip_lo.y} qd = {clip_lo.z,clip_hi.xyz}
__device__ __forceinline__ bool quadric_t(float4 r0, float4 r1, float4 r2, float4 qa, float4 qb, float4 qc, float4 qd, const Ray& r, float t_min, float t_max, float& t_out) {
    const V3 o = xf_point(r0, r1, r2, r.o), d = xf_vec(r0, r1, r2, r.d);
    const float A = qa.x, B = qa.y, C = qa.z, D = qa.w, E = qb.x, F = qb.y, G = qb.z, H = qb.w, I = qc.x, J = qc.y;
    const float a = A * d.x * d.x + B * d.y * d.y + C * d.z * d.z + D * d.x * d.y + E * d.x * d.z + F * d.y * d.z;
    const float half_b = A * o.x * d.x + B * o.y * d.y + C * o.z * d.z + 0.5f * (D * (o.x * d.y + o.y * d.x) + E * (o.x * d.z + o.z * d.x) + F * (o.y * d.z + o.z * d.y)) +
                         0.5f * (G * d.x + H * d.y + I * d.z);
    const float c = A * o.x * o.x + B * o.y * o.y + C * o.z * o.z + D * o.x * o.y + E * o.x * o.z + F * o.y * o.z + G * o.x + H * o.y + I * o.z + J;
    Ray lr;
    lr.o = o; lr.d = d;
    auto ok = [&](float root) {
        const V3 p = ray_at(lr, root);
        return !(root < t_min || root > t_max) && !(p.x < qc.z || p.x > qd.y) && !(p.y < qc.w || p.y > qd.z) && !(p.z < qd.x || p.z > qd.w);
    };
    float root;
    if (a == 0.f) {  // the quadric is linear along this ray: one root
        if (half_b == 0.f) return false;
        root = -c / (2.f * half_b);
        if (!ok(root)) return false;
    } else {
        const float discriminant = half_b * half_b - a * c;
        if (discriminant < 0.f) return false;
        const float sqrtd = sqrtf(discriminant);
        const float s = a < 0.f ? -1.f : 1.f;  // nearer root first, also where the form is negative along the ray
        root = (-half_b - s * sqrtd) / a;
        if (!ok(root)) {
            root = (-half_b + s * sqrtd) / a;
            if (!ok(root)) return false;
        }
    }
    t_out = root;
    return true;
}

// ---- closest hit over the traced primitives ------------------------------------------------
struct Best {
    float t;
    uint32_t ref;
};
// Linear scan keeps the LATER object on equal t (see om_dev_types.h); any visiting order gives
// the same winner with this rule.
__device__ __forceinline__ void accept(Best& b, float t, uint32_t ref) {
    if (t < b.t || b.ref == omd::REF_NONE || ref > b.ref) { b.t = t; b.ref = ref; }
}
#ifndef OM_RARE_PRIMS_OUT_OF_LINE
#define OM_RARE_PRIMS_OUT_OF_LINE 0  // 1: every primitive test but the sphere's is one out-of-line function (smaller hot loop; experiment)
#endif
// the primitive kinds that the benchmark scenes hold a handful of (cube, triangle, plane, parallelogram, dyn quadric)
template <bool SMEM>
#if OM_RARE_PRIMS_OUT_OF_LINE
__device__ __noinline__
#else
__device__ __forceinline__
#endif
Best test_prim_other(SceneView<SMEM> sv, const SceneHeader& S, uint32_t ref, Ray r, float t_min, Best best) {
    const uint32_t type = omd::ref_type(ref), i = omd::ref_index(ref);
    float t;
    if (type == omd::PT_CUBE) {
        const uint32_t u = S.off_hot[omd::PT_CUBE] + omd::U_XF * i;
        int axis;
        if (cube_t(sv.hot(u), sv.hot(u + 1), sv.hot(u + 2), r, t_min, best.t, t, axis)) accept(best, t, ref);
    } else if (type == omd::PT_PLANE) {
        const uint32_t u = S.off_hot[omd::PT_PLANE] + omd::U_PLANE * i;
        if (plane_t(sv.hot(u), sv.hot(u + 1), r, t_min, best.t, t)) accept(best, t, ref);
    } else if (type == omd::PT_TDYN) {
        const uint32_t u = S.off_hot[omd::PT_TDYN] + omd::U_TDYN * i;
        if (quadric_t(sv.hot(u), sv.hot(u + 1), sv.hot(u + 2), sv.hot(u + 3), sv.hot(u + 4), sv.hot(u + 5), sv.hot(u + 6), r, t_min, best.t, t)) accept(best, t, ref);
    } else {
        const uint32_t u = S.off_hot[type] + omd::U_BARY * i;
        if (bary_t(sv.hot(u), sv.hot(u + 1), sv.hot(u + 2), sv.hot(u + 3), type == omd::PT_TRI, r, t_min, best.t, t)) accept(best, t, ref);
    }
    return best;
}
template <bool SMEM>
__device__ __forceinline__ void test_prim(const SceneView<SMEM>& sv, const SceneHeader& S, uint32_t ref, const Ray& r, float t_min, Best& best) {
    if (omd::ref_type(ref) == omd::PT_SPHERE) {
        const uint32_t u = S.off_hot[omd::PT_SPHERE] + omd::U_XF * omd::ref_index(ref);
        float t;
        if (sphere_t(sv.hot(u), sv.hot(u + 1), sv.hot(u + 2), r, t_min, best.t, t)) accept(best, t, ref);
    } else best = test_prim_other(sv, S, ref, r, t_min, best);
}

struct Counters {
    uint32_t segments, node_visits, prim_tests, sdf_evals, march_steps, normal_evals, march_entered;
};

// ---- culling arithmetic ----------------------------------------------------------------------
// Box tests only decide what gets tested, never a result, and the boxes are padded (om_scene.cpp),
// so they use explicit FMAs in BOTH builds (fmaf() is not subject to -fmad=false).
struct RayInv {
    V3 inv;    // 1/d, clamped to +-1e30 so an axis-parallel ray gives huge finite values, not inf - inf
    V3 noinv;  // -o * inv
};
__device__ __forceinline__ float clamp_inv(float d) {
    const float r = 1.f / d;
    return copysignf(fminf(fabsf(r), 1e30f), d);
}
__device__ __forceinline__ RayInv ray_inv(const Ray& r) {
    RayInv q;
    q.inv = v3(clamp_inv(r.d.x), clamp_inv(r.d.y), clamp_inv(r.d.z));
    q.noinv = v3(-r.o.x * q.inv.x, -r.o.y * q.inv.y, -r.o.z * q.inv.z);
    return q;
}
// slab test of one box against the ray; returns entry distance, or +INF when missed
__device__ __forceinline__ float slab(float lx, float ly, float lz, float hx, float hy, float hz, const RayInv& q, float t_min, float t_max) {
    const float ax = fmaf(lx, q.inv.x, q.noinv.x), bx = fmaf(hx, q.inv.x, q.noinv.x);
    const float ay = fmaf(ly, q.inv.y, q.noinv.y), by = fmaf(hy, q.inv.y, q.noinv.y);
    const float az = fmaf(lz, q.inv.z, q.noinv.z), bz = fmaf(hz, q.inv.z, q.noinv.z);
    const float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), t_min));
    const float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), t_max));
    return tn <= tf ? tn : kInf;
}

// Same test against the box widened on every side by pad * D, D = distance from the ray origin to the farthest
// corner (SceneHeader::dyn_pad): the margin for the reference's round-off, which grows with that distance.
__device__ __forceinline__ float slab_padded(float lx, float ly, float lz, float hx, float hy, float hz, const RayInv& q, V3 o, float pad, float t_min, float t_max) {
    const float dx = fmaxf(fabsf(lx - o.x), fabsf(hx - o.x)), dy = fmaxf(fabsf(ly - o.y), fabsf(hy - o.y)), dz = fmaxf(fabsf(lz - o.z), fabsf(hz - o.z));
    const float e = pad * sqrtf(fmaf(dx, dx, fmaf(dy, dy, dz * dz)));
    const float ex = e * fabsf(q.inv.x), ey = e * fabsf(q.inv.y), ez = e * fabsf(q.inv.z);
    const float ax = fmaf(lx, q.inv.x, q.noinv.x), bx = fmaf(hx, q.inv.x, q.noinv.x);
    const float ay = fmaf(ly, q.inv.y, q.noinv.y), by = fmaf(hy, q.inv.y, q.noinv.y);
    const float az = fmaf(lz, q.inv.z, q.noinv.z), bz = fmaf(hz, q.inv.z, q.noinv.z);
    const float tn = fmaxf(fmaxf(fminf(ax, bx) - ex, fminf(ay, by) - ey), fmaxf(fminf(az, bz) - ez, t_min));
    const float tf = fminf(fminf(fmaxf(ax, bx) + ex, fmaxf(ay, by) + ey), fminf(fmaxf(az, bz) + ez, t_max));
    return tn <= tf ? tn : kInf;
}
// one BVH child box; the per-ray padding exists only in the global-memory variant (the host never stages such a scene)
template <bool SMEM>
__device__ __forceinline__ float node_slab(float lx, float ly, float lz, float hx, float hy, float hz, const RayInv& q, V3 o, float pad, float t_min, float t_max) {
    if (!SMEM && pad != 0.f) return slab_padded(lx, ly, lz, hx, hy, hz, q, o, pad, t_min, t_max);
    return slab(lx, ly, lz, hx, hy, hz, q, t_min, t_max);
}

constexpr unsigned FULL = 0xffffffffu;
constexpr int STACK_DEPTH = 64;  // local-memory stack: om_scene.cpp collapses the tree so that SceneHeader::stack_need <= 60

// ---- traversal stack ------------------------------------------------------------------------------------------------
// An entry stands for a child of a wide node that the ray's box test hit but that was not the nearest: the child and the
// entry distance, so that a pop can skip it once a closer hit is known.
//  SHARED = true: 4-byte entries in shared memory behind the scene image, entry e of thread t at word e * STRIDE + t
//    (conflict-free whatever the threads' depths are): {entry distance truncated to its upper 16 bits - never above the
//    true value, so it only ever culls less -, node index (14 bits), child slot (2 bits)}; the child reference is read
//    back from the node on the pop.  The host sizes the area from SceneHeader::stack_need and only picks this variant when
//    it fits beside the scene and the tree has at most 16384 nodes.
//  SHARED = false: {child, distance} pairs in local memory (scenes that live in global memory; the wavefront kernels).
template <bool SHARED, int STRIDE>
struct TravStack;
template <int STRIDE>
struct TravStack<true, STRIDE> {
    uint32_t* col;
    __device__ __forceinline__ void init(uint32_t hot_units) { col = reinterpret_cast<uint32_t*>(om_smem + hot_units) + threadIdx.x; }
    __device__ __forceinline__ void push(int sp, uint32_t key, uint32_t node, int) { col[sp * STRIDE] = (key & 0xFFFF0003u) | (node << 2); }
    template <bool SMEM>
    __device__ __forceinline__ bool pop(int sp, float best_t, const SceneView<SMEM>& sv, const SceneHeader& S, int& cur) const {
        const uint32_t e = col[sp * STRIDE];
        if (!(__uint_as_float(e & 0xFFFF0000u) <= best_t)) return false;
        cur = (int)sv.hot_word(S.off_nodes + omd::U_NODE4 * ((e >> 2) & 0x3FFFu) + 6u, e & 3u);
        return true;
    }
};
template <int STRIDE>
struct TravStack<false, STRIDE> {
    uint2 e[STACK_DEPTH];
    __device__ __forceinline__ void init(uint32_t) {}
    __device__ __forceinline__ void push(int sp, uint32_t key, uint32_t, int child) { e[sp] = make_uint2((uint32_t)child, key); }
    template <bool SMEM>
    __device__ __forceinline__ bool pop(int sp, float best_t, const SceneView<SMEM>&, const SceneHeader&, int& cur) const {
        const uint2 x = e[sp];
        if (!(__uint_as_float(x.y & 0xFFFFFFFCu) <= best_t)) return false;
        cur = (int)x.x;
        return true;
    }
};

// k-th entry of the linear part of a query: the always-test list (infinite planes, scene-sized
// objects), or in brute-force mode every traced primitive in the reference's scan order.
template <bool SMEM>
__device__ __forceinline__ uint32_t linear_ref(const SceneView<SMEM>& sv, const SceneHeader& S, bool brute, uint32_t k) {
    if (brute) {
        uint32_t type = 0;
        while (type < 4u && k >= S.n[type]) { k -= S.n[type]; type++; }
        if (type == 4u && k >= S.n[4]) { k -= S.n[4]; type = omd::PT_TDYN; }  // traced_objects come after the typed lists (hits.rs:280-285)
        return omd::make_ref(type, k);
    }
    const float4 q = sv.hot(S.off_large + (k >> 2));
    const uint32_t sel = k & 3u;
    return __float_as_uint(sel == 0 ? q.x : (sel == 1 ? q.y : (sel == 2 ? q.z : q.w)));
}

// One visit of a wide node (omb::Node4): the ray against the four child boxes.  The planes of each axis are fetched
// near-first by address from the sign of the ray direction (fma is monotonic in the plane, so this is the min / max of
// the two-plane form without computing it), every hit child gets the key {entry distance, slot} - one unsigned
// compare orders them -, a five-exchange network sorts the four keys, the hit children but the nearest go on the stack
// farthest first, and the nearest becomes `cur`.  Returns false when the ray misses all four.
// The per-ray padding (SceneHeader::dyn_pad, see slab_padded) exists only in the global-memory variant.
template <bool SMEM, class STK>
__device__ __forceinline__ bool visit4(const SceneView<SMEM>& sv, const SceneHeader& S, uint32_t node, const RayInv& q, V3 o, float pad, float t_min, float best_t,
                                       STK& stk, int& sp, int& cur) {
    const uint32_t u = S.off_nodes + omd::U_NODE4 * node;
    const uint32_t sx = __float_as_uint(q.inv.x) >> 31, sy = __float_as_uint(q.inv.y) >> 31, sz = __float_as_uint(q.inv.z) >> 31;
    const float4 nx = sv.hot(u + sx), fx = sv.hot(u + (sx ^ 1u));
    const float4 ny = sv.hot(u + 2u + sy), fy = sv.hot(u + 2u + (sy ^ 1u));
    const float4 nz = sv.hot(u + 4u + sz), fz = sv.hot(u + 4u + (sz ^ 1u));
    const float4 cf = sv.hot(u + 6u);
    // per-ray padding (global-memory variant): ONE margin per node, from a ball around its children's boxes that the host
    // packed (SceneHeader::off_node_far): |o - centre| + radius is at least the distance to the farthest point of any child's
    // box, which is what each child's own margin would use (the padding only ever has to be large enough) - one square
    // root and a dozen instructions instead of four roots and fifty
    float ex = 0.f, ey = 0.f, ez = 0.f;
    if (!SMEM && pad != 0.f) {
        const float4 cr = sv.hot(S.off_node_far + node);
        const float dx = cr.x - o.x, dy = cr.y - o.y, dz = cr.z - o.z;
        const float e = pad * (sqrtf(fmaf(dx, dx, fmaf(dy, dy, dz * dz))) + cr.w);
        ex = e * fabsf(q.inv.x); ey = e * fabsf(q.inv.y); ez = e * fabsf(q.inv.z);
    }
    // the margin rides in the constant term of the planes' fma: near planes move by -e, far planes by +e
    const float lox = q.noinv.x - ex, hix = q.noinv.x + ex, loy = q.noinv.y - ey, hiy = q.noinv.y + ey, loz = q.noinv.z - ez, hiz = q.noinv.z + ez;
    auto key = [&](float nxj, float fxj, float nyj, float fyj, float nzj, float fzj, uint32_t j) -> uint32_t {
        const float ax = fmaf(nxj, q.inv.x, lox), bx = fmaf(fxj, q.inv.x, hix);
        const float ay = fmaf(nyj, q.inv.y, loy), by = fmaf(fyj, q.inv.y, hiy);
        const float az = fmaf(nzj, q.inv.z, loz), bz = fmaf(fzj, q.inv.z, hiz);
        // (an empty slot's inverted box (+INF, -INF) gives ax = +INF or bx = -INF whatever the margin: tn <= tf fails)
        const float tn = fmaxf(fmaxf(ax, ay), fmaxf(az, t_min));
        const float tf = fminf(fminf(bx, by), fminf(bz, best_t));
        return (tn <= tf) ? ((__float_as_uint(fmaxf(tn, 0.f)) & 0xFFFFFFFCu) | j) : 0xFFFFFFFFu;
    };
    uint32_t k0 = key(nx.x, fx.x, ny.x, fy.x, nz.x, fz.x, 0u), k1 = key(nx.y, fx.y, ny.y, fy.y, nz.y, fz.y, 1u);
    uint32_t k2 = key(nx.z, fx.z, ny.z, fy.z, nz.z, fz.z, 2u), k3 = key(nx.w, fx.w, ny.w, fy.w, nz.w, fz.w, 3u);
    auto cx = [](uint32_t& a, uint32_t& b) { const uint32_t lo = min(a, b), hi = max(a, b); a = lo; b = hi; };
    cx(k0, k1); cx(k2, k3); cx(k0, k2); cx(k1, k3); cx(k1, k2);
    if (k0 == 0xFFFFFFFFu) return false;
    auto child = [&](uint32_t k) { return __float_as_int(sel4f(k & 3u, cf)); };
    if (k3 != 0xFFFFFFFFu) { stk.push(sp, k3, node, child(k3)); sp++; }
    if (k2 != 0xFFFFFFFFu) { stk.push(sp, k2, node, child(k2)); sp++; }
    if (k1 != 0xFFFFFFFFu) { stk.push(sp, k1, node, child(k1)); sp++; }
    cur = child(k0);
    return true;
}

// FrozenHittableList::hit, traced part (hits.rs:270-286), for a whole warp at once.
//
// Every lane walks its own ray through a list of ITEMS: first the linear entries, then the BVH
// (inner node = four child boxes, leaf = primitive).  A lane's next item is either an inner node or a
// leaf; each iteration the warp VOTES and runs only the kind that more lanes are waiting for, so
// the two bodies always execute with at least half of the busy lanes active (the first version
// tested a leaf the moment a lane reached it: ~2 lanes active, half of all issue slots).
// All 32 lanes must call this; `alive` says whether the lane has a ray.
template <bool SMEM, bool COUNT, class STK>
__device__ __forceinline__ Best closest_traced_warp(const SceneView<SMEM>& sv, const SceneHeader& S, const Ray& r, float t_min, float t_max, bool brute,
                                                    bool alive, Counters& cn, STK& stk) {
    Best best;
    best.t = t_max;
    best.ref = omd::REF_NONE;
    const uint32_t n_lin = brute ? (S.n[0] + S.n[1] + S.n[2] + S.n[3] + S.n[4] + S.n[omd::PT_TDYN]) : S.n_large;
    const bool have_bvh = !brute && S.n_nodes > 0u;
    const RayInv q = ray_inv(r);
    const float pad = S.dyn_pad;
    uint32_t k = 0;
    int sp = 0, cur = 0;
    bool act = alive, in_bvh = false;
    // take the next item after a leaf (or at the start): linear entries, then the BVH root, then the stack
    auto pop = [&]() {
        while (sp > 0) {
            sp--;
            if (stk.pop(sp, best.t, sv, S, cur)) return;
        }
        act = false;
    };
    auto next_item = [&]() {
        if (k < n_lin) { cur = ~(int)linear_ref(sv, S, brute, k); k++; return; }
        if (!in_bvh) { in_bvh = true; if (have_bvh) { cur = 0; return; } }
        pop();
    };
    if (act) next_item();
    for (;;) {
        const unsigned m_inner = __ballot_sync(FULL, act && cur >= 0);
        const unsigned m_leaf = __ballot_sync(FULL, act && cur < 0);
        if ((m_inner | m_leaf) == 0u) break;
        if (__popc(m_inner) >= __popc(m_leaf)) {
            if (act && cur >= 0) {
                if (COUNT) cn.node_visits++;
                if (!visit4(sv, S, (uint32_t)cur, q, r.o, pad, t_min, best.t, stk, sp, cur)) pop();
            }
        } else {
            if (act && cur < 0) {
                if (COUNT) cn.prim_tests++;
                test_prim(sv, S, (uint32_t)(~cur), r, t_min, best);
                next_item();
            }
        }
    }
    return best;
}

// ---- marched objects (marched.rs) --------------------------------------------------------------
// |Marched::sdf| of object `ref` at world point p (marched.rs:14-18)
__device__ __forceinline__ float torus_local_sdf(V3 p, float R, float rr) {  // marched.rs:134-138
    const float qx = sqrtf(p.x * p.x + p.z * p.z) - R;
    return sqrtf(qx * qx + p.y * p.y) - rr;
}
__device__ __forceinline__ float box_local_sdf(V3 p, V3 sizes) {  // marched.rs:86-89
    const V3 q = om::vabs(p) - sizes;
    const V3 qm = v3(fmaxf(q.x, 0.f), fmaxf(q.y, 0.f), fmaxf(q.z, 0.f));
    return om::length(qm) + fminf(fmaxf(q.x, fmaxf(q.y, q.z)), 0.f);
}
template <bool SMEM>
__device__ __forceinline__ V3 torus_to_local(const SceneView<SMEM>& sv, uint32_t u, V3 p) {  // marched.rs:142-144
    const float4 r0 = sv.hot(u), r1 = sv.hot(u + 1), r2 = sv.hot(u + 2), si = sv.hot(u + 6);
    const float qx = p.x * si.x, qy = p.y * si.y, qz = p.z * si.z, qw = 1.f * si.w;
    return v3(r0.x * qx + r0.y * qy + r0.z * qz + r0.w * qw, r1.x * qx + r1.y * qy + r1.z * qz + r1.w * qw, r2.x * qx + r2.y * qy + r2.z * qz + r2.w * qw);
}
// local_sdf of a dyn marched object (om_dyn_marched, include/ottomarcher.h): the postfix program at words
// [first, first + len) of the code pool, evaluated at local point p.  One f32 operation per instruction, in program
// order; the host has checked the program (omh::check_sdf_program), so there are no bounds tests here.
template <bool SMEM>
__device__ __noinline__ float sdf_program(SceneView<SMEM> sv, uint32_t off_code, uint32_t first, uint32_t len, V3 p) {
    float st[OM_SDF_STACK], reg[OM_SDF_REGS];
    int sp = 0;
    for (uint32_t pc = first; pc < first + len; pc++) {
        const uint32_t w = sv.hot_word(off_code, pc), op = w & 0xFFu, arg = w >> 8;
        switch (op) {
            case OM_SDF_X: st[sp++] = p.x; break;
            case OM_SDF_Y: st[sp++] = p.y; break;
            case OM_SDF_Z: st[sp++] = p.z; break;
            case OM_SDF_CONST: pc++; st[sp++] = __uint_as_float(sv.hot_word(off_code, pc)); break;
            case OM_SDF_ADD: sp--; st[sp - 1] = st[sp - 1] + st[sp]; break;
            case OM_SDF_SUB: sp--; st[sp - 1] = st[sp - 1] - st[sp]; break;
            case OM_SDF_MUL: sp--; st[sp - 1] = st[sp - 1] * st[sp]; break;
            case OM_SDF_DIV: sp--; st[sp - 1] = st[sp - 1] / st[sp]; break;
            case OM_SDF_MIN: sp--; st[sp - 1] = fminf(st[sp - 1], st[sp]); break;
            case OM_SDF_MAX: sp--; st[sp - 1] = fmaxf(st[sp - 1], st[sp]); break;
            case OM_SDF_NEG: st[sp - 1] = -st[sp - 1]; break;
            case OM_SDF_ABS: st[sp - 1] = fabsf(st[sp - 1]); break;
            case OM_SDF_SQRT: st[sp - 1] = sqrtf(st[sp - 1]); break;
            case OM_SDF_LEN2: sp--; st[sp - 1] = om::length(v3(st[sp - 1], st[sp], 0.f)); break;
            case OM_SDF_LEN3: sp -= 2; st[sp - 1] = om::length(v3(st[sp - 1], st[sp], st[sp + 1])); break;
            case OM_SDF_STORE: sp--; reg[arg & (OM_SDF_REGS - 1)] = st[sp]; break;
            default: st[sp++] = reg[arg & (OM_SDF_REGS - 1)]; break;  // OM_SDF_LOAD
        }
    }
    return st[0];
}
struct DynObj {  // hot record of a dyn marched object
    float4 w2l0, w2l1, w2l2;
    float world_f;
    uint32_t first, len;
};
template <bool SMEM>
__device__ __forceinline__ DynObj load_dyn(const SceneView<SMEM>& sv, const SceneHeader& S, uint32_t i) {
    const uint32_t u = S.off_hot[omd::PT_MDYN] + omd::U_MDYN * i;
    DynObj d;
    d.w2l0 = sv.hot(u); d.w2l1 = sv.hot(u + 1); d.w2l2 = sv.hot(u + 2);
    const float4 m = sv.hot(u + 6);
    d.world_f = m.x; d.first = __float_as_uint(m.y); d.len = __float_as_uint(m.z);
    return d;
}
// Mat4x4::dot of a point (w = 1): x*x + y*y + z*z + w*1 (mat4x4.rs:45-48, vec4.rs dot)
__device__ __forceinline__ V3 xf_point_w1(float4 r0, float4 r1, float4 r2, V3 p) {
    return v3(r0.x * p.x + r0.y * p.y + r0.z * p.z + r0.w * 1.f, r1.x * p.x + r1.y * p.y + r1.z * p.z + r1.w * 1.f, r2.x * p.x + r2.y * p.y + r2.z * p.z + r2.w * 1.f);
}
// DYN: the scene may hold dyn marched objects.  The kernel that renders scenes without marched structures and without dyn
// objects (k_render<.., MA = 0 or 1>) is instantiated with DYN = false and contains none of this code (see march()).
template <bool SMEM, bool DYN>
__device__ __forceinline__ float marched_sdf(const SceneView<SMEM>& sv, const SceneHeader& S, uint32_t type, uint32_t i, V3 p) {
    if (type == omd::PT_MSPHERE) {
        const float4 cr = sv.hot(S.off_hot[omd::PT_MSPHERE] + omd::U_MSPHERE * i);
        return om::length(p - v3(cr.x, cr.y, cr.z)) - cr.w;
    } else if (type == omd::PT_MBOX) {
        const uint32_t u = S.off_hot[omd::PT_MBOX] + omd::U_MBOX * i;
        const float4 c = sv.hot(u), s = sv.hot(u + 1);
        return box_local_sdf(p - v3(c.x, c.y, c.z), v3(s.x, s.y, s.z));
    } else if (!DYN || type == omd::PT_MTORUS) {
        const uint32_t u = S.off_hot[omd::PT_MTORUS] + omd::U_MTORUS * i;
        const float4 sz = sv.hot(u + 8);
        return torus_local_sdf(torus_to_local(sv, u, p), sz.x, sz.y) * sz.z;  // to_world_f marched.rs:148-150
    } else {  // trait default, marched.rs:14-18
        const DynObj d = load_dyn(sv, S, i);
        return sdf_program(sv, S.off_dyn_code, d.first, d.len, xf_point_w1(d.w2l0, d.w2l1, d.w2l2, p)) * d.world_f;
    }
}

// Marched::get_outward_local_normal (marched.rs:25-44); LSDF is the object's local sdf
template <class LSDF>
__device__ __forceinline__ V3 numeric_local_normal(V3 p, LSDF lsdf) {
    const float eps = 0.0000001f;
    const float x = lsdf(v3(p.x + eps, p.y, p.z)) - lsdf(v3(p.x - eps, p.y, p.z));
    const float y = lsdf(v3(p.x, p.y + eps, p.z)) - lsdf(v3(p.x, p.y - eps, p.z));
    const float z = lsdf(v3(p.x, p.y, p.z + eps)) - lsdf(v3(p.x, p.y, p.z - eps));
    const V3 normal = om::unit(v3(x, y, z));
    const V3 dir = om::unit(normal);                         // Ray::new normalises again (ray.rs:11-13)
    const V3 start = v3(0.f, 0.f, 0.f) + 0.f * dir;          // test_ray.at(0.)
    const float start_val = lsdf(start);
    const V3 end = v3(0.f, 0.f, 0.f) + 1.f * dir;            // test_ray.at(1.)
    const float end_val = lsdf(end);
    const float sign = (end_val > start_val) ? 1.f : -1.f;
    return normal * sign;
}
// Marched::get_outward_normal (marched.rs:19-24; MarchedSphere override :60-63).  Out of line: it
// runs once per marched HIT, not per step (see march()).
template <bool SMEM, bool DYN>
__device__ __noinline__ V3 marched_normal(SceneView<SMEM> sv, const SceneHeader& S, uint32_t ref, V3 p) {
    const uint32_t type = omd::ref_type(ref), i = omd::ref_index(ref);
    if (type == omd::PT_MSPHERE) {
        const float4 cr = sv.hot(S.off_hot[omd::PT_MSPHERE] + omd::U_MSPHERE * i);
        return om::unit(p - v3(cr.x, cr.y, cr.z));
    } else if (type == omd::PT_MBOX) {
        const uint32_t u = S.off_hot[omd::PT_MBOX] + omd::U_MBOX * i;
        const float4 c = sv.hot(u), s = sv.hot(u + 1);
        const V3 sizes = v3(s.x, s.y, s.z);
        const V3 n = numeric_local_normal(p - v3(c.x, c.y, c.z), [&](V3 q) { return box_local_sdf(q, sizes); });
        // to_world of a vector: n + n.w * center with n.w = 0 (marched.rs:96-98)
        return om::unit(v3(n.x + 0.f * c.x, n.y + 0.f * c.y, n.z + 0.f * c.z));
    } else if (!DYN || type == omd::PT_MTORUS) {
        const uint32_t u = S.off_hot[omd::PT_MTORUS] + omd::U_MTORUS * i;
        const float4 sz = sv.hot(u + 8);
        const V3 n = numeric_local_normal(torus_to_local(sv, u, p), [&](V3 q) { return torus_local_sdf(q, sz.x, sz.y); });
        const float4 r0 = sv.hot(u + 3), r1 = sv.hot(u + 4), r2 = sv.hot(u + 5), sc = sv.hot(u + 7);
        const V3 w = xf_vec(r0, r1, r2, n);  // L2W_TR . (n, 0)
        return om::unit(v3(w.x * sc.x, w.y * sc.y, w.z * sc.z));
    } else {  // trait defaults, marched.rs:19-44
        const DynObj d = load_dyn(sv, S, i);
        const V3 n = numeric_local_normal(xf_point_w1(d.w2l0, d.w2l1, d.w2l2, p), [&](V3 q) { return sdf_program(sv, S.off_dyn_code, d.first, d.len, q); });
        const uint32_t u = S.off_hot[omd::PT_MDYN] + omd::U_MDYN * i;
        const float4 r0 = sv.hot(u + 3), r1 = sv.hot(u + 4), r2 = sv.hot(u + 5);
        // to_world(n, 0): x*x + y*y + z*z + w*0
        return om::unit(v3(r0.x * n.x + r0.y * n.y + r0.z * n.z + r0.w * 0.f, r1.x * n.x + r1.y * n.y + r1.z * n.z + r1.w * 0.f, r2.x * n.x + r2.y * n.y + r2.z * n.z + r2.w * 0.f));
    }
}

// j-th marched object in the reference's order: spheres, boxes, tori (hits.rs:301-311)
template <bool DYN = true>
__device__ __forceinline__ uint32_t marched_ref(const SceneHeader& S, uint32_t j) {
    const uint32_t n_ms = S.n[omd::PT_MSPHERE], n_mb = S.n[omd::PT_MBOX];
    if (j < n_ms) return omd::make_ref(omd::PT_MSPHERE, j);
    if (j < n_ms + n_mb) return omd::make_ref(omd::PT_MBOX, j - n_ms);
    const uint32_t n_mt = S.n[omd::PT_MTORUS];
    if (!DYN || j < n_ms + n_mb + n_mt) return omd::make_ref(omd::PT_MTORUS, j - n_ms - n_mb);
    return omd::make_ref(omd::PT_MDYN, j - n_ms - n_mb - n_mt);  // marched_objects, after the typed lists (hits.rs:312-320)
}
constexpr int MSTACK = 40;  // scratch entries nearest_marched_accel needs
struct Near {
    uint32_t evals;  // SDFs evaluated (executed-work accounting)
    float d;       // min over the marched objects of |sdf(p)|
    uint32_t ref;  // the first object that attains it (strict `<`, hits.rs:304,345); REF_NONE if none
};
template <bool SMEM, bool DYN>
__device__ __noinline__ Near nearest_marched(SceneView<SMEM> sv, const SceneHeader& S, V3 p);
// min_j |sdf_j(p)| with help of the host-built structures (scenes with many marched objects):
//  - inside the candidate grid only the cell's list is evaluated, in the reference's order: the objects left out are
//    provably farther than a listed one everywhere in the cell (om_scene.cpp, build_marched_grid);
//  - elsewhere the distance BVH is walked with the exact minimum found so far as the bound.  A subtree is skipped only when
//    a LOWER bound of every |sdf| in it is strictly above that minimum, so what is skipped can neither be the minimum nor
//    tie with it; the tree hands objects out of order, so an equal value from a smaller index takes over — the first
//    object of the reference's scan that attains the minimum (om_scene.cpp, build_marched_bvh).
// When the walk runs out of scratch stack it scans every object instead.  (Scenes with dyn marched objects have neither
// structure - a user program has no analytic bounds - so this code never meets one.)
template <bool SMEM>
__device__ __noinline__ Near nearest_marched_accel(SceneView<SMEM> sv, const SceneHeader& S, V3 p, int2* stk) {
    Near n;
    n.d = kInf;
    n.ref = omd::REF_NONE;
    n.evals = 0;
    auto word = [&](uint32_t off, uint32_t i) { return sv.cold_word(off, i); };
    if (S.mgrid_dim[0] != 0u) {
        const float fx = (p.x - S.mgrid_lo[0]) * S.mgrid_inv_h, fy = (p.y - S.mgrid_lo[1]) * S.mgrid_inv_h, fz = (p.z - S.mgrid_lo[2]) * S.mgrid_inv_h;
        if (fx >= 0.f && fy >= 0.f && fz >= 0.f && fx < (float)S.mgrid_dim[0] && fy < (float)S.mgrid_dim[1] && fz < (float)S.mgrid_dim[2]) {
            const uint32_t cell = ((uint32_t)fz * S.mgrid_dim[1] + (uint32_t)fy) * S.mgrid_dim[0] + (uint32_t)fx;
            const float4 cc = sv.cold(S.off_mgrid_cells + (cell >> 1));
            const uint32_t first = __float_as_uint((cell & 1u) ? cc.z : cc.x), count = __float_as_uint((cell & 1u) ? cc.w : cc.y);
            n.evals = count;
            // A list holds indices of the reference's scan in increasing order, i.e. spheres, then boxes, then tori.  One loop per
            // type: the lanes of a warp have lists of different make-up, and a single loop with the type switch inside ran
            // all three bodies in most of its trips (4.2 lanes per instruction in the SDFs, ncu of C3).
            const uint32_t end = first + count, n_ms = S.n[omd::PT_MSPHERE], n_sb = n_ms + S.n[omd::PT_MBOX];
            uint32_t k = first, j = k < end ? word(S.off_mgrid_lists, k) : 0xFFFFFFFFu;
            auto take = [&](uint32_t type, uint32_t i) {
                const float d = fabsf(marched_sdf<SMEM, false>(sv, S, type, i, p));
                if (d < n.d) { n.d = d; n.ref = omd::make_ref(type, i); }
                k++;
                j = k < end ? word(S.off_mgrid_lists, k) : 0xFFFFFFFFu;
            };
            while (j < n_ms) take(omd::PT_MSPHERE, j);
            while (j < n_sb) take(omd::PT_MBOX, j - n_ms);
            while (j != 0xFFFFFFFFu) take(omd::PT_MTORUS, j - n_sb);
            return n;
        }
    }
    bool walk = S.n_mbvh_nodes != 0u;
    // A subtree is skipped when mbvh_f * dist(p, box) - slack > n.d.  The walk compares SQUARED distances with the squared
    // threshold instead (two IEEE square roots per node were a third of its instructions); the threshold is widened by 1e-5
    // relative, far more than the round-off of either side, so the squared test never skips what the plain one keeps.
    const float slack = S.mbvh_margin + 1e-5f * (fabsf(p.x) + fabsf(p.y) + fabsf(p.z));
    const float inv_f = 1.f / S.mbvh_f;
    auto lower2 = [&](float lx, float ly, float lz, float hx, float hy, float hz) {
        const float dx = fmaxf(fmaxf(lx - p.x, p.x - hx), 0.f), dy = fmaxf(fmaxf(ly - p.y, p.y - hy), 0.f), dz = fmaxf(fmaxf(lz - p.z, p.z - hz), 0.f);
        return fmaf(dx, dx, fmaf(dy, dy, dz * dz));
    };
    auto reach2 = [&](float d) { const float r = (d + slack) * inv_f; return r * r * 1.00001f; };  // (+INF stays +INF)
    float lim2 = reach2(n.d);
    uint32_t k = 0, best_j = 0xFFFFFFFFu;
    int sp = 1;
    stk[0] = make_int2(0, __float_as_int(0.f));  // the root, after the scene-sized objects of the always list
    while (walk) {
        uint32_t j = 0xFFFFFFFFu;
        if (k < S.n_mbvh_always) { j = word(S.off_mbvh_always, k); k++; }
        else {
            while (sp > 0) {
                sp--;
                const int2 e = stk[sp];
                if (__int_as_float(e.y) > lim2) continue;
                if (e.x < 0) { j = (uint32_t)(~e.x); break; }
                const uint32_t u = S.off_mbvh_nodes + omd::U_NODE * (uint32_t)e.x;
                const float4 n0 = sv.cold(u), n1 = sv.cold(u + 1), n2 = sv.cold(u + 2), n3 = sv.cold(u + 3);
                float dl = lower2(n0.x, n0.y, n0.z, n0.w, n1.x, n1.y), dr = lower2(n1.z, n1.w, n2.x, n2.y, n2.z, n2.w);
                int cl = __float_as_int(n3.x), cr = __float_as_int(n3.y);
                if (dr < dl) { const float t = dl; dl = dr; dr = t; const int c = cl; cl = cr; cr = c; }  // (cl, dl): the nearer child, popped first
                if (sp + 2 > MSTACK) { walk = false; break; }
                if (!(dr > lim2)) { stk[sp] = make_int2(cr, __float_as_int(dr)); sp++; }
                if (!(dl > lim2)) { stk[sp] = make_int2(cl, __float_as_int(dl)); sp++; }
            }
        }
        if (!walk) break;
        if (j == 0xFFFFFFFFu) return n;
        const uint32_t ref = marched_ref<false>(S, j);
        const float d = fabsf(marched_sdf<SMEM, false>(sv, S, omd::ref_type(ref), omd::ref_index(ref), p));
        n.evals++;
        if (d < n.d || (d == n.d && j < best_j && best_j != 0xFFFFFFFFu)) { n.d = d; n.ref = ref; best_j = j; lim2 = reach2(d); }
    }
    return nearest_marched<SMEM, false>(sv, S, p);
}
template <bool SMEM, bool DYN>
__device__ __noinline__ Near nearest_marched(SceneView<SMEM> sv, const SceneHeader& S, V3 p) {  // hits.rs:301-320
    Near n;
    n.d = kInf;
    n.ref = omd::REF_NONE;
    n.evals = S.n_marched;
    for (uint32_t j = 0; j < S.n_marched; j++) {
        const uint32_t ref = marched_ref<DYN>(S, j);
        const float d = fabsf(marched_sdf<SMEM, DYN>(sv, S, omd::ref_type(ref), omd::ref_index(ref), p));
        if (d < n.d) { n.d = d; n.ref = ref; }
    }
    return n;
}
// The same value through the host-built structures when the scene has them (warp-uniform choice)
template <bool SMEM>
__device__ __forceinline__ Near nearest_marched(const SceneView<SMEM>& sv, const SceneHeader& S, V3 p, int2* stk) {  // stk: MSTACK entries of scratch
    if ((S.mgrid_dim[0] | S.n_mbvh_nodes) != 0u) return nearest_marched_accel(sv, S, p, stk);
    return nearest_marched<SMEM, true>(sv, S, p);
}
template <bool SMEM, bool DYN = true>
__device__ __noinline__ float abs_sdf_of(SceneView<SMEM> sv, const SceneHeader& S, uint32_t ref, V3 p) {
    return fabsf(marched_sdf<SMEM, DYN>(sv, S, omd::ref_type(ref), omd::ref_index(ref), p));
}

// Phase alignment (OM_RENDER_PHASE_ALIGN): the 8 warps of one SM sub-partition (warp % 4; they share an L0 instruction
// cache) meet at a named barrier at the top of the path loop and again between the closest-hit query and shading, so that
// they stream the same part of this 70 KB kernel at the same time.  The count variant also tells the warps that ran out of
// tickets when the last warp of their group has left the loop.
#ifndef OM_MARCH_STEPS
#define OM_MARCH_STEPS 16  // bounded marching: sphere-tracing steps per path-loop iteration
#endif
#ifndef OM_BOUNDED_MARCH_ALL
#define OM_BOUNDED_MARCH_ALL 0  // 1: also in the kernel of scenes without marched structures (experiment)
#endif
#ifndef OM_GROUP_TICKETS
#define OM_GROUP_TICKETS 8
#endif
#ifndef OM_TILE_BLOCKS
#define OM_TILE_BLOCKS 1
#endif
#ifndef OM_PB_GROUPS
#define OM_PB_GROUPS 4
#endif
constexpr int PB_GROUPS = OM_PB_GROUPS;
constexpr int PB_THREADS = RENDER_THREADS / PB_GROUPS;
__device__ __forceinline__ unsigned phase_bar_id() { return 1u + ((threadIdx.x >> 5) % (unsigned)PB_GROUPS); }
__device__ __forceinline__ int phase_bar_count(int pred) {
    int r;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %2, 0;\n\tbar.red.popc.u32 %0, %1, %3, p;\n\t}" : "=r"(r) : "r"(phase_bar_id()), "r"(pred), "n"(PB_THREADS) : "memory");
    return r;
}
__device__ __forceinline__ void phase_bar() { asm volatile("bar.sync %0, %1;" ::"r"(phase_bar_id()), "n"(PB_THREADS) : "memory"); }

struct Hit {
    float t;
    uint32_t ref;  // REF_NONE = miss
    V3 point, normal;  // filled for marched hits only
};

// FrozenHittableList::hit, marched part (hits.rs:288-332) + unstuck (hits.rs:336-365).
// `closest` is the traced hit distance (or t_max).  Returns ref == REF_NONE when the march ends
// without a hit (the traced record stands).
// ACCEL: the scene has a candidate grid / distance BVH.  DYN: it has dyn marched objects (om_dyn_marched).  k_render is instantiated for either case (template MA) and the host
// picks: merely linking the structures' code into the kernel that renders C2 (one marched torus, never uses them) cost it
// 8-10 % (more spill traffic in the 64-register main body), so that kernel does not contain it.
template <bool SMEM, bool COUNT, bool ACCEL, bool DYN>
__device__ __noinline__ Hit march(SceneView<SMEM> sv, const SceneHeader& S, Ray r, float t_min, float t_max, float closest, Counters* cn) {
    int2 stk[ACCEL ? MSTACK : 1];  // scratch of the distance-BVH walk
    Hit out;
    out.t = closest;
    out.ref = omd::REF_NONE;
    out.point = v3(0.f, 0.f, 0.f);
    out.normal = v3(0.f, 0.f, 0.f);
    if (COUNT) cn->march_entered++;
    // ---- unstuck ----
    float t;
    {
        const float MIN_STEP_SIZE = HIT_SIZE / 2.f;
        float new_t = t_min;
        const Near n0 = ACCEL ? nearest_marched_accel(sv, S, ray_at(r, t_min), stk) : nearest_marched<SMEM, DYN>(sv, S, ray_at(r, t_min));
        if (COUNT) cn->sdf_evals += n0.evals;
        if (n0.ref == omd::REF_NONE) return out;  // `marched.is_none()` -> INF -> return rec (also every NaN ray)
        float aux = n0.d;
        int steps = 0;
        while (aux < HIT_SIZE && steps < UNSTUCK_MAX_STEPS) {
            new_t += MIN_STEP_SIZE;
            aux = abs_sdf_of<SMEM, DYN>(sv, S, n0.ref, ray_at(r, new_t));
            steps++;
            if (COUNT) cn->sdf_evals++;
        }
        t = new_t;
    }
    // ---- sphere tracing ----
    int max_march_iter = 1024;
    while (t < t_max && t < closest && max_march_iter > 0) {
        max_march_iter--;
        const V3 point = ray_at(r, t);
        const Near n = ACCEL ? nearest_marched_accel(sv, S, point, stk) : nearest_marched<SMEM, DYN>(sv, S, point);
        if (COUNT) { cn->march_steps++; cn->sdf_evals += n.evals; }
        if (n.ref == omd::REF_NONE) return out;  // `material.is_none()`
        if (n.d < HIT_SIZE) {
            // The reference evaluates the normal of the nearest object at every step
            // (hits.rs:306); only the value at the accepted point survives, so it is computed here.
            out.t = t;
            out.ref = n.ref;
            out.point = point;
            out.normal = marched_normal<SMEM, DYN>(sv, S, n.ref, point);
            if (COUNT) cn->normal_evals++;
            return out;
        }
        t += n.d;
    }
    return out;
}

// hits.rs:288-365 in instalments: the same steps in the same order as march() above, at most `steps` sphere-tracing
// iterations per call.  Returns true when the march is over: out.ref is the marched hit, or REF_NONE = the traced record stands.
template <bool SMEM, bool COUNT, bool ACCEL, bool DYN>
__device__ __noinline__ bool march_some(SceneView<SMEM> sv, const SceneHeader& S, Ray r, float t_min, float t_max, float closest, float& t, uint32_t& left, int steps,
                                        Hit& out, Counters* cn) {
    int2 stk[ACCEL ? MSTACK : 1];
    out.t = closest; out.ref = omd::REF_NONE; out.point = v3(0.f, 0.f, 0.f); out.normal = v3(0.f, 0.f, 0.f);
    if (left == 0xFFFFFFFFu) {  // unstuck
        if (COUNT) cn->march_entered++;
        const float MIN_STEP_SIZE = HIT_SIZE / 2.f;
        float new_t = t_min;
        const Near n0 = ACCEL ? nearest_marched_accel(sv, S, ray_at(r, t_min), stk) : nearest_marched<SMEM, DYN>(sv, S, ray_at(r, t_min));
        if (COUNT) cn->sdf_evals += n0.evals;
        if (n0.ref == omd::REF_NONE) return true;
        float aux = n0.d;
        int k = 0;
        while (aux < HIT_SIZE && k < UNSTUCK_MAX_STEPS) {
            new_t += MIN_STEP_SIZE;
            aux = abs_sdf_of<SMEM, DYN>(sv, S, n0.ref, ray_at(r, new_t));
            k++;
            if (COUNT) cn->sdf_evals++;
        }
        t = new_t;
        left = 1024u;
    }
    while (t < t_max && t < closest && left > 0u) {
        if (steps-- <= 0) return false;
        left--;
        const V3 point = ray_at(r, t);
        const Near n = ACCEL ? nearest_marched_accel(sv, S, point, stk) : nearest_marched<SMEM, DYN>(sv, S, point);
        if (COUNT) { cn->march_steps++; cn->sdf_evals += n.evals; }
        if (n.ref == omd::REF_NONE) return true;
        if (n.d < HIT_SIZE) {
            out.t = t; out.ref = n.ref; out.point = point;
            out.normal = marched_normal<SMEM, DYN>(sv, S, n.ref, point);
            if (COUNT) cn->normal_evals++;
            return true;
        }
        t += n.d;
    }
    return true;
}


// FrozenHittableList::hit (hits.rs:270-334); also first_hit (the BVH subsumes CameraHash).
// Warp-collective: all 32 lanes call it, `alive` marks the lanes that carry a ray.
// MA: 0 = the scene has no candidate grid / distance BVH and no dyn marched objects, 1 = it has the structures,
// 3 = it has dyn marched objects (om_dyn_marched), 2 = look at the header (parity hooks)
template <bool SMEM, bool COUNT, int MA, class STK>
__device__ __forceinline__ Hit closest_hit_warp(const SceneView<SMEM>& sv, const SceneHeader& S, const Ray& r, float t_min, float t_max, bool brute, bool alive,
                                                Counters& cn, STK& stack) {
    if (COUNT && alive) cn.segments++;
    const Best b = closest_traced_warp<SMEM, COUNT>(sv, S, r, t_min, t_max, brute, alive, cn, stack);
    Hit h;
    h.t = b.t;
    h.ref = b.ref;
    h.point = v3(0.f, 0.f, 0.f);
    h.normal = v3(0.f, 0.f, 0.f);
    if (S.n_marched != 0u && alive) {
        // A ray whose [t_min, min(t_max, closest)] span misses the padded union of the marched
        // bounds can neither start within HIT_SIZE of a marched surface nor reach one: the
        // reference's loop would end without touching `rec`.  Skipping it is result-identical.
        const RayInv q = ray_inv(r);
        const float te = slab(S.march_lo[0], S.march_lo[1], S.march_lo[2], S.march_hi[0], S.march_hi[1], S.march_hi[2], q, t_min, fminf(t_max, h.t));
        if (te < kInf) {
            auto run = [&]() -> Hit {
                if constexpr (MA == 0) return march<SMEM, COUNT, false, false>(sv, S, r, t_min, t_max, h.t, &cn);
                else if constexpr (MA == 1) return march<SMEM, COUNT, true, false>(sv, S, r, t_min, t_max, h.t, &cn);
                else if constexpr (MA == 3) return march<SMEM, COUNT, false, true>(sv, S, r, t_min, t_max, h.t, &cn);
                else return S.n[omd::PT_MDYN] != 0u ? march<SMEM, COUNT, false, true>(sv, S, r, t_min, t_max, h.t, &cn)
                            : ((S.mgrid_dim[0] | S.n_mbvh_nodes) != 0u ? march<SMEM, COUNT, true, false>(sv, S, r, t_min, t_max, h.t, &cn)
                                                                       : march<SMEM, COUNT, false, false>(sv, S, r, t_min, t_max, h.t, &cn));
            };
            const Hit m = run();
            if (m.ref != omd::REF_NONE) h = m;
        }
    }
    return h;
}

// ---- hit record of the winner (point, normal) and material ------------------------------------
struct Material {
    V3 albedo;
    float fuzz, ior;
    uint32_t mat_type;
    uint32_t bloom_lo, bloom_hi;
};
template <bool SMEM>
__device__ __forceinline__ Material load_material(const SceneView<SMEM>& sv, const SceneHeader& S, uint32_t ref) {
    const uint32_t u = S.off_mat + omd::U_MAT * (S.mat_base[omd::ref_type(ref)] + omd::ref_index(ref));
    const float4 m0 = sv.cold(u), m1 = sv.cold(u + 1);
    Material m;
    m.albedo = v3(m0.x, m0.y, m0.z); m.fuzz = m0.w; m.ior = m1.x; m.mat_type = __float_as_uint(m1.y);
    m.bloom_lo = __float_as_uint(m1.z); m.bloom_hi = __float_as_uint(m1.w);
    return m;
}
template <bool SMEM, bool DYN = true>
__device__ __forceinline__ void finish_hit(const SceneView<SMEM>& sv, const SceneHeader& S, const Ray& r, float t_min, Hit& h) {
    const uint32_t type = omd::ref_type(h.ref), i = omd::ref_index(h.ref);
    if (type == omd::PT_SPHERE) {  // traced.rs:57-61
        const uint32_t u = S.off_hot[omd::PT_SPHERE] + omd::U_XF * i;
        const float4 r0 = sv.hot(u), r1 = sv.hot(u + 1), r2 = sv.hot(u + 2);
        Ray lr; lr.o = xf_point(r0, r1, r2, r.o); lr.d = xf_vec(r0, r1, r2, r.d);
        const V3 local_point = ray_at(lr, h.t);
        const uint32_t c = S.off_sph_l2w + omd::U_XF * i;
        const float4 m0 = sv.cold(c), m1 = sv.cold(c + 1), m2 = sv.cold(c + 2);
        h.point = xf_point(m0, m1, m2, local_point);
        h.normal = om::unit(xf_vec(m0, m1, m2, local_point));
    } else if (type == omd::PT_CUBE) {  // traced.rs:290-297
        const uint32_t u = S.off_hot[omd::PT_CUBE] + omd::U_XF * i;
        const float4 r0 = sv.hot(u), r1 = sv.hot(u + 1), r2 = sv.hot(u + 2);
        Ray lr; lr.o = xf_point(r0, r1, r2, r.o); lr.d = xf_vec(r0, r1, r2, r.d);
        float t; int axis = 0;
        cube_t(r0, r1, r2, r, t_min, h.t, t, axis);  // recover the face axis of the accepted root
        const V3 local_point = ray_at(lr, h.t);
        const float s = copysignf(1.f, om::comp(local_point, axis));
        const V3 ln = v3(axis == 0 ? 1.f : 0.f, axis == 1 ? 1.f : 0.f, axis == 2 ? 1.f : 0.f) * s;
        const uint32_t c = S.off_cube_l2w + omd::U_XF * i;
        const float4 m0 = sv.cold(c), m1 = sv.cold(c + 1), m2 = sv.cold(c + 2);
        h.point = xf_point(m0, m1, m2, local_point);
        h.normal = xf_vec(m0, m1, m2, ln);  // not normalised (reference quirk)
    } else if (type == omd::PT_PLANE) {  // traced.rs:111-113
        const uint32_t u = S.off_hot[omd::PT_PLANE] + omd::U_PLANE * i;
        const float4 pn = sv.hot(u);
        const V3 n = v3(pn.x, pn.y, pn.z);
        const float div = om::dot(n, r.d);
        h.point = ray_at(r, h.t);
        h.normal = n * copysignf(1.f, -div);
    } else if (type == omd::PT_TDYN) {  // om_dyn_traced: point = M . p_local, normal = unit(M . grad q / 2), as Sphere and Cube carry theirs (traced.rs:57-61,290-297)
        const uint32_t u = S.off_hot[omd::PT_TDYN] + omd::U_TDYN * i;
        const float4 r0 = sv.hot(u), r1 = sv.hot(u + 1), r2 = sv.hot(u + 2), qa = sv.hot(u + 3), qb = sv.hot(u + 4), qc = sv.hot(u + 5);
        Ray lr; lr.o = xf_point(r0, r1, r2, r.o); lr.d = xf_vec(r0, r1, r2, r.d);
        const V3 p = ray_at(lr, h.t);
        const V3 half_grad = v3(qa.x * p.x + 0.5f * (qa.w * p.y + qb.x * p.z + qb.z), qa.y * p.y + 0.5f * (qa.w * p.x + qb.y * p.z + qb.w), qa.z * p.z + 0.5f * (qb.x * p.x + qb.y * p.y + qc.x));
        const uint32_t c = S.off_tdyn_l2w + omd::U_XF * i;
        const float4 m0 = sv.cold(c), m1 = sv.cold(c + 1), m2 = sv.cold(c + 2);
        h.point = xf_point(m0, m1, m2, p);
        h.normal = om::unit(xf_vec(m0, m1, m2, half_grad));
    } else if (type == omd::PT_TRI || type == omd::PT_PARA) {  // traced.rs:198-199
        const uint32_t u = S.off_hot[type] + omd::U_BARY * i;
        const float4 b = sv.hot(u + 1);
        const V3 n = v3(b.x, b.y, b.z);
        const float div = om::dot(n, r.d);
        h.point = ray_at(r, h.t);
        h.normal = n * copysignf(1.f, -div);
    } else {  // marched hit (hits.rs:322-328): the accepted point and the nearest object's normal there
        h.point = ray_at(r, h.t);
        h.normal = marched_normal<SMEM, DYN>(sv, S, h.ref, h.point);
    }
}

// ---- materials.rs --------------------------------------------------------------------------------
__device__ __forceinline__ V3 reflect(V3 v, V3 n) { return v - 2.f * om::dot(v, n) * n; }  // :98-100
__device__ __forceinline__ V3 refract(V3 uv, V3 n, float etai_over_etat) {                 // :102-108
    const float cos_theta = fminf(om::dot(-uv, n), 1.0f);
    const V3 r_out_perp = etai_over_etat * (uv + cos_theta * n);
    const float aux = -sqrtf(fabsf(1.0f - om::length_squared(r_out_perp)));
    return r_out_perp + aux * n;
}
__device__ __forceinline__ float reflectance(float cosv, float ref_idx) {  // :110-116
    const float r0 = (1.f - ref_idx) / (1.f + ref_idx);
    const float r0_2 = r0 * r0;
    const float cos_5 = (1.f - cosv) * (1.f - cosv) * (1.f - cosv) * (1.f - cosv) * (1.f - cosv);
    return r0_2 + (1.f - r0_2) * cos_5;
}
// Material::scatter (:39-95).  Returns the attenuation; `r` becomes the scattered ray.
__device__ __forceinline__ V3 scatter(const Material& m, Ray& r, V3 point, V3 normal, Rng& g) {
    V3 new_dir, att;
    if (m.mat_type != OM_MAT_DIELECTRIC) {
        // Lambertian (:52-61) and Metal (:62-66, no absorb test) both start their draws with
        // rand_in_unit_sphere (metal even when fuzz == 0): one shared rejection loop.
        const V3 s = rand_in_unit_sphere(g);
        if (m.mat_type == OM_MAT_LAMBERTIAN) {
            new_dir = normal + om::unit(s);
            const float eps = 1e-8f;
            if (fabsf(new_dir.x) < eps && fabsf(new_dir.y) < eps && fabsf(new_dir.z) < eps) new_dir = normal;
        } else {
            new_dir = reflect(r.d, normal) + m.fuzz * s;
        }
        att = m.albedo;
    } else {  // :68-95
        const V3 dir_unit = r.d;
        const bool front_face = om::dot(dir_unit, normal) < 0.f;
        const float refraction_ratio = front_face ? 1.0f / m.ior : m.ior;
        const V3 n = front_face ? normal : -normal;
        const float cos_theta = fminf(om::dot(-dir_unit, n), 1.0f);
        const float sin_theta = sqrtf(1.0f - cos_theta * cos_theta);
        const bool cannot_refract = (refraction_ratio * sin_theta) > 1.0f;
        const bool reflect_by_reflectance = reflectance(cos_theta, refraction_ratio) > rng_f32(g);
        new_dir = (cannot_refract || reflect_by_reflectance) ? reflect(dir_unit, n) : refract(dir_unit, n, refraction_ratio);
        att = v3(1.f, 1.f, 1.f);
    }
    r.o = point;
    r.d = om::unit(new_dir);  // Ray::new, ray.rs:11-13
    return att;
}

// ---- camera.rs:60-74, render_thread.rs:183-192 -------------------------------------------------------
__device__ __forceinline__ Ray camera_ray(const om_camera& c, float u, float v, V3 disc) {
    const V3 rand_in_lens = c.lens_radius * disc;
    const V3 up = v3(c.u_of_plane[0], c.u_of_plane[1], c.u_of_plane[2]), vp = v3(c.v_of_plane[0], c.v_of_plane[1], c.v_of_plane[2]);
    const V3 offset = up * rand_in_lens.x + vp * rand_in_lens.y;
    // uv_to_dir().dot((u,v,0,1)): rows (horizontal_i, vertical_i, 0, (llc-origin)_i)
    const V3 origin = v3(c.origin[0], c.origin[1], c.origin[2]);
    const V3 llc_o = v3(c.lower_left_corner[0], c.lower_left_corner[1], c.lower_left_corner[2]) - origin;
    V3 direction;
    if (kExact)
        direction = v3(c.horizontal[0] * u + c.vertical[0] * v + 0.f * 0.f + llc_o.x, c.horizontal[1] * u + c.vertical[1] * v + 0.f * 0.f + llc_o.y,
                       c.horizontal[2] * u + c.vertical[2] * v + 0.f * 0.f + llc_o.z);
    else
        direction = v3(c.horizontal[0] * u + c.vertical[0] * v + llc_o.x, c.horizontal[1] * u + c.vertical[1] * v + llc_o.y, c.horizontal[2] * u + c.vertical[2] * v + llc_o.z);
    Ray r;
    r.o = origin + offset;
    r.d = om::unit(om::unit(direction - offset));  // .unit() in get_ray, again in Ray::new
    return r;
}
__device__ __forceinline__ Ray primary_ray(const RenderArgs& A, uint32_t pxl, uint32_t sample, Rng& g) {
    const uint32_t line = pxl / A.width, col = pxl - A.width * line;
    const uint32_t code = A.jitter[sample];
    float jr[2];
    rng_take<2>(g, jr);
    const float i_rand = (jr[0] + (float)((code >> 1) & 1u)) / 2.f;
    const float j_rand = (jr[1] + (float)(code & 1u)) / 2.f;
    const float u = ((float)col + i_rand) / ((float)A.width - 1.f);
    const float v = 1.0f - ((float)line + j_rand) / ((float)A.height - 1.f);
    const V3 disc = rand_in_unit_disc(g);
    return camera_ray(A.cam, u, v, disc);
}
__device__ __forceinline__ Ray center_ray(const RenderArgs& A, uint32_t pxl) {
    const uint32_t line = pxl / A.width, col = pxl - A.width * line;
    const float u = ((float)col + 0.5f) / ((float)A.width - 1.f);
    const float v = 1.0f - ((float)line + 0.5f) / ((float)A.height - 1.f);
    return camera_ray(A.cam, u, v, v3(0.f, 0.f, 0.f));
}

// ---- Stats (render_thread.rs:9-40), tone map (utils.rs:12-17, vec3.rs:73-76) ---------------------------------
__device__ __forceinline__ uint32_t f32_as_u8(float f) {  // Rust `as u8`: saturating, NaN -> 0
    if (!(f == f) || f <= 0.f) return 0u;
    if (f >= 255.f) return 255u;
    return (uint32_t)f;
}
__device__ __forceinline__ float tone1(float c) { return fminf(fmaxf(sqrtf(c), 0.f), 0.999f) * 256.0f; }
__device__ __forceinline__ uint32_t tone_map_u8x3(V3 avg) {
    return f32_as_u8(tone1(avg.x)) | (f32_as_u8(tone1(avg.y)) << 8) | (f32_as_u8(tone1(avg.z)) << 16);
}
struct Stats {
    V3 sum;
    uint32_t n;
    float avg_depth;
    uint32_t bad_avgs;
    uint32_t bloom_lo, bloom_hi;
    uint32_t color;  // packed r | g<<8 | b<<16
};
__device__ __forceinline__ bool stats_add(Stats& s, V3 x, float depth, uint32_t bloom_lo, uint32_t bloom_hi) {
    const uint32_t old_color = s.color;
    s.sum = s.sum + x;
    s.n += 1;
    s.color = tone_map_u8x3(om::div(s.sum, (float)s.n));
    const uint32_t bad_run = (old_color == s.color) ? 1u : 0u;
    s.bad_avgs += bad_run;
    s.bad_avgs *= bad_run;
    s.avg_depth = (((float)s.n - 1.f) * s.avg_depth + depth) / (float)s.n;
    s.bloom_lo |= bloom_lo;
    s.bloom_hi |= bloom_hi;
    return s.bad_avgs >= 5u;
}
__device__ __forceinline__ uint32_t stats_color_of(V3 sum, uint32_t n) {
    return n == 0u ? 0u : tone_map_u8x3(om::div(sum, (float)n));  // Stats::new: (0,0,0)
}

// ---- work -> pixel maps ---------------------------------------------------------------------------
// Without a partition the image is walked in 8x4 tiles (one warp = one tile: coherent primary
// rays).  With the interleaved-chunk partition (main.rs:172-189) slot k is the k-th pixel of the
// handle's own 2730-pixel chunks.
__device__ __forceinline__ uint32_t work_to_pixel(const RenderArgs& A, uint32_t k) {
    if (A.live_list) return k < *A.live_count ? A.live_list[k] : 0xFFFFFFFFu;
    if (A.part_world > 1u) {
        const uint32_t c = (k / OM_CHUNK_PIXELS) * A.part_world + A.part_rank;
        const uint32_t p = c * OM_CHUNK_PIXELS + (k % OM_CHUNK_PIXELS);
        return p < A.npx ? p : 0xFFFFFFFFu;
    }
    const uint32_t tile = k >> 5, in = k & 31u;
#if OM_TILE_BLOCKS
    // 8 consecutive tickets = a block of 2 x 4 tiles = 16 x 16 pixels (the run a phase-aligned warp group draws from)
    uint32_t blk = tile >> 3;
    if (A.block_order) blk = __ldg(A.block_order + blk);
    const uint32_t i = tile & 7u, bx_n = (A.tiles_x + 1u) >> 1;
    const uint32_t tx = (blk % bx_n) * 2u + (i & 1u), ty = (blk / bx_n) * 4u + (i >> 1);
    const uint32_t x = tx * 8u + (in & 7u), y = ty * 4u + (in >> 3);
#else
    const uint32_t x = (tile % A.tiles_x) * 8u + (in & 7u), y = (tile / A.tiles_x) * 4u + (in >> 3);
#endif
    return (x < A.width && y < A.height) ? y * A.width + x : 0xFFFFFFFFu;
}

// ---- staging the hot image into shared memory: TMA 1-D bulk copy + mbarrier ------------------------
__device__ __forceinline__ void stage_scene_smem(const float4* hot_g, uint32_t units) {
    __shared__ __align__(8) unsigned long long bar;
    const uint32_t bar_s = (uint32_t)__cvta_generic_to_shared(&bar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_s));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t bytes = units * 16u;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_s), "r"(bytes) : "memory");
        uint32_t done = 0;
        while (done < bytes) {
            const uint32_t chunk = min(bytes - done, 32768u);
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(reinterpret_cast<char*>(om_smem) + done);
            const char* src = reinterpret_cast<const char*>(hot_g) + done;
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(chunk), "r"(bar_s)
                         : "memory");
            done += chunk;
        }
    }
    // every thread waits for phase 0 of the barrier
    uint32_t ok = 0;
    while (!ok) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar_s) : "memory");
    }
}

// Per-lane state of the path state machine (one path segment per loop iteration).
struct PathLane {
    Rng g;
    Ray ray;
    V3 color;
    uint32_t seg;
    float first_t;
    uint32_t first_lo, first_hi;
    bool fresh;
    bool marching;       // BOUNDED: the lane's ray is part-way through its march
    float m_t, pend_t;   // ... how far, and the traced hit that stands if the march finds nothing
    uint32_t m_left, pend_ref;
    __device__ __forceinline__ void init(const uint32_t* rk) {
        rng_init(g, rk, 0u, 0u);
        g.b0 = g.b1 = g.b2 = g.b3 = 0u;
        ray.o = v3(0.f, 0.f, 0.f); ray.d = v3(0.f, 0.f, 1.f);
        color = v3(1.f, 1.f, 1.f);
        seg = 0; first_t = kInf; first_lo = 0; first_hi = 0;
        fresh = true; marching = false; m_t = 0.f; pend_t = 0.f; m_left = 0xFFFFFFFFu; pend_ref = omd::REF_NONE;
    }
};
struct PathTotals {
    unsigned long long credited, traced;
    uint32_t warp_iters, lane_iters;  // COUNT: lane occupancy of the path loop
};

// One iteration of the path loop for a whole warp: every lane with todo > 0 resolves (at most) one segment of its pixel's current
// sample - new sample, closest hit, shade or sky, Stats::add at the end of a path.  Shared by k_render (a warp per tile) and
// k_render_lanes (a slot per lane).
template <bool SMEM, bool COUNT, int MA, class STK>
__device__ __forceinline__ void path_iteration(const RenderArgs& A, const SceneView<SMEM>& sv, const SceneHeader& S, bool brute, bool adaptive, bool align, uint32_t nseg,
                                               uint32_t pxl, Stats& st, uint32_t& todo, PathLane& L, Counters& cn, PathTotals& T, STK& stack) {
    constexpr bool BOUNDED = OM_BOUNDED_MARCH_ALL || MA == 1 || MA == 3;  // scenes built around marched objects (C3); C2's kernel keeps march()
    Rng& g = L.g;
    Ray& ray = L.ray;
    V3& color = L.color;
    uint32_t& seg = L.seg;
    float& first_t = L.first_t;
    uint32_t &first_lo = L.first_lo, &first_hi = L.first_hi;
    bool &fresh = L.fresh, &marching = L.marching;
    float &m_t = L.m_t, &pend_t = L.pend_t;
    uint32_t &m_left = L.m_left, &pend_ref = L.pend_ref;
    unsigned long long &credited = T.credited, &traced = T.traced;
    uint32_t &warp_iters = T.warp_iters, &lane_iters = T.lane_iters;
    const bool alive = todo > 0u;
    if (COUNT) { warp_iters++; lane_iters += alive ? 1u : 0u; }
    if (align) phase_bar_count(1);
    if (alive && fresh) {
        const uint32_t sample = A.sample_base + st.n;  // jitters[pixel.stats.n]
        rng_init(g, A.rk, pxl, sample);
        ray = primary_ray(A, pxl, sample, g);
        color = v3(1.f, 1.f, 1.f);
        seg = 0; first_t = kInf; first_lo = 0; first_hi = 0;
        fresh = false;
    }
    Hit h;
    bool resolved = alive;  // this lane's segment is resolved in this iteration (shade it)
    if constexpr (BOUNDED) {
        // Bounded marching: a lane whose ray enters the marched objects' bounds sphere-traces at most OM_MARCH_STEPS steps per
        // iteration and carries on in the next one, while the other lanes of the warp shade, start samples and traverse.
        // (One call of march() ran as long as the warp's longest march: 5.6 of 32 lanes per instruction on C3.)
        const bool trav = alive && !marching;
        if (COUNT && trav) cn.segments++;
        const Best b = closest_traced_warp<SMEM, COUNT>(sv, S, ray, A.tmin, A.tmax, brute, trav, cn, stack);
        h.t = pend_t; h.ref = pend_ref; h.point = v3(0.f, 0.f, 0.f); h.normal = v3(0.f, 0.f, 0.f);
        if (trav) {
            h.t = b.t; h.ref = b.ref;
            if (S.n_marched != 0u) {  // (see closest_hit_warp)
                const RayInv q = ray_inv(ray);
                if (slab(S.march_lo[0], S.march_lo[1], S.march_lo[2], S.march_hi[0], S.march_hi[1], S.march_hi[2], q, A.tmin, fminf(A.tmax, b.t)) < kInf) {
                    marching = true; m_left = 0xFFFFFFFFu; m_t = 0.f; pend_t = b.t; pend_ref = b.ref;
                }
            }
        }
        if (alive && marching) {
            Hit m;
            const bool over = march_some<SMEM, COUNT, MA == 1, MA == 3>(sv, S, ray, A.tmin, A.tmax, pend_t, m_t, m_left, OM_MARCH_STEPS, m, &cn);
            resolved = over;
            if (over) { marching = false; if (m.ref != omd::REF_NONE) h = m; else { h.t = pend_t; h.ref = pend_ref; } }
        }
    } else {
        h = closest_hit_warp<SMEM, COUNT, MA>(sv, S, ray, A.tmin, A.tmax, brute, alive, cn, stack);
    }
    if (align) phase_bar();
    if (resolved) {
        bool finished;
        V3 result = color;
        if (h.ref != omd::REF_NONE) {  // handle_hit Some(hr), render_thread.rs:110-116
            finish_hit<SMEM, MA == 3>(sv, S, ray, A.tmin, h);
            const Material m = load_material(sv, S, h.ref);
            if (seg == 0u) { first_t = h.t; first_lo = m.bloom_lo; first_hi = m.bloom_hi; }
            color = color * scatter(m, ray, h.point, h.normal, g);
            seg++;
            finished = seg >= nseg;
            if (finished) result = -v3(0.f, 0.f, 0.f);  // out of depth: -black (:142)
        } else {  // None: sky (:117-123)
            const float t = 0.5f * (ray.d.y + 1.0f);
            const V3 sky = (1.0f - t) * v3(1.f, 1.f, 1.f) + t * v3(0.5f, 0.7f, 1.0f);
            result = color * sky;
            finished = true;
        }
        if (finished) {
            bool done = stats_add(st, result, first_t, first_lo, first_hi);
            if (!adaptive) done = false;
            traced++;
            todo--;
            const uint32_t left = A.spp_budget - st.n;
            credited += (done ? (unsigned long long)left : 0ull) + 1ull;  // render_thread.rs:196-198
            if (done) todo = 0u;
            fresh = true;
        }
    }

}

// ============================================================================================
// k_render: persistent path-regeneration kernel.  One lane owns one pixel for the whole pass and
// runs a per-lane state machine (new sample -> segment -> shade -> ... -> Stats::add), so lanes
// never wait for each other's paths; the accumulators live in registers between the single
// coalesced read and write of the pixel's 32-byte state.
// ============================================================================================
template <bool SMEM, bool COUNT, int MA>
__global__ void __launch_bounds__(RENDER_THREADS, OM_RENDER_MIN_CTAS) k_render(const __grid_constant__ RenderArgs A) {
    if (A.spread_max != 0u && A.live_count != nullptr && *A.live_count <= A.spread_max) return;  // k_render_spread's pass
    if (SMEM) stage_scene_smem(A.hot_g, A.scene.hot_units);
    SceneView<SMEM> sv;
    sv.hot_g = A.hot_g;
    sv.cold_g = A.cold_g;
    const SceneHeader& S = A.scene;
    const bool brute = (A.flags & OM_RENDER_BRUTE_FORCE) != 0;
    const bool adaptive = (A.flags & OM_RENDER_ADAPTIVE) != 0;
    const bool align = (A.flags & OM_RENDER_PHASE_ALIGN) != 0;
    Counters cn = {0, 0, 0, 0, 0, 0, 0};
    PathTotals T = {0ull, 0ull, 0u, 0u};
    const uint32_t lane = threadIdx.x & 31u;
    TravStack<SMEM, RENDER_THREADS> stack;  // shared memory behind the scene image when the scene is staged, else local memory
    stack.init(S.hot_units);
    const uint32_t nseg = A.max_depth > 1u ? A.max_depth : 1u;  // segment 0 + `for _i in 1..depth`
#if OM_GROUP_TICKETS
    __shared__ unsigned grp_lock[PB_GROUPS], grp_next[PB_GROUPS], grp_left[PB_GROUPS];
    if (threadIdx.x < (unsigned)PB_GROUPS) { grp_lock[threadIdx.x] = 0u; grp_next[threadIdx.x] = 0u; grp_left[threadIdx.x] = 0u; }
    __syncthreads();
#endif

    for (;;) {
        // one ticket = one warp-sized run of 32 work slots (an 8x4 pixel tile)
        uint32_t ticket = 0;
#if OM_GROUP_TICKETS
        // the warps of a phase-aligned group draw from a shared run of OM_GROUP_TICKETS consecutive tickets (adjacent tiles:
        // similar traversal work, less waiting at the group's barriers); single tickets near the end of the launch
        if (lane == 0u) {
            if (align) {
                const uint32_t g = (threadIdx.x >> 5) % (unsigned)PB_GROUPS;
                while (atomicCAS(&grp_lock[g], 0u, 1u) != 0u) {}
                __threadfence_block();
                if (grp_left[g] == 0u) {
                    const uint32_t n_tickets = (A.n_work + 31u) / 32u;
                    const uint32_t want = (grp_next[g] + 4u * (uint32_t)OM_GROUP_TICKETS * gridDim.x * (unsigned)PB_GROUPS < n_tickets) ? (uint32_t)OM_GROUP_TICKETS : 1u;
                    grp_next[g] = (uint32_t)atomicAdd(&A.counters[CTR_TICKET], (unsigned long long)want);
                    grp_left[g] = want;
                }
                ticket = grp_next[g];
                grp_next[g] = ticket + 1u;
                grp_left[g] -= 1u;
                __threadfence_block();
                atomicExch(&grp_lock[g], 0u);
            } else ticket = (uint32_t)atomicAdd(&A.counters[CTR_TICKET], 1ull);
        }
#else
        if (lane == 0u) ticket = (uint32_t)atomicAdd(&A.counters[CTR_TICKET], 1ull);
#endif
        ticket = __shfl_sync(FULL, ticket, 0);
        const uint32_t base = ticket * 32u;
        if (base >= A.n_work) break;
        const uint32_t tile_t0 = (uint32_t)clock64();
        const uint32_t k = base + lane;
        const uint32_t pxl = k < A.n_work ? work_to_pixel(A, k) : 0xFFFFFFFFu;
        const bool valid = pxl != 0xFFFFFFFFu;

        // ---- load Stats ----
        Stats st;
        st.sum = v3(0.f, 0.f, 0.f); st.n = 0; st.avg_depth = 0.f; st.bad_avgs = 0; st.bloom_lo = 0; st.bloom_hi = 0; st.color = 0;
        uint32_t todo = 0;
        if (valid) {
            const float4 a = A.acc[pxl];
            const uint4 x = A.aux[pxl];
            st.sum = v3(a.x, a.y, a.z); st.n = __float_as_uint(a.w);
            st.avg_depth = __uint_as_float(x.x); st.bad_avgs = x.y; st.bloom_lo = x.z; st.bloom_hi = x.w;
            st.color = stats_color_of(st.sum, st.n);
            todo = A.spp_budget > st.n ? min(A.spp_budget - st.n, A.spp_pass) : 0u;
            if (A.n_target) { const uint32_t tg = min(A.spp_budget, A.n_target); todo = tg > st.n ? tg - st.n : 0u; }
            if (adaptive && st.bad_avgs >= 5u) todo = 0u;
        }
        const bool touched = todo > 0u;

        // ---- per-lane path state machine; the loop itself is warp-uniform ----
        PathLane L;
        L.init(A.rk);
        while (__any_sync(FULL, todo > 0u)) path_iteration<SMEM, COUNT, MA>(A, sv, S, brute, adaptive, align, nseg, pxl, st, todo, L, cn, T, stack);
        // ---- store Stats ----
        if (touched) {
            A.acc[pxl] = make_float4(st.sum.x, st.sum.y, st.sum.z, __uint_as_float(st.n));
            A.aux[pxl] = make_uint4(__float_as_uint(st.avg_depth), st.bad_avgs, st.bloom_lo, st.bloom_hi);
        }
        // what this tile cost, filed under its place in the image (the host orders the next launch's blocks by it)
        if (A.tile_cost && lane == 0u) {
            uint32_t blk = ticket >> 3;
            if (A.block_order) blk = __ldg(A.block_order + blk);
            A.tile_cost[blk * 8u + (ticket & 7u)] = ((uint32_t)clock64() - tile_t0) >> 4;  // 32-bit cycle counter: tiles are far shorter than its 2 s period
        }
    }

    // out of tickets: keep answering the group's barriers until all of its warps are out
    if (align) {
        while (phase_bar_count(0) > 0) phase_bar();
    }
    // ---- progress counters: warp-aggregate, one atomic per warp ----
    for (int o = 16; o > 0; o >>= 1) {
        T.credited += __shfl_xor_sync(FULL, T.credited, o);
        T.traced += __shfl_xor_sync(FULL, T.traced, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (T.credited) atomicAdd(&A.counters[CTR_CREDITED], T.credited);
        if (T.traced) atomicAdd(&A.counters[CTR_TRACED], T.traced);
    }
    if (COUNT) {
        unsigned long long v[7] = {cn.segments, cn.node_visits, cn.prim_tests, cn.sdf_evals, cn.march_steps, cn.normal_evals, cn.march_entered};
        for (int i = 0; i < 7; i++) {
            for (int o = 16; o > 0; o >>= 1) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
            if ((threadIdx.x & 31) == 0 && v[i]) atomicAdd(&A.counters[CTR_SEGMENTS + i], v[i]);
        }
        unsigned long long li = T.lane_iters;
        for (int o = 16; o > 0; o >>= 1) li += __shfl_xor_sync(0xffffffffu, li, o);
        if ((threadIdx.x & 31) == 0) { atomicAdd(&A.counters[CTR_WARP_ITERS], (unsigned long long)T.warp_iters); atomicAdd(&A.counters[CTR_LANE_ITERS], li); }
    }
}

// ============================================================================================
// k_render_lanes: k_render with a work slot per LANE instead of a tile per warp, for adaptive passes.  There a warp that owns
// a tile is held by the pixel that retires last while the others' lanes idle (lane occupancy of the adaptive passes 0.4-0.66;
// a single 256-spp adaptive launch reaches 800 Msamples/s where fixed spp reaches 1850).  Here a lane whose pixel is done
// stores it and takes the next unstarted slot at once.  The warps of a group (the 8 that share an SM sub-partition and, with
// phase alignment, its barriers) draw from a shared run of OM_LANE_RUN consecutive slots - a 16 x 16-pixel block in tile
// order (work_to_pixel) or a stretch of the live list, which is in tile order too: neighbouring pixels, similar work.
// A pixel's samples still run in order in one lane: bit-identical.  (For FIXED spp the same scheme loses to tiles: occupancy
// is 0.9 there already and mixing the pixels of a warp costs more than it gains, profiles/r2_experiments.md section 6.)
// ============================================================================================
#ifndef OM_LANE_RUN
#define OM_LANE_RUN 256
#endif
template <bool SMEM, int MA>
__global__ void __launch_bounds__(RENDER_THREADS, OM_RENDER_MIN_CTAS) k_render_lanes(const __grid_constant__ RenderArgs A) {
    if (A.spread_max != 0u && A.live_count != nullptr && *A.live_count <= A.spread_max) return;  // k_render_spread's pass
    if (SMEM) stage_scene_smem(A.hot_g, A.scene.hot_units);
    SceneView<SMEM> sv;
    sv.hot_g = A.hot_g;
    sv.cold_g = A.cold_g;
    const SceneHeader& S = A.scene;
    const bool brute = (A.flags & OM_RENDER_BRUTE_FORCE) != 0;
    const bool adaptive = (A.flags & OM_RENDER_ADAPTIVE) != 0;
    const bool align = (A.flags & OM_RENDER_PHASE_ALIGN) != 0;
    Counters cn = {0, 0, 0, 0, 0, 0, 0};
    PathTotals T = {0ull, 0ull, 0u, 0u};
    const uint32_t lane = threadIdx.x & 31u, lt_mask = (1u << lane) - 1u;
    TravStack<SMEM, RENDER_THREADS> stack;
    stack.init(S.hot_units);
    const uint32_t nseg = A.max_depth > 1u ? A.max_depth : 1u;
    const uint32_t n_slots = A.live_list ? min(A.n_work, *A.live_count) : A.n_work;
    __shared__ unsigned grp_lock[PB_GROUPS], grp_next[PB_GROUPS], grp_end[PB_GROUPS];
    if (threadIdx.x < (unsigned)PB_GROUPS) { grp_lock[threadIdx.x] = 0u; grp_next[threadIdx.x] = 0u; grp_end[threadIdx.x] = 0u; }
    __syncthreads();

    uint32_t pxl = 0xFFFFFFFFu, todo = 0;
    bool touched = false;    // the lane holds a pixel whose Stats changed
    bool exhausted = false;  // (warp-uniform) no slots left
    Stats st;
    st.sum = v3(0.f, 0.f, 0.f); st.n = 0; st.avg_depth = 0.f; st.bad_avgs = 0; st.bloom_lo = 0; st.bloom_hi = 0; st.color = 0;
    PathLane L;
    L.init(A.rk);
    for (;;) {
        // ---- lanes without work: store the finished pixel, take the next slot ----
        const unsigned m_need = __ballot_sync(FULL, todo == 0u);
        if (m_need != 0u) {
            if (todo == 0u && touched) {
                A.acc[pxl] = make_float4(st.sum.x, st.sum.y, st.sum.z, __uint_as_float(st.n));
                A.aux[pxl] = make_uint4(__float_as_uint(st.avg_depth), st.bad_avgs, st.bloom_lo, st.bloom_hi);
                touched = false;
            }
            if (!exhausted) {
                const uint32_t n_need = (uint32_t)__popc(m_need);
                uint32_t b1 = 0, m1 = 0, b2 = 0;
                if (lane == 0u) {  // n_need slots: what is left of the group's run, then the head of a new run
                    const uint32_t gi = (threadIdx.x >> 5) % (unsigned)PB_GROUPS;
                    while (atomicCAS(&grp_lock[gi], 0u, 1u) != 0u) {}
                    __threadfence_block();
                    const uint32_t have = grp_end[gi] - grp_next[gi];
                    m1 = min(n_need, have);
                    b1 = grp_next[gi];
                    grp_next[gi] = b1 + m1;
                    if (m1 < n_need) {
                        b2 = (uint32_t)atomicAdd(&A.counters[CTR_TICKET], (unsigned long long)OM_LANE_RUN);  // (this kernel's unit of CTR_TICKET: one slot)
                        grp_next[gi] = b2 + (n_need - m1);
                        grp_end[gi] = b2 + (uint32_t)OM_LANE_RUN;
                    }
                    __threadfence_block();
                    atomicExch(&grp_lock[gi], 0u);
                }
                b1 = __shfl_sync(FULL, b1, 0); m1 = __shfl_sync(FULL, m1, 0); b2 = __shfl_sync(FULL, b2, 0);
                exhausted = (m1 > 0u ? b1 : b2) >= n_slots;
                if (todo == 0u) {
                    const uint32_t r = (uint32_t)__popc(m_need & lt_mask);
                    const uint32_t k = r < m1 ? b1 + r : b2 + (r - m1);
                    pxl = k < n_slots ? work_to_pixel(A, k) : 0xFFFFFFFFu;
                    if (pxl != 0xFFFFFFFFu) {
                        const float4 a = A.acc[pxl];
                        const uint4 x = A.aux[pxl];
                        st.sum = v3(a.x, a.y, a.z); st.n = __float_as_uint(a.w);
                        st.avg_depth = __uint_as_float(x.x); st.bad_avgs = x.y; st.bloom_lo = x.z; st.bloom_hi = x.w;
                        st.color = stats_color_of(st.sum, st.n);
                        todo = A.spp_budget > st.n ? min(A.spp_budget - st.n, A.spp_pass) : 0u;
                        if (A.n_target) { const uint32_t tg = min(A.spp_budget, A.n_target); todo = tg > st.n ? tg - st.n : 0u; }
                        if (adaptive && st.bad_avgs >= 5u) todo = 0u;
                        touched = todo > 0u;
                        L.fresh = true;
                    }
                }
            }
        }
        if (!__any_sync(FULL, todo > 0u)) {
            if (exhausted) break;
            continue;
        }
        path_iteration<SMEM, false, MA>(A, sv, S, brute, adaptive, align, nseg, pxl, st, todo, L, cn, T, stack);
    }
    // out of slots: keep answering the group's barriers until all of its warps are out
    if (align) {
        while (phase_bar_count(0) > 0) phase_bar();
    }
    for (int o = 16; o > 0; o >>= 1) {
        T.credited += __shfl_xor_sync(FULL, T.credited, o);
        T.traced += __shfl_xor_sync(FULL, T.traced, o);
    }
    if (lane == 0u) {
        if (T.credited) atomicAdd(&A.counters[CTR_CREDITED], T.credited);
        if (T.traced) atomicAdd(&A.counters[CTR_TRACED], T.traced);
    }
}

// ============================================================================================
// k_render_spread: adaptive passes with few live pixels.  In k_render a lane owns a pixel and traces its samples one after
// the other, so a pass is as long as its hardest pixel's chain of spp_pass paths (depth-50 glass paths: 8 ms for a pass of
// 16 however few pixels are live, and the same on every GPU of a multi-GPU render).  Here `spread` (a power of two <= 32)
// consecutive lanes share one pixel of the live list: in each round lane i of the group traces sample n + i, then every
// lane of the group folds the round's results into its own copy of the pixel's Stats in sample order - the same adds in the
// same order as the sequential loop, so the Stats are bit-identical - and stops at the sample that retires the pixel
// (render_thread.rs:194-200); the samples traced past that point are thrown away and not counted.
// Plain structure (no phase alignment, no tile costs, marches run to their end): these passes are latency-, not issue-bound.
// ============================================================================================
template <bool SMEM, int MA>
__global__ void __launch_bounds__(RENDER_THREADS, OM_RENDER_MIN_CTAS) k_render_spread(const __grid_constant__ RenderArgs A) {
    if (A.live_count == nullptr || *A.live_count > A.spread_max) return;  // k_render's pass
    if (SMEM) stage_scene_smem(A.hot_g, A.scene.hot_units);
    SceneView<SMEM> sv;
    sv.hot_g = A.hot_g;
    sv.cold_g = A.cold_g;
    const SceneHeader& S = A.scene;
    const bool brute = (A.flags & OM_RENDER_BRUTE_FORCE) != 0;
    const bool adaptive = (A.flags & OM_RENDER_ADAPTIVE) != 0;
    Counters cn = {0, 0, 0, 0, 0, 0, 0};
    unsigned long long credited = 0, traced = 0;
    const uint32_t lane = threadIdx.x & 31u;
    TravStack<SMEM, RENDER_THREADS> stack;
    stack.init(S.hot_units);
    const uint32_t nseg = A.max_depth > 1u ? A.max_depth : 1u;
    const uint32_t SP = A.spread, sub = lane & (SP - 1u), lead = lane & ~(SP - 1u);
    const uint32_t n_live = *A.live_count;
    const unsigned long long n_slots = (unsigned long long)n_live * SP;
    for (;;) {
        uint32_t ticket = 0;
        if (lane == 0u) ticket = (uint32_t)atomicAdd(&A.counters[CTR_TICKET], 1ull);
        ticket = __shfl_sync(FULL, ticket, 0);
        if ((unsigned long long)ticket * 32ull >= n_slots) break;
        const uint32_t pi = (ticket * 32u + lane) / SP;
        const uint32_t pxl = pi < n_live ? A.live_list[pi] : 0xFFFFFFFFu;
        // every lane of a group keeps its own copy of the pixel's Stats (the folds below are the same on all of them)
        Stats st;
        st.sum = v3(0.f, 0.f, 0.f); st.n = 0; st.avg_depth = 0.f; st.bad_avgs = 0; st.bloom_lo = 0; st.bloom_hi = 0; st.color = 0;
        uint32_t todo = 0;
        if (pxl != 0xFFFFFFFFu) {
            const float4 a = A.acc[pxl];
            const uint4 x = A.aux[pxl];
            st.sum = v3(a.x, a.y, a.z); st.n = __float_as_uint(a.w);
            st.avg_depth = __uint_as_float(x.x); st.bad_avgs = x.y; st.bloom_lo = x.z; st.bloom_hi = x.w;
            st.color = stats_color_of(st.sum, st.n);
            todo = A.spp_budget > st.n ? min(A.spp_budget - st.n, A.spp_pass) : 0u;
            if (A.n_target) { const uint32_t tg = min(A.spp_budget, A.n_target); todo = tg > st.n ? tg - st.n : 0u; }
            if (adaptive && st.bad_avgs >= 5u) todo = 0u;
        }
        const bool touched = todo > 0u;
        while (__any_sync(FULL, todo > 0u)) {
            // ---- one round: lane `sub` of a group traces sample st.n + sub ----
            bool tracing = sub < todo;
            Rng g;
            rng_init(g, A.rk, 0u, 0u);
            g.b0 = g.b1 = g.b2 = g.b3 = 0u;
            Ray ray;
            ray.o = v3(0.f, 0.f, 0.f); ray.d = v3(0.f, 0.f, 1.f);
            V3 color = v3(1.f, 1.f, 1.f), result = v3(0.f, 0.f, 0.f);
            uint32_t seg = 0, first_lo = 0, first_hi = 0;
            float first_t = kInf;
            if (tracing) {
                const uint32_t sample = A.sample_base + st.n + sub;  // jitters[pixel.stats.n]
                rng_init(g, A.rk, pxl, sample);
                ray = primary_ray(A, pxl, sample, g);
            }
            while (__any_sync(FULL, tracing)) {
                Hit h = closest_hit_warp<SMEM, false, MA>(sv, S, ray, A.tmin, A.tmax, brute, tracing, cn, stack);
                if (tracing) {
                    if (h.ref != omd::REF_NONE) {  // handle_hit Some(hr), render_thread.rs:110-116
                        finish_hit<SMEM, MA == 3>(sv, S, ray, A.tmin, h);
                        const Material m = load_material(sv, S, h.ref);
                        if (seg == 0u) { first_t = h.t; first_lo = m.bloom_lo; first_hi = m.bloom_hi; }
                        color = color * scatter(m, ray, h.point, h.normal, g);
                        seg++;
                        if (seg >= nseg) { result = -v3(0.f, 0.f, 0.f); tracing = false; }  // out of depth: -black (:142)
                    } else {  // None: sky (:117-123)
                        const float t = 0.5f * (ray.d.y + 1.0f);
                        const V3 sky = (1.0f - t) * v3(1.f, 1.f, 1.f) + t * v3(0.5f, 0.7f, 1.0f);
                        result = color * sky;
                        tracing = false;
                    }
                }
            }
            // ---- fold the round into the Stats, sample by sample ----
            const uint32_t round_n = min(todo, SP);
            bool retired = false;
            for (uint32_t i = 0; i < SP; i++) {  // (warp-uniform trip count; groups with a shorter round skip)
                const uint32_t src = lead + i;
                const float rx = __shfl_sync(FULL, result.x, src), ry = __shfl_sync(FULL, result.y, src), rz = __shfl_sync(FULL, result.z, src);
                const float ft = __shfl_sync(FULL, first_t, src);
                const uint32_t flo = __shfl_sync(FULL, first_lo, src), fhi = __shfl_sync(FULL, first_hi, src);
                if (i < round_n && !retired) {
                    bool done = stats_add(st, v3(rx, ry, rz), ft, flo, fhi);
                    if (!adaptive) done = false;
                    if (sub == 0u) {
                        traced++;
                        const uint32_t left = A.spp_budget - st.n;
                        credited += (done ? (unsigned long long)left : 0ull) + 1ull;  // render_thread.rs:196-198
                    }
                    if (done) retired = true;
                }
            }
            todo = retired ? 0u : todo - round_n;
        }
        if (touched && sub == 0u) {
            A.acc[pxl] = make_float4(st.sum.x, st.sum.y, st.sum.z, __uint_as_float(st.n));
            A.aux[pxl] = make_uint4(__float_as_uint(st.avg_depth), st.bad_avgs, st.bloom_lo, st.bloom_hi);
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        credited += __shfl_xor_sync(FULL, credited, o);
        traced += __shfl_xor_sync(FULL, traced, o);
    }
    if (lane == 0u) {
        if (credited) atomicAdd(&A.counters[CTR_CREDITED], credited);
        if (traced) atomicAdd(&A.counters[CTR_TRACED], traced);
    }
}

// ============================================================================================
// parity hooks
// ============================================================================================
// primary-ray closest hit per pixel: (obj_id, t) — what ray_color reports as (obj_id, depthf)
template <bool SMEM>
__global__ void __launch_bounds__(CTA_THREADS, 2) k_primary(const __grid_constant__ RenderArgs A, uint32_t mode, uint32_t sample_index, unsigned long long* ids, float* tout) {
    if (SMEM) stage_scene_smem(A.hot_g, A.scene.hot_units);
    SceneView<SMEM> sv;
    sv.hot_g = A.hot_g;
    sv.cold_g = A.cold_g;
    Counters cn = {0, 0, 0, 0, 0, 0, 0};
    TravStack<SMEM, CTA_THREADS> stack;
    stack.init(A.scene.hot_units);
    // warp-uniform trip count: closest_hit_warp is a warp-collective
    for (uint32_t base = blockIdx.x * CTA_THREADS; base < A.npx; base += gridDim.x * CTA_THREADS) {
        const uint32_t pxl = base + threadIdx.x;
        const bool alive = pxl < A.npx;
        Ray ray;
        ray.o = v3(0.f, 0.f, 0.f); ray.d = v3(0.f, 0.f, 1.f);
        if (alive) {
            if (mode == 0u) ray = center_ray(A, pxl);
            else { Rng g; rng_init(g, A.rk, pxl, sample_index); ray = primary_ray(A, pxl, sample_index, g); }
        }
        const Hit h = closest_hit_warp<SMEM, false, 2>(sv, A.scene, ray, A.tmin, A.tmax, (A.flags & OM_RENDER_BRUTE_FORCE) != 0, alive, cn, stack);
        if (alive) {
            if (h.ref == omd::REF_NONE) { ids[pxl] = 0ull; tout[pxl] = kInf; }
            else { ids[pxl] = ((unsigned long long)(omd::ref_type(h.ref) + 1u) << 32) | (unsigned long long)(omd::ref_index(h.ref) + 1u); tout[pxl] = h.t; }
        }
    }
}

// FrozenHittableList::hit for arbitrary rays (dirs normalised like Ray::new)
template <bool SMEM>
__global__ void __launch_bounds__(CTA_THREADS, 2) k_trace_rays(const __grid_constant__ TraceArgs A) {
    if (SMEM) stage_scene_smem(A.hot_g, A.scene.hot_units);
    SceneView<SMEM> sv;
    sv.hot_g = A.hot_g;
    sv.cold_g = A.cold_g;
    Counters cn = {0, 0, 0, 0, 0, 0, 0};
    TravStack<SMEM, CTA_THREADS> stack;
    stack.init(A.scene.hot_units);
    for (uint32_t base = blockIdx.x * CTA_THREADS; base < A.n; base += gridDim.x * CTA_THREADS) {
        const uint32_t i = base + threadIdx.x;
        const bool alive = i < A.n;
        Ray ray;
        ray.o = v3(0.f, 0.f, 0.f); ray.d = v3(0.f, 0.f, 1.f);
        if (alive) {
            ray.o = v3(A.origins[3 * i], A.origins[3 * i + 1], A.origins[3 * i + 2]);
            ray.d = om::unit(v3(A.dirs[3 * i], A.dirs[3 * i + 1], A.dirs[3 * i + 2]));
        }
        Hit h = closest_hit_warp<SMEM, false, 2>(sv, A.scene, ray, A.tmin, A.tmax, (A.flags & OM_RENDER_BRUTE_FORCE) != 0, alive, cn, stack);
        if (!alive) continue;
        const bool hit = h.ref != omd::REF_NONE;
        if (hit) finish_hit(sv, A.scene, ray, A.tmin, h);
        if (A.ids) A.ids[i] = hit ? (((unsigned long long)(omd::ref_type(h.ref) + 1u) << 32) | (unsigned long long)(omd::ref_index(h.ref) + 1u)) : 0ull;
        if (A.t) A.t[i] = hit ? h.t : kInf;
        if (A.points) { A.points[3 * i] = hit ? h.point.x : 0.f; A.points[3 * i + 1] = hit ? h.point.y : 0.f; A.points[3 * i + 2] = hit ? h.point.z : 0.f; }
        if (A.normals) { A.normals[3 * i] = hit ? h.normal.x : 0.f; A.normals[3 * i + 1] = hit ? h.normal.y : 0.f; A.normals[3 * i + 2] = hit ? h.normal.z : 0.f; }
    }
}

}  // namespace OM_NS
