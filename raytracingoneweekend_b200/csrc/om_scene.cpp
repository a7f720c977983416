#include "om_scene.h"

#include <math.h>
#include <string.h>

#include <stdio.h>
#include <stdlib.h>

#include <algorithm>
#include <chrono>
#include <limits>
#include <thread>

#include "om_bvh.h"

namespace oms {
namespace {
using omb::Aabb;
const float kInf = std::numeric_limits<float>::infinity();

// Pad a box so that every ray the reference's f32 intersection routine reports as a hit also
// overlaps the box.  `extra` is a per-axis world-space margin computed by the callers from the
// routine's conditioning; on top of it comes a small relative margin for the slab arithmetic.
Aabb padded(const double lo[3], const double hi[3], const double extra[3]) {
    Aabb b;
    for (int a = 0; a < 3; a++) {
        const double ext = hi[a] - lo[a];
        const double mag = std::max(std::fabs(lo[a]), std::fabs(hi[a]));
        const double eps = 1e-4 * std::max(ext, mag) + 1e-5 + extra[a];
        b.lo[a] = nextafterf(float(lo[a] - eps), -kInf);
        b.hi[a] = nextafterf(float(hi[a] + eps), kInf);
    }
    return b;
}
double col_norm_min(const float* m) {  // smallest scale factor of the linear part (columns of R*S)
    double best = 1e300;
    for (int c = 0; c < 3; c++) best = std::min(best, std::sqrt(double(m[c]) * m[c] + double(m[4 + c]) * m[4 + c] + double(m[8 + c]) * m[8 + c]));
    return std::max(best, 1e-30);
}

// traced.rs:65-70 boxes the unit sphere's +-1 cube; the exact extent of an ellipsoid M*S^2 along
// axis i is the norm of row i of the linear part.
// Sphere::hit (traced.rs:38-56) works in the sphere's local frame, where a ray origin `reach`
// away sits at |o_l| = rho = reach / scale radii: `c = |o_l|^2 - 1` and `half_b^2 - a*c` then carry
// an absolute f32 error of order rho^2 * 2^-23 against the "1", i.e. the test accepts rays that pass
// within sqrt(1 + k*rho^2*2^-24) radii.  The box must contain that fattened sphere.
constexpr double kSphereK = 32.0 * 5.96e-8;
double sphere_fat(const om_sphere& s, double reach) {
    const double rho = reach / col_norm_min(s.m_local_to_world);
    return std::sqrt(1.0 + kSphereK * rho * rho) - 1.0;
}
Aabb sphere_box(const om_sphere& s, double reach) {
    const float* m = s.m_local_to_world;
    const double fat = sphere_fat(s, reach);
    double lo[3], hi[3], extra[3];
    for (int i = 0; i < 3; i++) {
        const double h = std::sqrt(double(m[4 * i]) * m[4 * i] + double(m[4 * i + 1]) * m[4 * i + 1] + double(m[4 * i + 2]) * m[4 * i + 2]);
        lo[i] = m[4 * i + 3] - h;
        hi[i] = m[4 * i + 3] + h;
        extra[i] = h * fat;
    }
    return padded(lo, hi, extra);
}
// traced.rs:301-304 uses +-1; the cube really spans +-0.5 (traced.rs:273-274).  A hit is accepted
// when the computed local point is within 1e-5 of the surface (traced.rs:284); the computed point
// itself is off by about rho * 2^-22 local units.
constexpr double kCubeSlope = 1e-6, kBarySlope = 4e-6;
Aabb cube_box(const om_cube& s, double reach) {
    const float* m = s.m_local_to_world;
    const double rho = reach / col_norm_min(m);
    const double fat = 2e-5 + rho * kCubeSlope;
    double lo[3], hi[3], extra[3];
    for (int i = 0; i < 3; i++) {
        const double h = 0.5 * (std::fabs(double(m[4 * i])) + std::fabs(double(m[4 * i + 1])) + std::fabs(double(m[4 * i + 2])));
        lo[i] = m[4 * i + 3] - h;
        hi[i] = m[4 * i + 3] + h;
        extra[i] = 2.0 * h * fat;
    }
    return padded(lo, hi, extra);
}
// traced.rs:208-219 is marked "may be wrong" in the reference; use the real vertices.  The hit
// point o + t*d is rounded at the scale of the ray's reach.
Aabb bary_box(const om_barycentric& b, bool parallelogram, double reach) {
    double lo[3], hi[3], extra[3];
    for (int i = 0; i < 3; i++) {
        const double o = b.origin[i], pu = o + double(b.u[i]) * b.u_length, pv = o + double(b.v[i]) * b.v_length;
        lo[i] = std::min(o, std::min(pu, pv));
        hi[i] = std::max(o, std::max(pu, pv));
        if (parallelogram) { const double puv = pu + double(b.v[i]) * b.v_length; lo[i] = std::min(lo[i], puv); hi[i] = std::max(hi[i], puv); }
        extra[i] = reach * kBarySlope;
    }
    return padded(lo, hi, extra);
}
// om_dyn_traced: the user's world box (Bounded::build_world_bounding_box), widened because the f32 quadric test - like Sphere::hit,
// whose arithmetic it shares - reports hits for rays that pass slightly outside the exact surface; the same fattening as a sphere
// of the object's smallest scale, applied to the box's half extents.  A box that is not finite sends the object to the always-test list.
bool dyn_traced_bounded(const om_dyn_traced& o) {
    for (int a = 0; a < 3; a++) if (!std::isfinite(o.bounds_lo[a]) || !std::isfinite(o.bounds_hi[a])) return false;
    return true;
}
Aabb dyn_traced_box(const om_dyn_traced& o, double reach) {
    const double rho = reach / col_norm_min(o.m_local_to_world);
    const double fat = std::sqrt(1.0 + kSphereK * rho * rho) - 1.0;
    double lo[3], hi[3], extra[3];
    for (int a = 0; a < 3; a++) {
        lo[a] = o.bounds_lo[a]; hi[a] = o.bounds_hi[a];
        extra[a] = (0.5 * (hi[a] - lo[a]) + 1e-3) * (fat + 0.01);
    }
    return padded(lo, hi, extra);
}
// Distance-proportional padding.  Each margin above that depends on `reach` is at most slope * D, D = distance from
// the ray origin to the object's centre (sphere: r*(sqrt(1+K*(D/r)^2)-1) <= sqrt(K)*D; cube and triangle: linear in D
// by construction), times the object's anisotropy (largest world half extent / smallest scale).  A traversal that
// widens every box by slope * (largest distance from the ray origin to a point of the box) therefore keeps every
// candidate a box built with the scene-wide `reach` would keep, and far fewer when the objects are small against it.
double aniso(const float* m) {
    double hmax = 0.0;
    for (int i = 0; i < 3; i++) hmax = std::max(hmax, std::sqrt(double(m[4 * i]) * m[4 * i] + double(m[4 * i + 1]) * m[4 * i + 1] + double(m[4 * i + 2]) * m[4 * i + 2]));
    return std::max(1.0, hmax / col_norm_min(m));
}
double half_area(const Aabb& b) {
    const double dx = double(b.hi[0]) - b.lo[0], dy = double(b.hi[1]) - b.lo[1], dz = double(b.hi[2]) - b.lo[2];
    return dx * dy + dy * dz + dz * dx;
}

// Scenes of 1e5 objects and more: the per-object loops of freeze() run on the host's threads (every iteration writes its own
// slots, so the images are byte-identical to the sequential ones).
template <class F>
void parallel_for(size_t n, F&& body) {
    const size_t hw = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    if (n < 65536 || hw < 2 || getenv("OM_FREEZE_SERIAL")) { for (size_t i = 0; i < n; i++) body(i); return; }
    std::vector<std::thread> th;
    const size_t chunk = (n + hw - 1) / hw;
    for (size_t t = 0; t < hw; t++) {
        const size_t a = t * chunk, b = std::min(n, a + chunk);
        if (a >= b) break;
        th.emplace_back([=, &body]() { for (size_t i = a; i < b; i++) body(i); });
    }
    for (auto& x : th) x.join();
}

struct Packer {
    std::vector<float>& v;
    explicit Packer(std::vector<float>& v_) : v(v_) {}
    uint32_t units() const { return uint32_t(v.size() / 4); }
    float* grow(size_t n_units) { const size_t at = v.size(); v.resize(at + 4 * n_units); return v.data() + at; }  // (valid until the next growth)
    void f4(float a, float b, float c, float d) { const float x[4] = {a, b, c, d}; v.insert(v.end(), x, x + 4); }
    void rows3(const float* m) { for (int r = 0; r < 3; r++) f4(m[4 * r], m[4 * r + 1], m[4 * r + 2], m[4 * r + 3]); }
    void u4(uint32_t a, uint32_t b, uint32_t c, uint32_t d) { float f[4]; uint32_t u[4] = {a, b, c, d}; memcpy(f, u, 16); f4(f[0], f[1], f[2], f[3]); }
};
float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

void pack_material(float* dst, const om_material& m, uint32_t pt, size_t index) {  // two units
    const uint64_t id = OM_OBJ_ID(pt + 1, index);  // ABI type tag = scan-order code + 1
    const uint64_t bloom = omh::bloom_get_hash(id);
    const float x[8] = {m.albedo[0], m.albedo[1], m.albedo[2], m.fuzz, m.ior, u2f(m.mat_type), u2f(uint32_t(bloom)), u2f(uint32_t(bloom >> 32))};
    memcpy(dst, x, sizeof x);
}

// ---- candidate grid over the marched objects (see SceneHeader::mgrid_*) -------------------------------------
// Bounds, over an axis-aligned cell, of the value the reference computes for |Marched::sdf| (marched.rs:14-18),
// in double with margins far above f32 rounding.
struct MBound { double lb, ub; };
double box_point_dist(const double lo[3], const double hi[3], const double c[3], bool farthest) {
    double s = 0;
    for (int a = 0; a < 3; a++) {
        double d;
        if (farthest) d = std::max(std::fabs(c[a] - lo[a]), std::fabs(c[a] - hi[a]));
        else d = c[a] < lo[a] ? lo[a] - c[a] : (c[a] > hi[a] ? c[a] - hi[a] : 0.0);
        s += d * d;
    }
    return std::sqrt(s);
}
MBound loosen(double lb, double ub, double scale) {
    MBound b;
    b.lb = std::max(0.0, lb * (1 - 1e-4) - 1e-5 * (1 + scale));
    b.ub = ub * (1 + 1e-4) + 1e-5 * (1 + scale);
    return b;
}
MBound msphere_bound(const om_marched_sphere& s, const double lo[3], const double hi[3]) {  // | |p - c| - r |
    const double c[3] = {s.center[0], s.center[1], s.center[2]}, r = s.radius;
    const double dmin = box_point_dist(lo, hi, c, false), dmax = box_point_dist(lo, hi, c, true);
    const double lb = (r >= dmin && r <= dmax) ? 0.0 : std::min(std::fabs(dmin - r), std::fabs(dmax - r));
    return loosen(lb, std::max(std::fabs(dmin - r), std::fabs(dmax - r)), std::fabs(r) + dmax);
}
MBound mbox_bound(const om_marched_box& s, const double lo[3], const double hi[3]) {  // marched.rs:86-89, convex in p
    auto sdf = [&](const double p[3]) {
        double q[3], o = 0, in = -1e300;
        for (int a = 0; a < 3; a++) { q[a] = std::fabs(p[a] - s.center[a]) - s.sizes[a]; o += std::max(q[a], 0.0) * std::max(q[a], 0.0); in = std::max(in, q[a]); }
        return std::sqrt(o) + std::min(in, 0.0);
    };
    double blo[3], bhi[3], gap = 0, mx = -1e300, deepest = 1e300, ext = 0;
    for (int a = 0; a < 3; a++) {
        blo[a] = s.center[a] - std::fabs(s.sizes[a]); bhi[a] = s.center[a] + std::fabs(s.sizes[a]);
        const double d = lo[a] > bhi[a] ? lo[a] - bhi[a] : (hi[a] < blo[a] ? blo[a] - hi[a] : 0.0);
        gap += d * d;
        deepest = std::min(deepest, std::fabs(double(s.sizes[a])));
        ext = std::max(ext, std::max(std::fabs(lo[a]), std::fabs(hi[a])));
    }
    for (int k = 0; k < 8; k++) { const double p[3] = {k & 1 ? hi[0] : lo[0], k & 2 ? hi[1] : lo[1], k & 4 ? hi[2] : lo[2]}; mx = std::max(mx, sdf(p)); }
    return loosen(std::sqrt(gap), std::max(mx, deepest), ext);  // cell touching the box: lb = 0; inside, |sdf| <= smallest half size
}
MBound mtorus_bound(const om_marched_torus& s, const double lo[3], const double hi[3]) {
    // q = W2L_TR * (p * w2l_scale) (marched.rs:142-144): affine; value = local_sdf(q) * min l2w scale (marched.rs:148-150)
    const float* m = s.m_world_to_local_translate_rotate;
    const float* si = s.m_world_to_local_scale;
    double pc[3], half = 0, smax = 0;
    for (int a = 0; a < 3; a++) { pc[a] = 0.5 * (lo[a] + hi[a]); half += 0.25 * (hi[a] - lo[a]) * (hi[a] - lo[a]); smax = std::max(smax, std::fabs(double(si[a]))); }
    // operator norm of the linear part <= |rotation rows| * max scale (rows of a TR matrix are unit, but be safe)
    double rown = 0;
    for (int r = 0; r < 3; r++) rown = std::max(rown, std::sqrt(double(m[4 * r]) * m[4 * r] + double(m[4 * r + 1]) * m[4 * r + 1] + double(m[4 * r + 2]) * m[4 * r + 2]));
    const double rho = std::sqrt(half) * smax * rown * 1.7320508;
    double q0[3], n0 = 0;
    for (int r = 0; r < 3; r++) {
        q0[r] = m[4 * r] * (pc[0] * si[0]) + m[4 * r + 1] * (pc[1] * si[1]) + m[4 * r + 2] * (pc[2] * si[2]) + m[4 * r + 3] * double(si[3]);
        n0 += q0[r] * q0[r];
    }
    n0 = std::sqrt(n0);
    const double ball = std::fabs(double(s.sizes[0])) + std::fabs(double(s.sizes[1]));
    const double mins = std::fabs(double(fminf(s.m_local_to_world_scale[0], fminf(s.m_local_to_world_scale[1], s.m_local_to_world_scale[2]))));
    return loosen(mins * std::max(0.0, n0 - rho - ball), mins * (n0 + rho + ball), mins * (n0 + rho + ball));
}

void build_marched_grid(const omh::SceneLists& L, omd::SceneHeader& h, Packer& cold) {
    const size_t ns = L.marched_spheres.size(), nb = L.marched_boxes.size(), nt = L.marched_tori.size(), n = ns + nb + nt;
    for (int a = 0; a < 3; a++) { h.mgrid_lo[a] = 0.f; h.mgrid_dim[a] = 0u; }
    h.mgrid_inv_h = 0.f; h.off_mgrid_cells = 0u; h.off_mgrid_lists = 0u;
    if (n < 6) return;  // a handful of objects: the plain loop is as fast
    // world extent (centre, radius) of each object; the grid covers the typical ones, not a scene-sized ground
    std::vector<double> cx(n), cy(n), cz(n), rad(n);
    for (size_t i = 0; i < ns; i++) { const auto& s = L.marched_spheres[i]; cx[i] = s.center[0]; cy[i] = s.center[1]; cz[i] = s.center[2]; rad[i] = std::fabs(double(s.radius)); }
    for (size_t i = 0; i < nb; i++) { const auto& s = L.marched_boxes[i]; const size_t j = ns + i; cx[j] = s.center[0]; cy[j] = s.center[1]; cz[j] = s.center[2]; rad[j] = std::sqrt(double(s.sizes[0]) * s.sizes[0] + double(s.sizes[1]) * s.sizes[1] + double(s.sizes[2]) * s.sizes[2]); }
    for (size_t i = 0; i < nt; i++) {
        const auto& s = L.marched_tori[i]; const size_t j = ns + nb + i;
        const float* m = s.m_local_to_world_translate_rotate; const float* sc = s.m_local_to_world_scale;
        cx[j] = double(m[3]) * sc[0]; cy[j] = double(m[7]) * sc[1]; cz[j] = double(m[11]) * sc[2];
        rad[j] = (std::fabs(double(s.sizes[0])) + std::fabs(double(s.sizes[1]))) * std::max(std::fabs(double(sc[0])), std::max(std::fabs(double(sc[1])), std::fabs(double(sc[2]))));
    }
    std::vector<double> sorted(rad);
    std::sort(sorted.begin(), sorted.end());
    const double med = std::max(sorted[n / 2], 1e-6);
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    size_t typical = 0;
    for (size_t j = 0; j < n; j++) {
        if (!(rad[j] <= 20 * med) || !std::isfinite(cx[j] + cy[j] + cz[j])) continue;
        typical++;
        const double c[3] = {cx[j], cy[j], cz[j]};
        for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], c[a] - rad[j]); hi[a] = std::max(hi[a], c[a] + rad[j]); }
    }
    if (typical < 4) return;
    // rays spend many march steps around the objects too (arriving from the camera, leaving towards the sky): the grid
    // covers the objects' box plus half its largest extent on every side
    double vol = 1, ext = 0;
    for (int a = 0; a < 3; a++) ext = std::max(ext, hi[a] - lo[a]);
    for (int a = 0; a < 3; a++) { lo[a] -= std::max(2 * med, 0.5 * ext); hi[a] += std::max(2 * med, 0.5 * ext); vol *= (hi[a] - lo[a]); }
    double edge = std::max(1.5 * med, std::cbrt(vol / 100000.0));
    uint32_t dim[3];
    for (int a = 0; a < 3; a++) dim[a] = uint32_t(std::min(1024.0, std::max(1.0, std::ceil((hi[a] - lo[a]) / edge))));
    const float flo[3] = {float(lo[0]), float(lo[1]), float(lo[2])};
    const float inv_h = float(1.0 / edge);
    const size_t ncell = size_t(dim[0]) * dim[1] * dim[2];
    std::vector<uint32_t> first(ncell), count(ncell), lists;
    std::vector<MBound> bnd(n);
    for (uint32_t iz = 0; iz < dim[2]; iz++) for (uint32_t iy = 0; iy < dim[1]; iy++) for (uint32_t ix = 0; ix < dim[0]; ix++) {
        // the device maps p to a cell with f32 arithmetic: (p - lo) * inv_h, truncated; make the cell's box a bit larger
        const uint32_t id[3] = {ix, iy, iz};
        double clo[3], chi[3];
        for (int a = 0; a < 3; a++) {
            const double e = 1.0 / double(inv_h), slack = 1e-4 * e + 1e-5 * (1 + std::fabs(double(flo[a])) + e * dim[a]);
            clo[a] = double(flo[a]) + id[a] * e - slack; chi[a] = double(flo[a]) + (id[a] + 1) * e + slack;
        }
        double u = 1e300;
        for (size_t j = 0; j < n; j++) {
            bnd[j] = j < ns ? msphere_bound(L.marched_spheres[j], clo, chi) : (j < ns + nb ? mbox_bound(L.marched_boxes[j - ns], clo, chi) : mtorus_bound(L.marched_tori[j - ns - nb], clo, chi));
            if (bnd[j].ub == bnd[j].ub) u = std::min(u, bnd[j].ub);
        }
        const size_t cell = (size_t(iz) * dim[1] + iy) * dim[0] + ix;
        first[cell] = uint32_t(lists.size());
        for (size_t j = 0; j < n; j++)
            if (!(bnd[j].lb > u)) lists.push_back(uint32_t(j));  // (a NaN bound keeps the object)
        count[cell] = uint32_t(lists.size()) - first[cell];
    }
    for (int a = 0; a < 3; a++) { h.mgrid_lo[a] = flo[a]; h.mgrid_dim[a] = dim[a]; }
    h.mgrid_inv_h = inv_h;
    h.off_mgrid_cells = cold.units();
    for (size_t c = 0; c < ncell; c += 2) cold.u4(first[c], count[c], c + 1 < ncell ? first[c + 1] : 0u, c + 1 < ncell ? count[c + 1] : 0u);
    h.off_mgrid_lists = cold.units();
    for (size_t k = 0; k < lists.size(); k += 4) cold.u4(lists[k], k + 1 < lists.size() ? lists[k + 1] : 0u, k + 2 < lists.size() ? lists[k + 2] : 0u, k + 3 < lists.size() ? lists[k + 3] : 0u);
}

// ---- distance BVH over the marched objects (see SceneHeader::mbvh_*) ----------------------------------------
// Every object gets a proxy box B_j and a factor f_j in (0,1] with |sdf_j(p)| >= f_j * dist(p, B_j) for every p:
//   sphere | |p-c| - r |          >= |p-c| - |r|                >= dist(p, c +- |r|)
//   box    (marched.rs:86-89)     exact distance outside, |.| >= 0 inside: B = c +- |sizes|
//   torus  local_sdf(q) * min scale (marched.rs:134-150): exact in local space, and the local->world map stretches
//          lengths by at most the largest scale, so f = min scale / max scale over a ball around the torus.
// The device walks the tree with the exact minimum found so far as the pruning bound (nearest_marched).
void build_marched_bvh(const omh::SceneLists& L, omd::SceneHeader& h, Packer& cold) {
    const size_t ns = L.marched_spheres.size(), nb = L.marched_boxes.size(), nt = L.marched_tori.size(), n = ns + nb + nt;
    if (n < 6) return;
    std::vector<Aabb> box(n);
    std::vector<double> f(n, 1.0);
    std::vector<bool> ok(n, true);
    auto set_box = [&](size_t j, const double c[3], const double r[3]) {
        for (int a = 0; a < 3; a++) {
            const double m = 1e-5 * (1 + std::fabs(c[a]) + r[a]);  // proxies err on the large side
            box[j].lo[a] = nextafterf(float(c[a] - r[a] - m), -kInf); box[j].hi[a] = nextafterf(float(c[a] + r[a] + m), kInf);
            if (!std::isfinite(box[j].lo[a]) || !std::isfinite(box[j].hi[a])) ok[j] = false;
        }
    };
    for (size_t i = 0; i < ns; i++) {
        const auto& s = L.marched_spheres[i];
        const double c[3] = {s.center[0], s.center[1], s.center[2]}, r[3] = {std::fabs(double(s.radius)), std::fabs(double(s.radius)), std::fabs(double(s.radius))};
        set_box(i, c, r);
    }
    for (size_t i = 0; i < nb; i++) {
        const auto& s = L.marched_boxes[i];
        const double c[3] = {s.center[0], s.center[1], s.center[2]}, r[3] = {std::fabs(double(s.sizes[0])), std::fabs(double(s.sizes[1])), std::fabs(double(s.sizes[2]))};
        set_box(ns + i, c, r);
    }
    for (size_t i = 0; i < nt; i++) {
        const auto& s = L.marched_tori[i];
        const size_t j = ns + nb + i;
        const float* m = s.m_local_to_world_translate_rotate; const float* sc = s.m_local_to_world_scale;
        double smin = 1e300, smax = 0;
        for (int a = 0; a < 3; a++) { smin = std::min(smin, std::fabs(double(sc[a]))); smax = std::max(smax, std::fabs(double(sc[a]))); }
        // rows of the rotation part should be unit; a sloppy matrix only makes the ball larger
        double rown = 1.0;
        for (int r = 0; r < 3; r++) rown = std::max(rown, std::sqrt(double(m[4 * r]) * m[4 * r] + double(m[4 * r + 1]) * m[4 * r + 1] + double(m[4 * r + 2]) * m[4 * r + 2]));
        const double c[3] = {double(m[3]) * sc[0], double(m[7]) * sc[1], double(m[11]) * sc[2]};
        const double rad = (std::fabs(double(s.sizes[0])) + std::fabs(double(s.sizes[1]))) * smax * rown * 1.7320508;
        const double r[3] = {rad, rad, rad};
        set_box(j, c, r);
        // value = local_sdf * min_val(l2w scale) (signed min, marched.rs:149); a non-positive or wildly uneven scale is not bounded here
        const double mv = std::min(double(sc[0]), std::min(double(sc[1]), double(sc[2])));
        f[j] = (mv > 0 && smax > 0) ? mv / (smax * rown) : 0.0;
        if (!(f[j] > 1e-3) || !std::isfinite(rad)) ok[j] = false;
    }
    // scene-sized or unbounded objects are evaluated at every query (like the traced always-test list)
    Aabb all;
    for (int a = 0; a < 3; a++) { all.lo[a] = kInf; all.hi[a] = -kInf; }
    size_t nok = 0;
    for (size_t j = 0; j < n; j++) if (ok[j]) { nok++; for (int a = 0; a < 3; a++) { all.lo[a] = std::min(all.lo[a], box[j].lo[a]); all.hi[a] = std::max(all.hi[a], box[j].hi[a]); } }
    if (nok < 4) return;
    const double total = half_area(all);
    std::vector<uint32_t> always, refs;
    std::vector<Aabb> boxes;
    double fmin = 1.0, ext = 0.0;
    for (size_t j = 0; j < n; j++) {
        if (!ok[j] || (half_area(box[j]) >= 0.25 * total && always.size() < 16)) { always.push_back(uint32_t(j)); continue; }
        refs.push_back(uint32_t(j)); boxes.push_back(box[j]);
        fmin = std::min(fmin, f[j]);
        for (int a = 0; a < 3; a++) ext = std::max(ext, std::max(std::fabs(double(box[j].lo[a])), std::fabs(double(box[j].hi[a]))));
    }
    if (refs.size() < 4 || always.size() > 16) return;
    omb::BuildResult bvh;
    omb::build(refs, boxes, bvh);
    if (bvh.nodes.empty() || bvh.depth > 40) return;
    h.mbvh_f = float(fmin * (1 - 1e-4));
    h.mbvh_margin = float(1e-5 * (1 + ext));
    h.n_mbvh_nodes = uint32_t(bvh.nodes.size());
    h.n_mbvh_always = uint32_t(always.size());
    h.off_mbvh_always = cold.units();
    for (size_t k = 0; k < always.size(); k += 4) cold.u4(always[k], k + 1 < always.size() ? always[k + 1] : 0u, k + 2 < always.size() ? always[k + 2] : 0u, k + 3 < always.size() ? always[k + 3] : 0u);
    h.off_mbvh_nodes = cold.units();
    for (const omb::Node& nd : bvh.nodes) {
        cold.f4(nd.lmin[0], nd.lmin[1], nd.lmin[2], nd.lmax[0]);
        cold.f4(nd.lmax[1], nd.lmax[2], nd.rmin[0], nd.rmin[1]);
        cold.f4(nd.rmin[2], nd.rmax[0], nd.rmax[1], nd.rmax[2]);
        cold.u4(uint32_t(nd.left), uint32_t(nd.right), 0u, 0u);
    }
}
}  // namespace

std::string freeze(const omh::SceneLists& L, const om_camera* cam, uint32_t flags, FrozenScene& out) {
    using namespace omd;
    SceneHeader& h = out.hdr;
    memset(&h, 0, sizeof h);
    out.hot.clear();
    out.cold.clear();
    // OM_FREEZE_TIMING=1 (development): seconds spent per phase, to stderr
    const bool timing = getenv("OM_FREEZE_TIMING") != nullptr;
    auto t_prev = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!timing) return;
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[om freeze] %-22s %.3f s\n", what, std::chrono::duration<double>(now - t_prev).count());
        t_prev = now;
    };
    h.n[PT_SPHERE] = uint32_t(L.spheres.size());   h.n[PT_CUBE] = uint32_t(L.cubes.size());
    h.n[PT_TRI] = uint32_t(L.triangles.size());    h.n[PT_PLANE] = uint32_t(L.infinite_planes.size());
    h.n[PT_PARA] = uint32_t(L.parallelograms.size());
    h.n[PT_MSPHERE] = uint32_t(L.marched_spheres.size()); h.n[PT_MBOX] = uint32_t(L.marched_boxes.size());
    h.n[PT_MTORUS] = uint32_t(L.marched_tori.size());
    h.n[PT_MDYN] = uint32_t(L.dyn_marched.size());
    h.n[PT_TDYN] = uint32_t(L.dyn_traced.size());
    out.n_total = 0;
    for (uint32_t i = 0; i < PT_COUNT; i++) out.n_total += h.n[i];

    // ---- bounded traced primitives -> BVH; unbounded / scene-sized ones -> always-test list ----
    // `reach`: how far a ray origin can be from an object it still has to find.  Origins are the
    // camera or points on other objects, so the diagonal of everything scene-sized objects aside,
    // and at least the reference's t_max of 100 (main.rs:206).
    double reach = 100.0;
    {
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        auto grow_pt = [&](double x, double y, double z, double r) {
            const double p[3] = {x, y, z};
            for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], p[a] - r); hi[a] = std::max(hi[a], p[a] + r); }
        };
        std::vector<double> radii(L.spheres.size());
        parallel_for(L.spheres.size(), [&](size_t i) { radii[i] = col_norm_min(L.spheres[i].m_local_to_world); });
        double med = 0.0;
        if (!radii.empty()) { std::vector<double> t = radii; std::nth_element(t.begin(), t.begin() + t.size() / 2, t.end()); med = t[t.size() / 2]; }
        for (size_t i = 0; i < L.spheres.size(); i++)
            if (L.spheres.size() < 4 || radii[i] <= 64.0 * med) grow_pt(L.spheres[i].m_local_to_world[3], L.spheres[i].m_local_to_world[7], L.spheres[i].m_local_to_world[11], radii[i]);
        for (const om_cube& o : L.cubes) grow_pt(o.m_local_to_world[3], o.m_local_to_world[7], o.m_local_to_world[11], 1.0);
        for (const om_barycentric& o : L.triangles) grow_pt(o.origin[0], o.origin[1], o.origin[2], std::max(o.u_length, o.v_length));
        for (const om_barycentric& o : L.parallelograms) grow_pt(o.origin[0], o.origin[1], o.origin[2], double(o.u_length) + o.v_length);
        if (cam) grow_pt(cam->origin[0], cam->origin[1], cam->origin[2], 0.0);
        if (lo[0] <= hi[0]) {
            const double dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
            const double diag = std::sqrt(dx * dx + dy * dy + dz * dz);
            if (std::isfinite(diag)) reach = std::max(reach, diag);
        }
    }
    // Static padding is fine while it stays a small share of the objects' own size; when the typical sphere's box
    // would at least double (tiny spheres in a huge scene), switch to boxes without the distance-dependent part and
    // let the traversal add it per ray (SceneHeader::dyn_pad).
    bool dyn = (flags & 2u) != 0;
    if (!(flags & 6u) && L.spheres.size() >= 64) {
        std::vector<double> fat(L.spheres.size());
        parallel_for(L.spheres.size(), [&](size_t i) { fat[i] = sphere_fat(L.spheres[i], reach); });
        std::nth_element(fat.begin(), fat.begin() + fat.size() / 2, fat.end());
        dyn = fat[fat.size() / 2] > 0.5;
    }
    if (dyn) {
        double slope = 0.0;
        for (const om_sphere& o : L.spheres) slope = std::max(slope, std::sqrt(kSphereK) * aniso(o.m_local_to_world));
        for (const om_cube& o : L.cubes) slope = std::max(slope, 2.0 * kCubeSlope * aniso(o.m_local_to_world));
        if (!L.triangles.empty() || !L.parallelograms.empty()) slope = std::max(slope, kBarySlope);
        for (const om_dyn_traced& o : L.dyn_traced) slope = std::max(slope, std::sqrt(kSphereK) * aniso(o.m_local_to_world));
        h.dyn_pad = float(slope * 1.01);
        if (!(h.dyn_pad > 0.f) || !std::isfinite(h.dyn_pad)) { dyn = false; h.dyn_pad = 0.f; }
    }
    if (dyn) reach = 0.0;
    std::vector<uint32_t> refs;
    std::vector<Aabb> boxes;
    refs.reserve(L.spheres.size() + L.cubes.size() + L.triangles.size() + L.parallelograms.size() + L.dyn_traced.size());
    boxes.reserve(refs.capacity());
    refs.resize(L.spheres.size());
    boxes.resize(L.spheres.size());
    parallel_for(L.spheres.size(), [&](size_t i) { refs[i] = make_ref(PT_SPHERE, uint32_t(i)); boxes[i] = sphere_box(L.spheres[i], reach); });
    for (size_t i = 0; i < L.cubes.size(); i++) { refs.push_back(make_ref(PT_CUBE, uint32_t(i))); boxes.push_back(cube_box(L.cubes[i], reach)); }
    for (size_t i = 0; i < L.triangles.size(); i++) { refs.push_back(make_ref(PT_TRI, uint32_t(i))); boxes.push_back(bary_box(L.triangles[i], false, reach)); }
    for (size_t i = 0; i < L.parallelograms.size(); i++) { refs.push_back(make_ref(PT_PARA, uint32_t(i))); boxes.push_back(bary_box(L.parallelograms[i], true, reach)); }
    for (size_t i = 0; i < L.dyn_traced.size(); i++)
        if (dyn_traced_bounded(L.dyn_traced[i])) { refs.push_back(make_ref(PT_TDYN, uint32_t(i))); boxes.push_back(dyn_traced_box(L.dyn_traced[i], reach)); }
    std::vector<uint32_t> large;
    for (size_t i = 0; i < L.infinite_planes.size(); i++) large.push_back(make_ref(PT_PLANE, uint32_t(i)));  // Bounded default: +-INF (bounding_box.rs:69-73)
    for (size_t i = 0; i < L.dyn_traced.size(); i++) if (!dyn_traced_bounded(L.dyn_traced[i])) large.push_back(make_ref(PT_TDYN, uint32_t(i)));
    if (!(flags & 1u) && refs.size() > 4) {
        // An object whose box is a large share of the whole scene (the r=1000 ground sphere,
        // main.rs:41) overlaps every subtree: test it once per ray instead.
        Aabb all;
        for (int a = 0; a < 3; a++) { all.lo[a] = kInf; all.hi[a] = -kInf; }
        for (const Aabb& b : boxes) for (int a = 0; a < 3; a++) { all.lo[a] = std::min(all.lo[a], b.lo[a]); all.hi[a] = std::max(all.hi[a], b.hi[a]); }
        const double total = half_area(all);
        std::vector<uint32_t> r2; std::vector<Aabb> b2;
        for (size_t i = 0; i < refs.size(); i++) {
            const bool nonfinite = !std::isfinite(half_area(boxes[i]));
            if ((nonfinite || half_area(boxes[i]) >= 0.25 * total) && large.size() < 64) large.push_back(refs[i]);
            else { r2.push_back(refs[i]); b2.push_back(boxes[i]); }
        }
        refs.swap(r2); boxes.swap(b2);
    }
    std::sort(large.begin(), large.end());
    lap("boxes + always list");
    omb::BuildResult bvh;
    bool built = false;
    if ((flags & 16u) && refs.size() >= 2) {
        char why[256] = "";
        built = omb::build_gpu(refs, boxes, bvh, 60, why, sizeof why);
        // a tree too deep for the traversal stacks or a non-finite box falls to the host builder; a CUDA failure is an error
        if (!built && strstr(why, ": ")) return std::string(why);
    }
    if (!built) omb::build(refs, boxes, bvh);
    // the traversal walks the tree four children at a time (half the visits of the binary tree, same primitives tested)
    omb::Wide4 wide;
    omb::collapse4(bvh, wide, 4);
    if (wide.stack_need > 60u) omb::collapse4(bvh, wide, 2);  // a degenerate tree: one sibling per level, as the binary traversal (depth <= 62)
    h.n_nodes = uint32_t(wide.nodes.size());
    h.n_large = uint32_t(large.size());
    h.n_bvh_prims = uint32_t(refs.size());
    h.bvh_depth = wide.depth;
    h.stack_need = wide.stack_need;

    lap(built ? "bvh (device)" : "bvh (host)");
    if (timing) fprintf(stderr, "[freeze] binary nodes %zu depth %u -> wide nodes %zu depth %u stack_need %u\n", bvh.nodes.size(), bvh.depth, wide.nodes.size(), wide.depth, wide.stack_need);
    // ---- hot image ----------------------------------------------------------------------------
    out.hot.reserve(28 * wide.nodes.size() + 12 * (L.spheres.size() + L.cubes.size()) + 16 * (L.triangles.size() + L.parallelograms.size()) + 4096);
    Packer hot(out.hot);
    h.off_nodes = hot.units();
    {
        static_assert(sizeof(omb::Node4) == 16 * U_NODE4, "a wide node is packed as it lies in memory: six plane quads, then the child references");
        float* dst = hot.grow(size_t(U_NODE4) * wide.nodes.size());
        parallel_for(wide.nodes.size(), [&](size_t i) { memcpy(dst + 4 * U_NODE4 * i, &wide.nodes[i], sizeof(omb::Node4)); });
    }
    h.off_node_far = hot.units();
    if (h.dyn_pad != 0.f) {
        float* dst = hot.grow(wide.nodes.size());
        parallel_for(wide.nodes.size(), [&](size_t i) {
            const omb::Node4& nd = wide.nodes[i];
            float* o4 = dst + 4 * i;
            double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
            for (int j = 0; j < 4; j++) {
                if (!(nd.lox[j] <= nd.hix[j])) continue;  // empty slot
                lo[0] = std::min(lo[0], double(nd.lox[j])); hi[0] = std::max(hi[0], double(nd.hix[j]));
                lo[1] = std::min(lo[1], double(nd.loy[j])); hi[1] = std::max(hi[1], double(nd.hiy[j]));
                lo[2] = std::min(lo[2], double(nd.loz[j])); hi[2] = std::max(hi[2], double(nd.hiz[j]));
            }
            if (!(lo[0] <= hi[0])) { o4[0] = o4[1] = o4[2] = o4[3] = 0.f; return; }
            const float c[3] = {float(0.5 * (lo[0] + hi[0])), float(0.5 * (lo[1] + hi[1])), float(0.5 * (lo[2] + hi[2]))};
            double r2 = 0;  // from the ROUNDED centre to the farthest corner
            for (int a = 0; a < 3; a++) { const double d = std::max(std::fabs(lo[a] - double(c[a])), std::fabs(hi[a] - double(c[a]))); r2 += d * d; }
            o4[0] = c[0]; o4[1] = c[1]; o4[2] = c[2];
            o4[3] = std::nextafter(float(std::sqrt(r2) * (1.0 + 1e-6)), std::numeric_limits<float>::infinity());
        });
    }
    h.off_hot[PT_SPHERE] = hot.units();
    {
        float* dst = hot.grow(3 * L.spheres.size());
        parallel_for(L.spheres.size(), [&](size_t i) { memcpy(dst + 12 * i, L.spheres[i].m_world_to_local, 48); });  // rows 0-2
    }
    h.off_hot[PT_CUBE] = hot.units();
    for (const om_cube& s : L.cubes) hot.rows3(s.m_world_to_local);
    auto pack_bary = [&](const om_barycentric& b) {
        // only rows 0 and 2 of base_inv and the x/z of v_in_base are read (traced.rs:155-167)
        hot.f4(b.origin[0], b.origin[1], b.origin[2], b.u_length);
        hot.f4(b.uxv[0], b.uxv[1], b.uxv[2], b.v_length);
        hot.f4(b.base_inv[0], b.base_inv[1], b.base_inv[2], b.v_in_base[0]);
        hot.f4(b.base_inv[6], b.base_inv[7], b.base_inv[8], b.v_in_base[2]);
    };
    h.off_hot[PT_TRI] = hot.units();
    for (const om_barycentric& b : L.triangles) pack_bary(b);
    h.off_hot[PT_PLANE] = hot.units();
    for (const om_infinite_plane& p : L.infinite_planes) { hot.f4(p.normal[0], p.normal[1], p.normal[2], 0.f); hot.f4(p.center[0], p.center[1], p.center[2], 0.f); }
    h.off_hot[PT_PARA] = hot.units();
    for (const om_barycentric& b : L.parallelograms) pack_bary(b);
    h.off_hot[PT_TDYN] = hot.units();
    for (const om_dyn_traced& o : L.dyn_traced) {
        hot.rows3(o.m_world_to_local);
        hot.f4(o.q[0], o.q[1], o.q[2], o.q[3]);
        hot.f4(o.q[4], o.q[5], o.q[6], o.q[7]);
        hot.f4(o.q[8], o.q[9], o.clip_lo[0], o.clip_lo[1]);
        hot.f4(o.clip_lo[2], o.clip_hi[0], o.clip_hi[1], o.clip_hi[2]);
    }
    h.off_large = hot.units();
    for (size_t i = 0; i < large.size(); i += 4) {
        uint32_t q[4] = {REF_NONE, REF_NONE, REF_NONE, REF_NONE};
        for (size_t k = 0; k < 4 && i + k < large.size(); k++) q[k] = large[i + k];
        hot.u4(q[0], q[1], q[2], q[3]);
    }
    // marched objects and their bounds
    double mlo[3] = {1e300, 1e300, 1e300}, mhi[3] = {-1e300, -1e300, -1e300};
    auto grow = [&](const double lo[3], const double hi[3]) { for (int a = 0; a < 3; a++) { mlo[a] = std::min(mlo[a], lo[a]); mhi[a] = std::max(mhi[a], hi[a]); } };
    const double HIT = 0.001;  // hits.rs:113
    h.off_hot[PT_MSPHERE] = hot.units();
    for (const om_marched_sphere& s : L.marched_spheres) {
        hot.f4(s.center[0], s.center[1], s.center[2], s.radius);
        double lo[3], hi[3];
        const double r = std::fabs(double(s.radius)) + 4 * HIT;
        for (int a = 0; a < 3; a++) { lo[a] = s.center[a] - r * 1.001; hi[a] = s.center[a] + r * 1.001; }
        grow(lo, hi);
    }
    h.off_hot[PT_MBOX] = hot.units();
    for (const om_marched_box& s : L.marched_boxes) {
        hot.f4(s.center[0], s.center[1], s.center[2], 0.f);
        hot.f4(s.sizes[0], s.sizes[1], s.sizes[2], 0.f);
        double lo[3], hi[3];
        for (int a = 0; a < 3; a++) { const double r = (std::fabs(double(s.sizes[a])) + 4 * HIT) * 1.001; lo[a] = s.center[a] - r; hi[a] = s.center[a] + r; }
        grow(lo, hi);
    }
    h.off_hot[PT_MTORUS] = hot.units();
    for (const om_marched_torus& s : L.marched_tori) {
        hot.rows3(s.m_world_to_local_translate_rotate);
        hot.rows3(s.m_local_to_world_translate_rotate);
        hot.f4(s.m_world_to_local_scale[0], s.m_world_to_local_scale[1], s.m_world_to_local_scale[2], s.m_world_to_local_scale[3]);
        hot.f4(s.m_local_to_world_scale[0], s.m_local_to_world_scale[1], s.m_local_to_world_scale[2], s.m_local_to_world_scale[3]);
        const float mins = fminf(s.m_local_to_world_scale[0], fminf(s.m_local_to_world_scale[1], s.m_local_to_world_scale[2]));  // marched.rs:148-150
        hot.f4(s.sizes[0], s.sizes[1], mins, 0.f);
        // |sdf| < HIT  <=>  local distance < HIT / min_scale; local torus inside a ball of radius R + r
        // around the TR translation; world point = (TR q) * scale componentwise (marched.rs:142-147)
        const double maxs = std::max(std::fabs(double(s.m_local_to_world_scale[0])), std::max(std::fabs(double(s.m_local_to_world_scale[1])), std::fabs(double(s.m_local_to_world_scale[2]))));
        const double rad = (std::fabs(double(s.sizes[0])) + std::fabs(double(s.sizes[1])) + 4 * HIT / std::max(1e-12, double(std::fabs(mins)))) * 1.001;
        double lo[3], hi[3];
        for (int a = 0; a < 3; a++) {
            const double c = s.m_local_to_world_translate_rotate[4 * a + 3], sc = s.m_local_to_world_scale[a];
            const double p0 = (c - rad) * sc, p1 = (c + rad) * sc;
            lo[a] = std::min(p0, p1) - 4 * HIT * maxs; hi[a] = std::max(p0, p1) + 4 * HIT * maxs;
        }
        grow(lo, hi);
    }
    // Arc<dyn Marched> as data: transforms + where the object's local_sdf program sits in the code pool
    h.off_hot[PT_MDYN] = hot.units();
    for (size_t i = 0; i < L.dyn_marched.size(); i++) {
        const om_dyn_marched& s = L.dyn_marched[i];
        hot.rows3(s.m_world_to_local);
        hot.rows3(s.m_local_to_world);
        hot.f4(s.world_f, u2f(L.dyn_code_first[i]), u2f(L.dyn_code_first[i + 1] - L.dyn_code_first[i]), 0.f);
        double lo[3], hi[3];
        for (int a = 0; a < 3; a++) {  // the user's box (or +-INF = draw always), padded like the others
            const double pad = 4 * HIT + 1e-3 * std::max(std::fabs(double(s.bounds_lo[a])), std::fabs(double(s.bounds_hi[a])));
            lo[a] = std::isfinite(s.bounds_lo[a]) ? s.bounds_lo[a] - pad : -1e300;
            hi[a] = std::isfinite(s.bounds_hi[a]) ? s.bounds_hi[a] + pad : 1e300;
        }
        grow(lo, hi);
    }
    h.off_dyn_code = hot.units();
    for (size_t i = 0; i < L.dyn_code.size(); i += 4) {
        uint32_t w[4] = {0, 0, 0, 0};
        for (size_t k = 0; k < 4 && i + k < L.dyn_code.size(); k++) w[k] = L.dyn_code[i + k];
        hot.u4(w[0], w[1], w[2], w[3]);
    }
    h.off_sph_mat = hot.units();
    for (size_t i = 0; i < L.spheres.size(); i += 16) {
        uint32_t w[4] = {0, 0, 0, 0};
        for (size_t k = 0; k < 16 && i + k < L.spheres.size(); k++) w[k >> 2] |= (L.spheres[i + k].material.mat_type & 0xFFu) << (8 * (k & 3));
        hot.u4(w[0], w[1], w[2], w[3]);
    }
    h.n_marched = h.n[PT_MSPHERE] + h.n[PT_MBOX] + h.n[PT_MTORUS] + h.n[PT_MDYN];
    for (int a = 0; a < 3; a++) {
        if (h.n_marched == 0) { h.march_lo[a] = kInf; h.march_hi[a] = -kInf; }
        else {  // (a dyn marched object without bounds makes the union infinite: every ray marches)
            h.march_lo[a] = mlo[a] < -3e38 ? -kInf : nextafterf(float(mlo[a]), -kInf);
            h.march_hi[a] = mhi[a] > 3e38 ? kInf : nextafterf(float(mhi[a]), kInf);
        }
    }
    h.hot_units = hot.units();

    // ---- cold image ---------------------------------------------------------------------------
    out.cold.reserve(12 * (L.spheres.size() + L.cubes.size()) + 8 * out.n_total + 4096);
    Packer cold(out.cold);
    h.off_sph_l2w = cold.units();
    {
        float* dst = cold.grow(3 * L.spheres.size());
        parallel_for(L.spheres.size(), [&](size_t i) { memcpy(dst + 12 * i, L.spheres[i].m_local_to_world, 48); });
    }
    h.off_cube_l2w = cold.units();
    for (const om_cube& s : L.cubes) cold.rows3(s.m_local_to_world);
    h.off_tdyn_l2w = cold.units();
    for (const om_dyn_traced& o : L.dyn_traced) cold.rows3(o.m_local_to_world);
    h.off_mat = cold.units();
    uint32_t base = 0;
    auto mats = [&](uint32_t pt, auto const& list) {
        h.mat_base[pt] = base;
        float* dst = cold.grow(2 * list.size());
        parallel_for(list.size(), [&](size_t i) { pack_material(dst + 8 * i, list[i].material, pt, i); });
        base += uint32_t(list.size());
    };
    mats(PT_SPHERE, L.spheres); mats(PT_CUBE, L.cubes); mats(PT_TRI, L.triangles); mats(PT_PLANE, L.infinite_planes);
    mats(PT_PARA, L.parallelograms); mats(PT_MSPHERE, L.marched_spheres); mats(PT_MBOX, L.marched_boxes); mats(PT_MTORUS, L.marched_tori);
    mats(PT_MDYN, L.dyn_marched);
    mats(PT_TDYN, L.dyn_traced);
    lap("pack hot + cold images");
    // The candidate grid and the distance BVH need analytic bounds of every |sdf|; a user program has none, so a scene
    // with dyn marched objects is marched by the plain scan (hits.rs:301-320).
    if (L.dyn_marched.empty()) {
        build_marched_grid(L, h, cold);
        if (!(flags & 8u)) build_marched_bvh(L, h, cold);
    }
    if (cold.units() == 0) cold.f4(0, 0, 0, 0);
    if (hot.units() == 0) hot.f4(0, 0, 0, 0);
    h.hot_units = hot.units();
    h.cold_units = cold.units();
    lap("marched structures");
    if (timing) {  // the images' checksum: a refactoring of this file must not move it
        auto fnv = [](const std::vector<float>& v) { uint64_t x = 1469598103934665603ull; const unsigned char* b = reinterpret_cast<const unsigned char*>(v.data()); for (size_t i = 0; i < v.size() * 4; i++) { x ^= b[i]; x *= 1099511628211ull; } return x; };
        fprintf(stderr, "[om freeze] hot %zu floats fnv %016llx, cold %zu floats fnv %016llx\n", out.hot.size(), (unsigned long long)fnv(out.hot), out.cold.size(), (unsigned long long)fnv(out.cold));
    }
    return std::string();
}

}  // namespace oms
