"""A second, independent pin of the oracle (CPU only).

The reference has no tests and cannot be built here, so `oracle/` (C++) could share a transcription error with the CUDA
code written beside it.  This file restates the closed forms of the hot path a second time, in numpy, DIRECTLY from the
Rust sources cited at each function (/root/reference/src/...), without importing, reading or following anything under
oracle/ except its C entry points as the thing under test.  Two statements per function where it pays:
  * float64: the mathematics.  Decisions (hit / miss, which root, which face, inside / outside) must agree exactly away
    from a stated guard band around each threshold; t, point, normal within 1e-5 relative.
  * float32, one numpy operation per Rust operation, in the Rust's order: must agree BIT FOR BIT (numpy rounds every
    elementwise f32 operation separately, like rustc without fast-math; sqrt and division are correctly rounded on both).
Integer functions (BloomFilter, scramble, tone map to u8) are restated with Python ints and must agree exactly.
Inputs are random and well-conditioned (>= 10^5 per function); the objects' matrices come from the constructors as data."""
import ctypes as C

import numpy as np
import pytest

F = np.float32
TMIN, TMAX = 0.001, 100.0
N_OBJ, N_RAYS = 64, 2000   # 128 000 rays per primitive type


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def rel_err(got, want, floor=1e-3):
    return np.abs(got.astype(np.float64) - want) / np.maximum(np.abs(want), floor)


# ---- vec3.rs:29-40,236-240 / ray.rs:11-16 / mat4x4.rs:49-57 / vec4.rs:27-29, one operation at a time ---------------------
def dot3(a, b):
    return a[..., 0] * b[..., 0] + a[..., 1] * b[..., 1] + a[..., 2] * b[..., 2]


def unit(v):  # `*self / self.length()`, and Div<f32> multiplies by the reciprocal
    one = v.dtype.type(1.0)
    return v * (one / np.sqrt(dot3(v, v)))[..., None]


def ray_at(o, d, t):  # self.orig + t*self.dir
    return o + t[..., None] * d


def dot_p3(m, p):  # rows . (x, y, z, 1)
    one = p.dtype.type(1.0)
    return np.stack([m[i, 0] * p[..., 0] + m[i, 1] * p[..., 1] + m[i, 2] * p[..., 2] + m[i, 3] * one for i in range(3)], axis=-1)


def dot_v3(m, v):  # rows . (x, y, z, 0)
    zero = v.dtype.type(0.0)
    return np.stack([m[i, 0] * v[..., 0] + m[i, 1] * v[..., 1] + m[i, 2] * v[..., 2] + m[i, 3] * zero for i in range(3)], axis=-1)


def rays_towards(rng, centre, radius, n, inside_frac=0.15):
    """origins 2..15 radii away (some inside), aimed near the object; dirs not normalised (Ray::new does it)"""
    o = centre + rng.normal(size=(n, 3)) * radius * rng.uniform(2.0, 15.0, size=(n, 1)) / np.sqrt(3.0)
    ins = rng.random(n) < inside_frac
    o[ins] = centre + rng.uniform(-0.3, 0.3, size=(int(ins.sum()), 3)) * radius
    tgt = centre + rng.uniform(-1.4, 1.4, size=(n, 3)) * radius
    d = tgt - o
    d[np.linalg.norm(d, axis=1) < 1e-3] = (0.3, 0.5, -0.8)
    return o.astype(F), d.astype(F)


def random_trs(api, rng, scale_lo=0.3, scale_hi=2.5):
    t, r, s = rng.uniform(-6, 6, 3), rng.uniform(0, 6.28, 3), rng.uniform(scale_lo, scale_hi, 3)
    m = (api.Mat4x4.new_translate(t) ^ api.Mat4x4.new_rotate_x(r[0]) ^ api.Mat4x4.new_rotate_y(r[1]) ^ api.Mat4x4.new_rotate_z(r[2])
         ^ api.Mat4x4.new_scale(s))
    return m, t, float(s.max())


def mat(a16, dtype):
    return np.array(a16[:], dtype=np.float32).reshape(4, 4).astype(dtype)


# ---- Sphere::hit, traced.rs:38-63 -----------------------------------------------------------------------------------------
def sphere_hit(w2l, l2w, o, d_raw, dtype):
    o, d = o.astype(dtype), unit(d_raw.astype(dtype))
    lo, ld = dot_p3(w2l, o), dot_v3(w2l, d)
    a = dot3(ld, ld)
    half_b = dot3(lo, ld)
    c = dot3(lo, lo) - dtype(1.0)
    disc = half_b * half_b - a * c
    with np.errstate(invalid="ignore"):
        sq = np.sqrt(disc)
        r1 = (-half_b - sq) / a
        r2 = (-half_b + sq) / a
    tmin, tmax = dtype(TMIN), dtype(TMAX)
    bad1 = (r1 < tmin) | (r1 > tmax)
    root = np.where(bad1, r2, r1)
    hit = (disc >= 0) & ~(bad1 & ((r2 < tmin) | (r2 > tmax)))
    lp = ray_at(lo, ld, root)
    point = dot_p3(l2w, lp)
    normal = unit(dot_v3(l2w, lp))
    # how close each decision is to flipping (relative), for the float64 guard band
    margin = np.fmin.reduce([np.abs(disc) / np.maximum(half_b * half_b, 1e-30), np.abs(r1 - tmin), np.abs(r2 - tmin),
                                np.abs(r1 - tmax) / tmax, np.abs(r2 - tmax) / tmax]) if dtype is np.float64 else None
    # conditioning of the accepted root in f32: disc = half_b^2 - a*c cancels for grazing rays and far origins
    # ... and -half_b - sqrt(disc) cancels once the origin is many radii away in the local frame
    # ... and `c = |o_l|^2 - 1` cancels for an origin on the surface (every scattered ray starts there: the GPU tests cover those bit for bit)
    cond = np.where((np.sqrt(dot3(lo, lo)) <= 25.0) & (np.abs(c) > 0.05 * dot3(lo, lo)), np.abs(disc) / np.maximum(half_b * half_b, 1e-30), 0.0) if dtype is np.float64 else None
    if dtype is np.float64:  # t is a difference of these two terms: its f32 error is relative to THEM, not to t
        with np.errstate(invalid="ignore"):
            cond = (cond, np.maximum(np.abs(root), (np.abs(half_b) + sq) / a))
    return hit, root, point, normal, margin, cond


def test_sphere_hit(api, O, A):
    rng = np.random.default_rng(101)
    matl = api.Material.new_lambertian((0.5, 0.5, 0.5))
    n_checked = n_tol = 0
    for _ in range(N_OBJ):
        m, centre, smax = random_trs(api, rng)
        s = api.Sphere.new(m, matl)
        osc = O.OracleScene().add("spheres", (A.Sphere * 1)(s))
        o, d = rays_towards(rng, centre, smax, N_RAYS)
        ids, t, pts, nrm = osc.trace_rays(o, d, TMIN, TMAX)
        got_hit = ids != 0
        # float32, operation by operation: bit-exact
        h32, t32, p32, n32, _, _ = sphere_hit(mat(s.m_world_to_local, F), mat(s.m_local_to_world, F), o, d, F)
        assert (h32 == got_hit).all()
        assert (bits(t32[h32]) == bits(t[h32])).all() and (bits(p32[h32]) == bits(pts[h32])).all() and (bits(n32[h32]) == bits(nrm[h32])).all()
        # float64: the mathematics
        h64, t64, p64, n64, margin, (cond, t_scale) = sphere_hit(mat(s.m_world_to_local, np.float64), mat(s.m_local_to_world, np.float64), o, d, np.float64)
        safe = margin > 1e-4
        assert (h64[safe] == got_hit[safe]).all()
        both = safe & h64 & (cond > 0.05)   # 1e-5: origin within 25 local radii, not within 2.5 % of the surface, and the f32 discriminant keeps >= 5 % of half_b^2 (not grazing)
        assert (np.abs(t[both] - t64[both]) / t_scale[both]).max() <= 1e-5   # relative to max(|t|, (|half_b| + sqrt(disc)) / a)
        scale = np.maximum(1.0, np.maximum(t64[both], np.abs(p64[both]).max(axis=1)))[:, None]   # a point is only as good as t * |d|
        assert (np.abs(pts[both] - p64[both]) / scale).max() <= 1e-5 and np.abs(nrm[both] - n64[both]).max() <= 1e-4
        graz = safe & h64 & ~(cond > 0.05)   # grazing or far: same root, looser t (cancellation amplifies the f32 round-off)
        assert graz.sum() == 0 or rel_err(t[graz], t64[graz]).max() <= 2e-3
        n_tol += int(both.sum())
        assert (ids[got_hit] == A.obj_id(A.TYPE_SPHERE, 0)).all()
        n_checked += int(safe.sum())
        assert 0.2 < got_hit.mean() < 0.95   # the sweep exercises both outcomes
    assert n_checked >= 100000 and n_tol >= 20000   # decisions on > 10^5 rays, the 1e-5 bar on the well-conditioned hits among them


# ---- ray_plane_intersect / InfinitePlane::hit, traced.rs:91-114 -------------------------------------------------------------
def plane_t(n, c, o, d, dtype):
    div = dot3(n, d)
    par = np.abs(div) < dtype(0.000001)
    num = -dot3(n, o - c)
    with np.errstate(divide="ignore", invalid="ignore"):
        t = np.where(par, dtype(np.inf), num / div)
    return t, np.where(par, dtype(0.0), div)


def plane_hit(n, c, o, d_raw, dtype):
    o, d = o.astype(dtype), unit(d_raw.astype(dtype))
    t, div = plane_t(n, c, o, d, dtype)
    hit = ~(np.isinf(t) & (t > 0)) & ~(t < dtype(TMIN)) & ~(t > dtype(TMAX))
    normal = n * np.copysign(dtype(1.0), -div)[..., None]
    return hit, t, ray_at(o, d, t), normal, div


def test_infinite_plane_hit(api, O, A):
    rng = np.random.default_rng(102)
    matl = api.Material.new_metal((0.5, 0.5, 0.5))
    n_checked = 0
    for _ in range(N_OBJ):
        centre, nraw = rng.uniform(-5, 5, 3), rng.normal(size=3)
        p = api.InfinitePlane.new(centre, nraw, matl)
        osc = O.OracleScene().add("infinite_planes", (A.InfinitePlane * 1)(p))
        o, d = rays_towards(rng, centre, 3.0, N_RAYS, inside_frac=0.0)
        d[::7] = np.cross(np.array(p.normal[:]), rng.normal(size=3)).astype(F)   # rays parallel to the plane: the 1e-6 cutoff
        ids, t, pts, nrm = osc.trace_rays(o, d, TMIN, TMAX)
        got_hit = ids != 0
        n32, c32 = np.array(p.normal[:], F), np.array(p.center[:], F)
        h32, t32, p32, nn32, _ = plane_hit(n32, c32, o, d, F)
        assert (h32 == got_hit).all()
        assert (bits(t32[h32]) == bits(t[h32])).all() and (bits(p32[h32]) == bits(pts[h32])).all() and (bits(nn32[h32]) == bits(nrm[h32])).all()
        h64, t64, p64, nn64, div = plane_hit(n32.astype(np.float64), c32.astype(np.float64), o, d, np.float64)
        with np.errstate(invalid="ignore"):
            safe = (np.abs(div) > 0.02) & (np.abs(t64 - TMIN) > 1e-4) & (np.abs(t64 - TMAX) > 1e-2)   # n.d cancels for rays nearly in the plane
        assert (h64[safe] == got_hit[safe]).all()
        both = safe & h64
        t_scale = np.maximum(np.abs(t64), np.linalg.norm(o.astype(np.float64) - c32.astype(np.float64), axis=1))   # t = -n.(o - c) / n.d: relative to |o - c|
        assert (np.abs(t[both] - t64[both]) / t_scale[both]).max() <= 1e-5 and (nrm[both] == nn64[both].astype(F)).all()
        n_checked += int(safe.sum())
        assert not got_hit[::7].any()       # parallel rays never hit
    assert n_checked >= 100000


# ---- Barycentric::hit_aux, traced.rs:156-200 (Parallelogram = 0, Triangle = 1) ------------------------------------------------
def bary_hit(b, triangle, o, d_raw, dtype):
    origin, uxv = np.array(b.origin[:], F).astype(dtype), np.array(b.uxv[:], F).astype(dtype)
    base_inv = np.array(b.base_inv[:], F).reshape(3, 3).astype(dtype)
    v_in_base = np.array(b.v_in_base[:], F).astype(dtype)
    u_length, v_length = dtype(F(b.u_length)), dtype(F(b.v_length))
    o, d = o.astype(dtype), unit(d_raw.astype(dtype))
    root, div = plane_t(uxv, origin, o, d, dtype)
    plane_ok = ~np.isinf(root) & ~(root < dtype(TMIN)) & ~(root > dtype(TMAX))
    point = ray_at(o, d, np.where(plane_ok, root, dtype(0.0)))
    pfo = point - origin
    rx, ry = dot3(base_inv[0], pfo), dot3(base_inv[2], pfo)      # coords2d.x(), coords2d.z()
    ux, uy, vx, vy = dtype(1.0), dtype(0.0), v_in_base[0], v_in_base[2]
    det = ux * vy - vx * uy
    l1 = ((rx * vy - vx * ry) / det) / u_length
    l2 = (-(rx * uy - ux * ry) / det) / v_length
    l3 = dtype(1.0) - l1 - l2
    inside = (l1 > 0) & (l2 > 0) & (l1 < 1) & (l2 < 1)
    if triangle:
        inside &= (l3 > 0) & (l3 < 1)
    normal = uxv * np.copysign(dtype(1.0), -div)[..., None]
    edge = np.fmin.reduce([np.abs(l1), np.abs(l2), np.abs(1 - l1), np.abs(1 - l2)] + ([np.abs(l3), np.abs(1 - l3)] if triangle else []))
    return plane_ok & inside, root, point, normal, div, edge


@pytest.mark.parametrize("triangle", [False, True])
def test_barycentric_hit(api, O, A, triangle):
    rng = np.random.default_rng(103 + int(triangle))
    matl = api.Material.new_lambertian((0.5, 0.5, 0.5))
    n_checked = n_hits = 0
    for _ in range(N_OBJ):
        p0 = rng.uniform(-4, 4, 3)
        e1, e2 = rng.normal(size=3) * rng.uniform(0.5, 2.5), rng.normal(size=3) * rng.uniform(0.5, 2.5)
        if np.linalg.norm(np.cross(e1, e2)) < 0.2 * np.linalg.norm(e1) * np.linalg.norm(e2):
            e2 = np.cross(e1, rng.normal(size=3))   # keep the edges well apart (det = v_in_base.z away from 0)
        ctor = api.Triangle if triangle else api.Parallelogram
        b = ctor.new3points(p0, p0 + e1, p0 + e2, matl).pod
        kind = "triangles" if triangle else "parallelograms"
        osc = O.OracleScene().add(kind, (A.Barycentric * 1)(b))
        centre = p0 + 0.5 * (e1 + e2) * (0.66 if triangle else 1.0)
        o, d = rays_towards(rng, centre, 0.3 * (np.linalg.norm(e1) + np.linalg.norm(e2)), N_RAYS, inside_frac=0.0)
        ids, t, pts, nrm = osc.trace_rays(o, d, TMIN, TMAX)
        got_hit = ids != 0
        h32, t32, p32, n32, _, _ = bary_hit(b, triangle, o, d, F)
        assert (h32 == got_hit).all()
        assert (bits(t32[h32]) == bits(t[h32])).all() and (bits(p32[h32]) == bits(pts[h32])).all() and (bits(n32[h32]) == bits(nrm[h32])).all()
        h64, t64, p64, n64, div, edge = bary_hit(b, triangle, o, d, np.float64)
        with np.errstate(invalid="ignore"):
            safe = (np.abs(div) > 0.02) & (np.abs(t64 - TMIN) > 1e-4) & (np.abs(t64 - TMAX) > 1e-2) & (edge > 1e-4)
        assert (h64[safe] == got_hit[safe]).all()
        both = safe & h64
        if both.any():
            t_scale = np.maximum(np.abs(t64), np.linalg.norm(o.astype(np.float64) - np.array(b.origin[:], np.float64), axis=1))
            assert (np.abs(t[both] - t64[both]) / t_scale[both]).max() <= 1e-5
            assert (nrm[both] == n64[both].astype(F)).all()
        assert (ids[got_hit] == A.obj_id(A.TYPE_TRIANGLE if triangle else A.TYPE_PARALLELOGRAM, 0)).all()
        n_checked += int(safe.sum()); n_hits += int(got_hit.sum())
    assert n_checked >= 100000 and n_hits >= 8000


# ---- Cube::hit, traced.rs:265-299 ------------------------------------------------------------------------------------------------
def cube_hit(w2l, l2w, o, d_raw, dtype):
    o, d = o.astype(dtype), unit(d_raw.astype(dtype))
    lo, ld = dot_p3(w2l, o), dot_v3(w2l, d)
    n = lo.shape[0]
    smallest = np.full(n, np.inf, dtype)
    idx = np.full(n, -1)
    band = np.full(n, np.inf)       # distance of every decision from its threshold (float64 guard band)
    half, tmin, tmax = dtype(0.5), dtype(TMIN), dtype(TMAX)
    for i in range(3):
        with np.errstate(divide="ignore", invalid="ignore"):
            live = ~(np.abs(ld[:, i]) < dtype(0.00001))
            t1 = (half - lo[:, i]) / ld[:, i]
            t2 = (-half - lo[:, i]) / ld[:, i]
            t = np.where((t1 >= 0) & (t2 >= 0), np.minimum(t1, t2), np.maximum(t1, t2))
            cand = live & ~((t > smallest) | (t > tmax) | (t < tmin))
            f = np.abs(ray_at(lo, ld, np.where(cand, t, dtype(0.0)))).max(axis=1)
            on_surface = np.abs(f - half) <= dtype(0.00001)
            take = cand & on_surface
            band = np.fmin(band, np.where(live, np.fmin(np.abs(np.abs(f - half) - 1e-5) * 1e3, np.fmin(np.abs(t - tmin), np.abs(t - smallest))), 0.0))
        smallest = np.where(take, t, smallest)
        idx = np.where(take, i, idx)
    hit = idx >= 0
    lp = ray_at(lo, ld, np.where(hit, smallest, dtype(0.0)))
    axis = np.eye(3, dtype=dtype)[np.maximum(idx, 0)]
    sign = np.copysign(dtype(1.0), np.take_along_axis(lp, np.maximum(idx, 0)[:, None], axis=1)[:, 0])
    ln = axis * sign[:, None]
    with np.errstate(divide="ignore", invalid="ignore"):  # t = (+-0.5 - o_i) / d_i on the chosen axis: its f32 error is relative to these terms
        k = np.maximum(idx, 0)[:, None]
        t_scale = np.maximum(np.abs(smallest), (0.5 + np.abs(lo).max(axis=1)) / np.abs(np.take_along_axis(ld, k, axis=1)[:, 0]))
    return hit, smallest, dot_p3(l2w, lp), dot_v3(l2w, ln), band, t_scale   # the normal is NOT normalised (reference quirk)


def test_cube_hit(api, O, A):
    rng = np.random.default_rng(105)
    matl = api.Material.new_metal((0.5, 0.5, 0.5))
    n_checked = n_hits = 0
    for _ in range(N_OBJ):
        m, centre, smax = random_trs(api, rng, 0.5, 2.5)
        c = api.Cube.new(m, matl)
        osc = O.OracleScene().add("cubes", (A.Cube * 1)(c))
        o, d = rays_towards(rng, centre, 0.7 * smax, N_RAYS, inside_frac=0.1)
        ids, t, pts, nrm = osc.trace_rays(o, d, TMIN, TMAX)
        got_hit = ids != 0
        h32, t32, p32, n32, _, _ = cube_hit(mat(c.m_world_to_local, F), mat(c.m_local_to_world, F), o, d, F)
        assert (h32 == got_hit).all()
        assert (bits(t32[h32]) == bits(t[h32])).all() and (bits(p32[h32]) == bits(pts[h32])).all() and (bits(n32[h32]) == bits(nrm[h32])).all()
        h64, t64, p64, n64, band, t_scale = cube_hit(mat(c.m_world_to_local, np.float64), mat(c.m_local_to_world, np.float64), o, d, np.float64)
        safe = band > 2e-3   # |f - 0.5| at least 2e-6 away from the 1e-5 surface threshold, roots 2e-3 away from t_min and from each other
        assert (h64[safe] == got_hit[safe]).all()
        both = safe & h64
        assert (np.abs(t[both] - t64[both]) / t_scale[both]).max() <= 1e-5 and np.abs(nrm[both] - n64[both]).max() <= 1e-5 * np.abs(n64[both]).max()
        assert (ids[got_hit] == A.obj_id(A.TYPE_CUBE, 0)).all()
        n_checked += int(safe.sum()); n_hits += int((safe & got_hit).sum())
    assert n_checked >= 100000 and n_hits >= 8000


# ---- marched.rs: local_sdf of the three shapes (:57,:86-89,:134-138), to_local / to_world / to_world_f, the trait's sdf (:14-18),
# ---- get_outward_local_normal with eps = 1e-7 and the sign fix (:25-44), get_outward_normal (:19-24; sphere override :60-63) ------------
def length3(x, y, z):
    return np.sqrt(x * x + y * y + z * z)


class MSphere:
    def __init__(self, pod, dtype):
        self.c, self.r, self.T = np.array(pod.center[:], F).astype(dtype), dtype(F(pod.radius)), dtype

    def to_local(self, p, w):   # *p - p.w() * Vec4::new_v3(center)
        return p + (-(self.T(w) * self.c))

    def to_world_vec(self, n):  # *p + p.w() * center with w = 0
        return n + self.T(0.0) * self.c

    def to_world_f(self, f):
        return f

    def local_sdf(self, p):
        return length3(p[..., 0], p[..., 1], p[..., 2]) - self.r


class MBox(MSphere):
    def __init__(self, pod, dtype):
        self.c, self.s, self.T = np.array(pod.center[:], F).astype(dtype), np.array(pod.sizes[:], F).astype(dtype), dtype

    def local_sdf(self, p):
        q = np.abs(p) + (-self.s)
        m = np.maximum(q, self.T(0.0))
        return length3(m[..., 0], m[..., 1], m[..., 2]) + np.minimum(np.maximum(q[..., 0], np.maximum(q[..., 1], q[..., 2])), self.T(0.0))


class MTorus:
    def __init__(self, pod, dtype):
        g = lambda a, shape: np.array(a[:], F).reshape(shape).astype(dtype)
        self.l2w, self.w2l = g(pod.m_local_to_world_translate_rotate, (4, 4)), g(pod.m_world_to_local_translate_rotate, (4, 4))
        self.s, self.si, self.sz, self.T = g(pod.m_local_to_world_scale, 4), g(pod.m_world_to_local_scale, 4), g(pod.sizes, 3), dtype

    def _dot4(self, m, x, y, z, w):   # Mat4x4::dot: every row a full Vec4 dot
        return np.stack([m[i, 0] * x + m[i, 1] * y + m[i, 2] * z + m[i, 3] * w for i in range(3)], axis=-1)

    def to_local(self, p, w):     # TR^-1 . (p * scale^-1)
        return self._dot4(self.w2l, p[..., 0] * self.si[0], p[..., 1] * self.si[1], p[..., 2] * self.si[2], self.T(w) * self.si[3])

    def to_world_vec(self, n):    # (TR . (n, 0)) * scale
        return self._dot4(self.l2w, n[..., 0], n[..., 1], n[..., 2], self.T(0.0)) * self.s[:3]

    def to_world_f(self, f):      # f * scale.xyz().min_val()
        return f * np.minimum(self.s[0], np.minimum(self.s[1], self.s[2]))

    def local_sdf(self, p):
        qx = length3(p[..., 0], p[..., 2], self.T(0.0)) - self.sz[0]
        return length3(qx, p[..., 1], self.T(0.0)) - self.sz[1]


def m_sdf(ob, p):
    return ob.to_world_f(ob.local_sdf(ob.to_local(p, 1.0)))


def m_local_normal(ob, p):
    T = ob.T
    eps = T(F(0.0000001))
    e = [np.array(v, dtype=T) for v in ((eps, 0, 0), (0, eps, 0), (0, 0, eps))]
    with np.errstate(invalid="ignore", divide="ignore"):
        comp = [ob.local_sdf(p + e[i]) - ob.local_sdf(p + (-e[i])) for i in range(3)]
        normal = unit(np.stack(comp, axis=-1))
        d = unit(normal)                                   # Ray::new normalises again
        zero = np.zeros_like(d)
        start_val = ob.local_sdf(zero + T(0.0) * d)       # test_ray.at(0.)
        end_val = ob.local_sdf(zero + T(1.0) * d)         # test_ray.at(1.)
        sign = np.where(end_val > start_val, T(1.0), T(-1.0))
        return normal * sign[..., None]


def m_normal(ob, p):
    if type(ob) is MSphere:
        return unit(p + (-ob.c))
    with np.errstate(invalid="ignore", divide="ignore"):
        return unit(ob.to_world_vec(m_local_normal(ob, ob.to_local(p, 1.0))))


def same_f32(a, b):   # bit-equal, or NaN on both sides (x86 and numpy agree on the NaN they make, but its bits are not the point)
    a, b = np.asarray(a, F), np.asarray(b, F)
    return ((bits(a) == bits(b)) | (np.isnan(a) & np.isnan(b))).all()


def test_marched_sdf_and_numeric_normal(api, O, A):
    rng = np.random.default_rng(106)
    L = O.lib()
    matl = api.Material.new_dielectric(1.5)
    n_sdf = n_nrm = n_sign = 0
    for k in range(60):
        kind = k % 3
        if kind == 0:
            pod = api.MarchedSphere(rng.uniform(-3, 3, 3), rng.uniform(0.2, 1.5), matl); mk = MSphere; ext = pod.radius
        elif kind == 1:
            pod = api.MarchedBox(rng.uniform(-3, 3, 3), rng.uniform(0.2, 1.2, 3), matl); mk = MBox; ext = max(pod.sizes[:])
        else:
            m, _, _ = random_trs(api, rng, 0.6, 1.8)
            pod = api.MarchedTorus.new(m, (rng.uniform(0.4, 0.9), rng.uniform(0.05, 0.25), 0.0), matl); mk = MTorus; ext = 2.0
        o32, o64 = mk(pod, F), mk(pod, np.float64)
        # world points: a cloud around the object, and points ON its surface (found with the f64 sdf) where the normal matters
        if kind == 2:
            centre = np.array([pod.m_local_to_world_translate_rotate[3], pod.m_local_to_world_translate_rotate[7], pod.m_local_to_world_translate_rotate[11]]) \
                * np.array(pod.m_local_to_world_scale[:3])
        else:
            centre = np.array(pod.center[:])
        P = (centre + rng.uniform(-1.6, 1.6, (2000, 3)) * ext)
        g = np.zeros_like(P)
        for _ in range(12):   # a few Newton steps along the f64 numeric gradient bring the second half onto the surface
            h = 1e-4
            for i in range(3):
                dp = np.zeros(3); dp[i] = h
                g[:, i] = (m_sdf(o64, P + dp) - m_sdf(o64, P - dp)) / (2 * h)
            s = m_sdf(o64, P)
            P[1000:] -= (s[:, None] * g / np.maximum((g * g).sum(1, keepdims=True), 1e-6))[1000:]
        P32 = P.astype(F)
        want_s, want_n = m_sdf(o32, P32), m_normal(o32, P32)
        got_s, got_n = np.zeros(len(P32), F), np.zeros((len(P32), 3), F)
        out = (C.c_float * 3)()
        for i in range(len(P32)):
            pp = (C.c_float * 3)(*P32[i])
            got_s[i] = L.omo_marched_sdf(kind, C.byref(pod), pp)
            L.omo_marched_normal(kind, C.byref(pod), pp, out)
            got_n[i] = out[:]
        assert same_f32(got_s, want_s), "sdf differs from the f32 statement of marched.rs"
        assert same_f32(got_n, want_n), "normal differs from the f32 statement of marched.rs"
        # float64: the value of the sdf
        s64 = m_sdf(o64, P32.astype(np.float64))
        assert (np.abs(got_s - s64) <= 1e-5 * np.maximum(1.0, np.abs(P32).max(axis=1)) * 4).all()
        # float64: inside / outside is the sign of the sdf, exact away from the surface
        away = np.abs(s64) > 1e-4
        assert ((got_s > 0) == (s64 > 0))[away].all()
        # and the analytic direction of the sphere's normal
        if kind == 0:
            assert np.abs(got_n - unit(P32.astype(np.float64) - centre)).max() <= 1e-5
        n_sdf += len(P32); n_nrm += int(np.isfinite(got_n).all(axis=1).sum()); n_sign += int(away.sum())
    assert n_sdf >= 100000 and n_sign >= 50000 and n_nrm >= 40000


# ---- materials.rs:98-116 -------------------------------------------------------------------------------------------------------------
def reflect(v, n):            # *v - 2.*v.dot(*n)*(*n)
    T = v.dtype.type
    return v + (-(n * (T(2.0) * dot3(v, n))[..., None]))


def refract(uv, n, eta):
    T = uv.dtype.type
    cos_theta = np.minimum(dot3(-uv, n), T(1.0))
    perp = (uv + n * cos_theta[..., None]) * eta[..., None]
    aux = -np.sqrt(np.abs(T(1.0) - dot3(perp, perp)))
    return perp + n * aux[..., None]


def reflectance(cos, ri):
    T = cos.dtype.type
    r0 = (T(1.0) - ri) / (T(1.0) + ri)
    r0_2 = r0 * r0
    m = T(1.0) - cos
    return r0_2 + (T(1.0) - r0_2) * (m * m * m * m * m)


def test_reflect_refract_reflectance(O):
    rng = np.random.default_rng(107)
    L = O.lib()
    n_pts = 100000
    v = unit(rng.normal(size=(n_pts, 3))).astype(F)
    n = unit(rng.normal(size=(n_pts, 3))).astype(F)
    n[dot3(v, n) > 0] *= -1                               # the normal faces the incoming ray, as scatter_dielectric arranges
    eta = np.where(rng.random(n_pts) < 0.5, F(1.0) / F(1.5), F(1.5)).astype(F)
    cos = rng.uniform(0, 1, n_pts).astype(F)
    want_rf, want_rr, want_sc = reflect(v, n), refract(v, n, eta), reflectance(cos, eta)
    o3 = (C.c_float * 3)()
    for i in range(n_pts):
        vv, nn = (C.c_float * 3)(*v[i]), (C.c_float * 3)(*n[i])
        L.omo_reflect(vv, nn, o3)
        assert bits(np.array(o3[:], F)).tolist() == bits(want_rf[i]).tolist()
        L.omo_refract(vv, nn, float(eta[i]), o3)
        assert bits(np.array(o3[:], F)).tolist() == bits(want_rr[i]).tolist()
        assert bits(F(L.omo_reflectance(float(cos[i]), float(eta[i])))) == bits(want_sc[i])
    # float64: the laws themselves.  A reflection keeps the length and flips the normal component; a refraction that is
    # physically possible obeys Snell (sin_t = eta * sin_i) and has unit length.
    v64, n64, e64 = v.astype(np.float64), n.astype(np.float64), eta.astype(np.float64)
    r = want_rf.astype(np.float64)
    assert np.abs(np.linalg.norm(r, axis=1) - 1).max() <= 1e-5 and np.abs(dot3(r, n64) + dot3(v64, n64)).max() <= 1e-5
    sin_i = np.sqrt(np.maximum(0, 1 - dot3(v64, n64) ** 2))
    ok = e64 * sin_i < 0.999
    t = want_rr.astype(np.float64)[ok]
    sin_t = np.linalg.norm(np.cross(t, n64[ok]), axis=1)
    assert np.abs(np.linalg.norm(t, axis=1) - 1).max() <= 1e-5 and np.abs(sin_t - (e64 * sin_i)[ok]).max() <= 1e-5
    r0 = ((1 - e64) / (1 + e64)) ** 2
    assert np.abs(want_sc - (r0 + (1 - r0) * (1 - cos.astype(np.float64)) ** 5)).max() <= 1e-6


# ---- Stats::add render_thread.rs:23-39, normalize_color utils.rs:12-17, to_u8x3 vec3.rs:73-76 ---------------------------------------------
def as_u8(x):        # Rust `as u8`: saturating, NaN -> 0
    x = np.where(np.isnan(x), 0.0, x)
    return np.clip(np.trunc(x), 0, 255).astype(np.uint8)


def tone_u8(avg):    # clamp(c.sqrt(), 0, 0.999) * 256 ; f32::max / min drop a NaN operand
    with np.errstate(invalid="ignore"):
        s = np.sqrt(avg)
    s = np.fmin(np.fmax(s, F(0.0)), F(0.999))
    return as_u8(s * F(256.0))


MASK = (1 << 64) - 1


def scramble(obj_id):     # utils.rs:45-56
    id1 = obj_id & 0xFFFFFFFF
    id1 ^= (id1 << 13) & MASK; id1 ^= id1 >> 7; id1 ^= (id1 << 17) & MASK
    id2 = obj_id >> 32
    id2 ^= (id2 << 13) & MASK; id2 ^= id2 >> 17; id2 ^= (id2 << 5) & MASK
    return ((id2 << 32) & MASK) ^ id1 ^ ((id1 * id2) & MASK)


BLOOM = [(456894789, 348764781, 17287318477382145149), (56456, 2345, 10520185020478678957), (12337, 7878, 6100366985798845493),
         (7438554325, 2554, 2581451885731034521), (12345, 123123044, 2015400956511055807), (6373412378, 12452, 8800267423223100703),
         (3453453, 7874856378, 7039701875810786467), (999465, 143, 3008457310659543551), (14444, 111345, 5935720376112203207)]


def bloom_hash(i):        # utils.rs:93-107
    if i == 0:
        return 0
    ret = 0
    for m, a, p in BLOOM:
        ret |= 1 << (scramble(((i * m + a) & MASK) % p) % 64)
    return ret


def test_bloom_filter_and_scramble_in_python_ints(O, A):
    rng = np.random.default_rng(108)
    L = O.lib()
    ids = [0, 1, 2, MASK, 1 << 32, (1 << 32) - 1, A.obj_id(1, 0), A.obj_id(8, 0)]
    ids += [A.obj_id(int(t), int(i)) for t, i in zip(rng.integers(1, 10, 60000), rng.integers(0, 1 << 20, 60000))]
    ids += [int(x) for x in rng.integers(0, 1 << 63, 40000, dtype=np.uint64)] + [int(x) | (1 << 63) for x in rng.integers(0, 1 << 63, 2000, dtype=np.uint64)]
    assert len(ids) >= 100000
    for i in ids:
        assert L.omo_scramble(i) == scramble(i)
        assert L.omo_bloom_hash(i) == bloom_hash(i)
    assert scramble(0) == 0 and bloom_hash(0) == 0
    assert all(1 <= bin(bloom_hash(i)).count("1") <= 9 for i in ids[1:2000])


def test_stats_add_recurrence_and_tone_map(O, A):
    rng = np.random.default_rng(109)
    L = O.lib()
    n_px, n_steps = 2000, 60
    # tone map alone, incl. the edges: 0, just under / over 1, huge, negative (sqrt -> NaN -> 0), -0
    cols = np.concatenate([rng.uniform(0, 1.2, (100000, 3)), [[0, 0.25, 1.0], [0.998, 0.999 ** 2, 1e30], [-1.0, -0.0, 4.0]]]).astype(F)
    want = tone_u8(cols)
    out = (C.c_uint8 * 3)()
    for i in range(len(cols)):
        L.omo_tone_map((C.c_float * 3)(*cols[i]), out)
        assert out[0] == want[i, 0] and out[1] == want[i, 1] and out[2] == want[i, 2]
    assert tone_u8(np.array([0.25], F))[0] == 128 and tone_u8(np.array([1.0], F))[0] == 255
    # the recurrence: n_px independent pixels fed n_steps samples; flat pixels retire (bad_avgs reaches 5), noisy ones reset
    px = (A.Pixel * n_px)()
    s_sum = np.zeros((n_px, 3), F); s_n = np.zeros(n_px, np.uint32); s_col = np.zeros((n_px, 3), np.uint8)
    s_bad = np.zeros(n_px, np.uint32); s_dep = np.zeros(n_px, F); s_bloom = [0] * n_px
    flat = rng.random(n_px) < 0.5
    base = rng.uniform(0, 1, (n_px, 3)).astype(F)
    done_seen = 0
    for step in range(n_steps):
        x = np.where(flat[:, None], base, rng.uniform(0, 1.5, (n_px, 3)).astype(F)).astype(F)
        x[step % 7::97] = -0.0                                   # ray_color's "-black" for an exhausted path
        depth = np.where(rng.random(n_px) < 0.2, np.inf, rng.uniform(0.1, 90, n_px)).astype(F)
        oid = [0 if np.isinf(depth[i]) else A.obj_id(1 + (i + step) % 8, (i * 7 + step) % 500) for i in range(n_px)]
        old = s_col.copy()
        s_sum = s_sum + x
        s_n = s_n + 1
        nf = s_n.astype(F)
        s_col = tone_u8(s_sum * (F(1.0) / nf)[:, None])          # Div<f32> = multiply by the reciprocal
        bad_run = (old == s_col).all(axis=1).astype(np.uint32)
        s_bad = (s_bad + bad_run) * bad_run
        with np.errstate(invalid="ignore"):
            s_dep = ((nf - F(1.0)) * s_dep + depth) / nf
        for i in range(n_px):
            s_bloom[i] |= bloom_hash(oid[i])
            d = L.omo_stats_add(C.byref(px[i]), (C.c_float * 3)(*x[i]), float(depth[i]), oid[i])
            assert d == int(s_bad[i] >= 5)
            done_seen += d
        got = np.frombuffer(px, dtype=A.PIXEL_DTYPE)
        assert (bits(got["sum"]) == bits(s_sum)).all() and (got["n"] == s_n).all() and (got["color"] == s_col).all()
        assert (got["bad_avgs"] == s_bad).all() and same_f32(got["avg_depth"], s_dep)
        assert got["bloom"].tolist() == s_bloom
    assert done_seen > 10000 and (s_bad[~flat] < 5).mean() > 0.5     # both regimes were exercised
    assert n_px * n_steps >= 100000


# ---- render_thread.rs:183-192 + Camera::get_ray camera.rs:60-74 + rand_in_unit_disc vec3.rs:108-113 ------------------------------------
# rand::thread_rng is replaced by a counter RNG on both sides of this build; the test takes the uniform stream AS DATA
# (omo_rand_stream) and restates what the reference does with consecutive draws.
def rand_stream(O, seed, a, b, n, stream=0):
    out = np.zeros(n, F)
    O.lib().omo_rand_stream(seed, a, b, stream, n, out.ctypes.data)
    return out


def f32v(a):
    return np.array(a[:], F)


def camera_ray_f32(cam, u, v, disc):
    lens = disc * F(cam.lens_radius)                                 # self.lens_radius * Vec3::rand_in_unit_disc()
    offset = f32v(cam.u_of_plane) * lens[0] + f32v(cam.v_of_plane) * lens[1]
    llc_o = f32v(cam.lower_left_corner) + (-f32v(cam.origin))
    h, vv = f32v(cam.horizontal), f32v(cam.vertical)
    direction = h * u + vv * v + F(0.0) * F(0.0) + llc_o * F(1.0)     # rows (h_i, v_i, 0, llc_o_i) . (u, v, 0, 1)
    return f32v(cam.origin) + offset, unit(unit(direction + (-offset)))   # .unit() in get_ray, again in Ray::new


def test_primary_rays_from_the_uniform_stream(api, O, A):
    W, H, spp, seed = 37, 23, 8, 77
    cam = api.Camera.new((13, 2, 3), (0, 0, 0), (0, 1, 0), 20.0, W / H, 0.1, 10.0)
    par = api.render_params(W, H, spp, seed=seed)
    jit = np.zeros(spp, np.uint8)
    O.lib().omo_jitter_table(spp, seed, jit.ctypes.data)
    assert sorted(jit.tolist()) == sorted([(((s // 2) & 1) << 1) | (s & 1) for s in range(spp)])   # a shuffle of the 2x2 strata table
    # Camera::new (camera.rs:38-59) in float64
    vfov = np.deg2rad(20.0); hh = np.tan(vfov / 2) * 10.0
    w = unit(np.array([13.0, 2, 3])); uu = unit(np.cross([0, 1.0, 0], w)); vv = unit(np.cross(w, uu))
    assert np.abs(f32v(cam.horizontal) - 2 * hh * (W / H) * uu).max() <= 1e-5 and np.abs(f32v(cam.vertical) - 2 * hh * vv).max() <= 1e-5
    assert np.abs(f32v(cam.lower_left_corner) - (np.array([13.0, 2, 3]) - hh * (W / H) * uu - hh * vv - 10.0 * w)).max() <= 2e-5
    assert cam.lens_radius == F(0.05)
    n = 0
    for sample in (0, 3, 7):
        go, gd = O.primary_rays(cam, par, 1, sample)
        for pxl in range(0, W * H, 3):
            r = rand_stream(O, seed, pxl, sample, 64)
            line, col = pxl // W, pxl - W * (pxl // W)
            i_rand = (r[0] + F((jit[sample] >> 1) & 1)) / F(2.0)
            j_rand = (r[1] + F(jit[sample] & 1)) / F(2.0)
            u = (F(col) + i_rand) / (F(W) - F(1.0))
            v = F(1.0) - (F(line) + j_rand) / (F(H) - F(1.0))
            k = 2
            while True:                                                # rand_in_unit_disc: rand_range(-1, 1) = rand() * (max - min) + min
                p = np.array([r[k] * (F(1.0) - F(-1.0)) + F(-1.0), r[k + 1] * (F(1.0) - F(-1.0)) + F(-1.0), F(0.0)], F)
                k += 2
                if dot3(p, p) < F(1.0):
                    break
            wo, wd = camera_ray_f32(cam, u, v, p)
            assert bits(wo).tolist() == bits(go[pxl]).tolist() and bits(wd).tolist() == bits(gd[pxl]).tolist()
            n += 1
    assert n >= 800


# ---- Material::scatter materials.rs:39-95 with rand_in_unit_sphere / rand_unit_vector vec3.rs:92-100 -----------------------------------------
def scatter_f32(m, in_dir, point, normal, r):
    k = [0]

    def rnd():
        k[0] += 1
        return r[k[0] - 1]

    def in_unit_sphere():
        while True:
            p = np.array([rnd() * (F(1.0) - F(-1.0)) + F(-1.0) for _ in range(3)], F)
            if dot3(p, p) < F(1.0):
                return p
    if m.mat_type == 0:      # LAMBERTIAN
        new_dir = normal + unit(in_unit_sphere())
        if (np.abs(new_dir) < F(1e-8)).all():
            new_dir = normal
        att = f32v(m.albedo)
    elif m.mat_type == 1:    # METAL: no absorb test
        new_dir = reflect(in_dir, normal) + in_unit_sphere() * F(m.fuzz)
        att = f32v(m.albedo)
    else:                    # DIELECTRIC
        front = dot3(in_dir, normal) < F(0.0)
        ratio = F(1.0) / F(m.ior) if front else F(m.ior)
        nn = normal if front else -normal
        cos_theta = np.minimum(dot3(-in_dir, nn), F(1.0))
        sin_theta = np.sqrt(F(1.0) - cos_theta * cos_theta)
        cannot = ratio * sin_theta > F(1.0)
        by_reflectance = reflectance(np.array(cos_theta, F), np.array(ratio, F)) > rnd()
        new_dir = reflect(in_dir, nn) if (cannot or by_reflectance) else refract(in_dir, nn, np.array(ratio, F))
        att = np.array([1, 1, 1], F)
    return att, point, unit(new_dir)


def test_scatter_draw_order_and_arithmetic(api, O, A):
    rng = np.random.default_rng(110)
    L = O.lib()
    mats = [api.Material.new_lambertian((0.3, 0.6, 0.9)), api.Material.new_metal_fuzz((0.8, 0.7, 0.6), 0.35), api.Material.new_metal((0.9, 0.9, 0.9)),
            api.Material.new_dielectric(1.5)]
    att, oo, od = (C.c_float * 3)(), (C.c_float * 3)(), (C.c_float * 3)()
    n_reflect = n = 0
    for i in range(4000):
        m = mats[i % 4]
        d = unit(rng.normal(size=3)).astype(F)
        nrm = unit(rng.normal(size=3)).astype(F)
        if i % 8 < 6 and dot3(d, nrm) > 0:
            nrm = -nrm                                                  # mostly front faces; the rest leave from inside
        pt = rng.uniform(-5, 5, 3).astype(F)
        L.omo_scatter(C.byref(m), (C.c_float * 3)(0, 0, 0), (C.c_float * 3)(*d), (C.c_float * 3)(*pt), (C.c_float * 3)(*nrm), 5, i, 3, att, oo, od)
        wa, wo, wd = scatter_f32(m, d, pt, nrm, rand_stream(O, 5, i, 3, 96))
        assert bits(f32v(att)).tolist() == bits(wa).tolist() and bits(f32v(oo)).tolist() == bits(wo).tolist()
        assert bits(f32v(od)).tolist() == bits(wd).tolist(), (i, m.mat_type)
        if m.mat_type == 2:
            n_reflect += int(dot3(wd, nrm) * dot3(d, nrm) < 0)
        n += 1
    assert 50 < n_reflect < 950     # the dielectric both reflects and refracts in the sweep


# ---- FrozenHittableList::hit, marched part hits.rs:288-332, and unstuck hits.rs:336-365 ----------------------------------------------------
def march_f32(objs, o, d, closest, t_min=F(TMIN), t_max=F(TMAX), hit_size=F(0.001)):
    """objs in the reference's order (spheres, boxes, tori).  Vectorised over rays; returns (hit, t, point, winner index)."""
    n = len(o)
    def nearest(p):
        dist = np.full(n, np.inf, F); who = np.full(n, -1)
        for j, ob in enumerate(objs):
            dj = np.abs(m_sdf(ob, p))
            better = dj < dist
            dist = np.where(better, dj, dist); who = np.where(better, j, who)
        return dist, who
    # unstuck: nearest object at r.at(t_min); step 0.0005 while |sdf of THAT object| < HIT_SIZE
    step = hit_size / F(2.0)
    new_t = np.full(n, t_min, F)
    aux, who0 = nearest(ray_at(o, d, new_t))
    for _ in range(1 << 20):   # the reference's loop has no cap (hits.rs:360-363); a ray skimming inside the 1e-3 shell takes hundreds of steps
        stuck = aux < hit_size
        if not stuck.any():
            break
        new_t = np.where(stuck, new_t + step, new_t)
        p = ray_at(o, d, new_t)
        each = np.stack([np.abs(m_sdf(ob, p)) for ob in objs], axis=0)
        aux = np.where(stuck, each[who0, np.arange(n)], aux)
    t = new_t.copy()
    hit = np.zeros(n, bool); winner = np.full(n, -1)
    live = np.ones(n, bool)
    for _ in range(1024):
        live &= (t < t_max) & (t < closest)
        if not live.any():
            break
        dist, who = nearest(ray_at(o, d, t))
        got = live & (dist < hit_size)
        hit |= got; winner = np.where(got, who, winner)
        live &= ~got
        t = np.where(live, t + dist, t)
    return hit, t, ray_at(o, d, t), winner


def test_march_and_unstuck(api, O, A):
    rng = np.random.default_rng(111)
    matl = api.Material.new_dielectric(1.5)
    m, _, _ = random_trs(api, rng, 0.7, 1.4)
    pods = [api.MarchedSphere((-2.0, 0.5, 0.3), 0.8, matl), api.MarchedBox((1.5, -0.5, 0.0), (0.6, 0.4, 0.7), matl),
            api.MarchedTorus.new(api.Mat4x4.new_translate((0.0, 1.6, -0.4)) ^ api.Mat4x4.new_rotate_x(0.6) ^ api.Mat4x4.new_scale((1.2, 0.9, 1.1)), (0.6, 0.18, 0.0), matl)]
    wall = api.Sphere.new_with_radius((0.0, 0.0, -6.0), 3.0, api.Material.new_lambertian((0.5, 0.5, 0.5)))   # a traced object behind: the `closest_so_far` bound
    osc = O.OracleScene()
    osc.add("marched_spheres", (A.MarchedSphere * 1)(pods[0])).add("marched_boxes", (A.MarchedBox * 1)(pods[1])).add("marched_tori", (A.MarchedTorus * 1)(pods[2]))
    osc.add("spheres", (A.Sphere * 1)(wall))
    objs = [MSphere(pods[0], F), MBox(pods[1], F), MTorus(pods[2], F)]
    ids_of = [A.obj_id(A.TYPE_MARCHED_SPHERE, 0), A.obj_id(A.TYPE_MARCHED_BOX, 0), A.obj_id(A.TYPE_MARCHED_TORUS, 0)]
    aims = [np.array([-2.0, 0.5, 0.3]), np.array([1.5, -0.5, 0.0]), np.array([0.0, 1.6 * 0.9, -0.4 * 1.1])]   # the three objects' centres
    both = [rays_towards(rng, c, 0.8, 2000, inside_frac=0.0) for c in aims]
    o, d = np.concatenate([b[0] for b in both]), np.concatenate([b[1] for b in both])
    total = stuck_starts = 0
    for gen in range(2):
        dn = unit(d.astype(F))
        ids, t, pts, nrm = osc.trace_rays(o, d, TMIN, TMAX)
        th, tt, _, _, _, _ = sphere_hit(mat(wall.m_world_to_local, F), mat(wall.m_local_to_world, F), o, d, F)
        closest = np.where(th, tt, F(TMAX)).astype(F)
        mh, mt, mp, who = march_f32(objs, o.astype(F), dn, closest)
        want_ids = np.where(mh, np.array(ids_of, np.uint64)[np.maximum(who, 0)], np.where(th, np.uint64(A.obj_id(A.TYPE_SPHERE, 0)), np.uint64(0)))
        assert (ids == want_ids).all()                                       # a marched hit always wins; else the traced record stands
        assert (bits(t[mh]) == bits(mt[mh])).all() and (bits(pts[mh]) == bits(mp[mh])).all()
        for j, ob in enumerate(objs):                                        # the normal of the nearest object at the accepted point
            sel = mh & (who == j)
            if sel.any():
                assert same_f32(nrm[sel], m_normal(ob, mp[sel]))
        total += int(mh.sum())
        # second generation: rays that START on the marched surfaces (what every scattered ray does) -> unstuck
        keep = np.nonzero(mh)[0][:3000]
        o = mp[keep].astype(F)
        d = rng.normal(size=(len(keep), 3)).astype(F)
        stuck_starts = max(stuck_starts, len(keep))
    assert total >= 3000 and stuck_starts >= 1500 and th.any() and (closest < F(TMAX)).any()
