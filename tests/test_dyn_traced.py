"""`Arc<dyn Traced>` as data (om_dyn_traced; hits.rs:39,91-95,280-285): a user type described as a quadric in its own frame.
CPU: the oracle's statement of it equals the oracle's typed Sphere bit for bit for the unit sphere, and the mathematics of a clipped
cylinder in float64.  GPU: hits and rendered pixels equal the oracle's bit for bit in every pipeline; dyn objects are scanned after the
parallelograms and win ties like a later object; a user ellipsoid renders the same radiance as the typed Sphere."""
import ctypes as C

import numpy as np
import pytest


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def trs(api, t, r, s):
    return (api.Mat4x4.new_translate(t) ^ api.Mat4x4.new_rotate_x(r[0]) ^ api.Mat4x4.new_rotate_y(r[1]) ^ api.Mat4x4.new_rotate_z(r[2]) ^ api.Mat4x4.new_scale(s))


def rays(rng, centre, radius, n):
    o = centre + rng.normal(size=(n, 3)) * radius * rng.uniform(2.0, 9.0, size=(n, 1)) / np.sqrt(3.0)
    inside = rng.random(n) < 0.15
    o[inside] = centre + rng.uniform(-0.3, 0.3, size=(int(inside.sum()), 3)) * radius
    d = centre + rng.uniform(-1.3, 1.3, size=(n, 3)) * radius - o
    return o.astype(np.float32), d.astype(np.float32)


def test_oracle_quadric_equals_typed_sphere_and_f64_cylinder(api, O, A):
    rng = np.random.default_rng(31)
    matl = api.Material.new_metal_fuzz((0.8, 0.7, 0.6), 0.1)
    n_hits = 0
    for _ in range(40):
        t, r, s = rng.uniform(-5, 5, 3), rng.uniform(0, 6.28, 3), rng.uniform(0.3, 2.5, 3)
        m = trs(api, t, r, s)
        typed = O.OracleScene().add("spheres", (A.Sphere * 1)(api.Sphere.new(m, matl)))
        user = O.OracleScene().add_dyn_traced((A.DynTraced * 1)(api.DynTraced.ellipsoid(m, matl).pod))
        o, d = rays(rng, t, float(s.max()), 3000)
        i1, t1, p1, n1 = typed.trace_rays(o, d)
        i2, t2, p2, n2 = user.trace_rays(o, d)
        assert ((i1 != 0) == (i2 != 0)).all()
        h = i1 != 0
        assert (i2[h] == A.obj_id(A.TYPE_DYN_TRACED, 0)).all()
        assert (bits(t1[h]) == bits(t2[h])).all() and (bits(p1[h]) == bits(p2[h])).all() and (bits(n1[h]) == bits(n2[h])).all()
        n_hits += int(h.sum())
    assert n_hits > 30000
    # a clipped cylinder x^2 + z^2 = 1, |y| <= h, under a rigid map + uniform scale: closed form in float64
    for _ in range(20):
        t, r, sc, hh = rng.uniform(-3, 3, 3), rng.uniform(0, 6.28, 3), rng.uniform(0.5, 2.0), rng.uniform(0.4, 1.5)
        m = trs(api, t, r, (sc, sc, sc))
        cyl = api.DynTraced.cylinder(m, matl, half_height=hh)
        osc = O.OracleScene().add_dyn_traced((A.DynTraced * 1)(cyl.pod))
        o, d = rays(rng, t, sc * max(1.0, hh), 3000)
        ids, tt, pts, nrm = osc.trace_rays(o, d)
        w2l = np.array(cyl.pod.m_world_to_local[:], np.float64).reshape(4, 4)
        l2w = np.array(cyl.pod.m_local_to_world[:], np.float64).reshape(4, 4)
        dn = d.astype(np.float64) / np.linalg.norm(d.astype(np.float64), axis=1, keepdims=True)
        lo = o.astype(np.float64) @ w2l[:3, :3].T + w2l[:3, 3]
        ld = dn @ w2l[:3, :3].T
        a = ld[:, 0] ** 2 + ld[:, 2] ** 2
        hb = lo[:, 0] * ld[:, 0] + lo[:, 2] * ld[:, 2]
        c = lo[:, 0] ** 2 + lo[:, 2] ** 2 - 1
        disc = hb * hb - a * c
        with np.errstate(invalid="ignore", divide="ignore"):
            sq = np.sqrt(disc)
            r1, r2 = (-hb - sq) / a, (-hb + sq) / a
            ok1 = (disc >= 0) & (r1 >= 0.001) & (r1 <= 100) & (np.abs(lo[:, 1] + r1 * ld[:, 1]) <= hh)
            ok2 = (disc >= 0) & (r2 >= 0.001) & (r2 <= 100) & (np.abs(lo[:, 1] + r2 * ld[:, 1]) <= hh)
            want_t = np.where(ok1, r1, r2)
            want_hit = ok1 | ok2
            margin = np.fmin.reduce([np.abs(disc) / np.maximum(hb * hb, 1e-30), np.abs(r1 - 0.001), np.abs(r2 - 0.001), np.abs(np.abs(lo[:, 1] + r1 * ld[:, 1]) - hh),
                                     np.abs(np.abs(lo[:, 1] + r2 * ld[:, 1]) - hh), np.abs(a) * 100])
        safe = margin > 1e-3
        assert ((ids != 0) == want_hit)[safe].all()
        both = safe & want_hit
        scale = np.maximum(np.abs(want_t), (np.abs(hb) + np.where(disc > 0, sq, 0)) / np.maximum(a, 1e-30))
        assert (np.abs(tt[both] - want_t[both]) / scale[both]).max() <= 1e-5
        lp = lo[both] + want_t[both, None] * ld[both]
        wn = np.stack([lp[:, 0], np.zeros(len(lp)), lp[:, 2]], axis=1) @ l2w[:3, :3].T   # M . grad/2 with grad/2 = (x, 0, z)
        wn /= np.linalg.norm(wn, axis=1, keepdims=True)
        assert np.abs(nrm[both] - wn).max() <= 2e-4 and both.sum() > 300


def build_scene(api, rng):
    """every traced type + user quadrics: bounded ellipsoids, a clipped cylinder, a cone cap, an unbounded paraboloid (always-test list),
    and a dyn ellipsoid that COINCIDES with a typed sphere (the tie must go to the dyn object, scanned later)"""
    w = api.HittableList()
    lam, met, gla = api.Material.new_lambertian((0.6, 0.5, 0.4)), api.Material.new_metal_fuzz((0.8, 0.8, 0.9), 0.2), api.Material.new_dielectric(1.5)
    w += api.Sphere.new_with_radius((0, -100.5, 0), 100.0, lam)
    w += api.Sphere.new_with_radius((-2.2, 0.2, 0.5), 0.7, gla)
    twin = trs(api, (2.0, 0.3, -0.5), (0.3, 0.2, 0.1), (0.8, 0.5, 0.6))
    w += api.Sphere.new(twin, met)
    w += api.DynTraced.ellipsoid(twin, lam)                               # same surface as the typed sphere before it
    w += api.Cube.new(trs(api, (0.3, 0.0, 2.0), (0.2, 0.7, 0.0), (0.8, 0.8, 0.8)), met)
    w += api.Triangle.new3points((-1.5, -0.4, 2.5), (-0.5, 1.2, 2.2), (-2.5, 0.9, 2.0), lam)
    w += api.Parallelogram.new3points((1.2, -0.4, 2.6), (2.4, -0.3, 2.2), (1.3, 0.8, 2.4), met)
    w += api.InfinitePlane.new((0, 0, -6.0), (0, 0, 1), lam)
    for k in range(5):
        w += api.DynTraced.ellipsoid(trs(api, rng.uniform(-2.5, 2.5, 3) + (0, 0.8, 0), rng.uniform(0, 6.28, 3), rng.uniform(0.2, 0.6, 3)), [lam, met, gla][k % 3])
    w += api.DynTraced.cylinder(trs(api, (-0.6, 0.2, -1.2), (0.4, 0.0, 0.3), (0.5, 0.5, 0.5)), gla, half_height=1.2)
    cone = trs(api, (1.0, 1.4, 0.6), (3.14, 0.0, 0.2), (0.6, 0.9, 0.6))       # x^2 + z^2 = y^2 for 0 <= y <= 1
    w += api.DynTraced(cone, (1, -1, 1, 0, 0, 0, 0, 0, 0, 0), met, clip=((-2, 0, -2), (2, 1, 2)), bounds=api.DynTraced.world_box(cone, (-1, 0, -1), (1, 1, 1)))
    bowl = trs(api, (0.0, -0.45, -2.5), (0.0, 0.0, 0.0), (1.0, 1.0, 1.0))   # y = x^2 + z^2 below y = 0.6, no world box: "draw always"
    w += api.DynTraced(bowl, (1, 0, 1, 0, 0, 0, 0, -1, 0, 0), lam, clip=((-3, -1, -3), (3, 0.6, 3)))
    return w


def oracle_of(O, A, w):
    o = O.OracleScene()
    for k, arr in w.arrays().items():
        if arr is not None:
            o.add(k, arr)
    o.add_dyn_traced((A.DynTraced * len(w.dyn_traced))(*[d.pod for d in w.dyn_traced]))
    return o


def pixels_equal(a, b):
    for f in ("sum", "avg_depth"):
        assert (a[f].view(np.uint32) == b[f].view(np.uint32)).all(), f
    for f in ("n", "bad_avgs", "bloom", "color"):
        assert (a[f] == b[f]).all(), f


@pytest.mark.gpu
def test_dyn_traced_hits_and_pixels_bit_exact(api, O, A):
    rng = np.random.default_rng(5)
    w = build_scene(api, rng)
    fz = w.freeze()
    osc = oracle_of(O, A, w)
    assert len(fz.dyn_traced()) == len(w.dyn_traced) == 9 and fz.stats()["always_test"] >= 3   # ground, plane, bowl
    n = 30000
    o = (np.array([0.0, 1.0, 7.0]) + rng.normal(0, 0.8, (n, 3))).astype(np.float32)
    tgt = rng.uniform(-3, 3, (n, 3)).astype(np.float32)
    want = osc.trace_rays(o, tgt - o)
    for flags in (0, A.RENDER_BRUTE_FORCE, A.RENDER_NO_SMEM_SCENE, A.RENDER_BRUTE_FORCE | A.RENDER_NO_SMEM_SCENE):
        got = fz.hit(o, tgt - o, flags=flags)
        assert (got[0] == want[0]).all(), flags
        for g, x in zip(got[1:], want[1:]):
            assert (bits(g) == bits(x)).all(), flags
    ids = want[0]
    assert (ids >> np.uint64(32) == 10).sum() > 2000                                   # the user objects are hit a lot
    assert not (ids == A.obj_id(A.TYPE_SPHERE, 2)).any() and (ids == A.obj_id(A.TYPE_DYN_TRACED, 0)).any()   # the coinciding pair: the later (dyn) object wins every tie
    cam = api.Camera.new((0, 1.2, 7.5), (0, 0.3, 0), (0, 1, 0), 40.0, 1.5, 0.05, 7.0)
    ref = osc.render(cam, api.render_params(72, 48, 6, seed=9))
    for flags in (A.RENDER_PERSISTENT, A.RENDER_WAVEFRONT, A.RENDER_POOL, A.RENDER_PERSISTENT | A.RENDER_BRUTE_FORCE):
        r = api.Renderer(fz, cam, api.render_params(72, 48, 6, seed=9, flags=flags))
        r.launch(0); r.sync()
        pixels_equal(r.read_pixels(), ref["pixels"])
        assert r.samples_traced() == ref["traced"]
        r.close()
    # adaptive, in passes
    ref = osc.render(cam, api.render_params(48, 32, 40, seed=2, flags=A.RENDER_ADAPTIVE))
    r = api.Renderer(fz, cam, api.render_params(48, 32, 40, seed=2, flags=A.RENDER_ADAPTIVE))
    for _ in range(5):
        r.launch(8)
    r.sync()
    pixels_equal(r.read_pixels(), ref["pixels"])
    r.close()


@pytest.mark.gpu
def test_user_ellipsoid_renders_like_the_typed_sphere(api, A):
    """The same scene with every Sphere handed over as an om_dyn_traced ellipsoid instead: same radiance and depth bit for bit
    (only the object ids, hence the bloom masks, differ)."""
    fz = api.FrozenHittableList.generate("rtiow_classic", 1)
    spheres = fz.lists()["spheres"]
    w = api.HittableList()
    for s in spheres:
        m = api.Mat4x4(s.m_local_to_world)
        d = api.DynTraced.ellipsoid(m, s.material, bounded=max(abs(x) for x in s.m_local_to_world[:3]) < 50)
        d.pod.m_world_to_local[:] = s.m_world_to_local[:]     # the very same inverse the typed sphere carries
        w += d
    fu = w.freeze()
    cam = api.Camera.new((13, 2, 3), (0, 0, 0), (0, 1, 0), 20.0, 1.5, 0.1, 10.0)
    out = []
    for scene in (fz, fu):
        r = api.Renderer(scene, cam, api.render_params(120, 80, 8, seed=3))
        r.launch(0); r.sync()
        out.append(r.read_pixels().copy())
        r.close()
    a, b = out
    assert (a["sum"].view(np.uint32) == b["sum"].view(np.uint32)).all() and (a["avg_depth"].view(np.uint32) == b["avg_depth"].view(np.uint32)).all()
    assert (a["color"] == b["color"]).all() and (a["n"] == b["n"]).all()


def test_dyn_traced_validation(api, A):
    from raytracingoneweekend_b200 import _lib
    matl = api.Material.new_lambertian((0.5, 0.5, 0.5))
    m = api.Mat4x4.new_translate((0, 0, 0))
    bad = [api.DynTraced(m, (0,) * 9 + (1,), matl), api.DynTraced(m, (1, 1, 1, 0, 0, 0, 0, 0, 0, float("nan")), matl),
           api.DynTraced(m, (1, 1, 1, 0, 0, 0, 0, 0, 0, -1), matl, clip=((1, 0, 0), (0, 1, 1)))]
    for d in bad:
        w = api.HittableList()
        w += d
        with pytest.raises(api.OmError) as e:
            w.freeze()
        assert e.value.status == A.OM_ERR_INVALID
    w = api.HittableList()
    with pytest.raises(api.OmError) as e:     # a native trait object still cannot cross
        w += object()
    assert e.value.status == A.OM_ERR_UNSUPPORTED
