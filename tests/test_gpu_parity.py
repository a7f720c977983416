"""GPU parity tests: the CUDA path, called through the C ABI, against the oracle on the same
seeded inputs.  Default build (the reference's arithmetic, FMA only in culling): bit-exact.
OM_RENDER_FAST_MATH build: tolerances stated in each test."""
import ctypes
import os

import numpy as np
import pytest

from conftest import bits, oracle_scene_from

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def cam_rtiow(api, aspect=1.5, aperture=0.1):
    return api.Camera.new((13, 2, 3), (0, 0, 0), (0, 1, 0), 20.0, aspect, aperture, 10.0)  # main.rs:136-142


def rays_at_scene(n, seed=0, spread=6.0):
    rng = np.random.default_rng(seed)
    orig = np.tile(np.array([13, 2, 3], np.float32), (n, 1)) + rng.normal(0, 0.5, (n, 3)).astype(np.float32)
    tgt = rng.uniform(-spread, spread, (n, 3)).astype(np.float32)
    tgt[:, 1] = rng.uniform(0, 1.5, n)
    return orig, (tgt - orig).astype(np.float32)


def assert_hits_equal(g, o):
    gi, gt, gp, gn = g
    oi, ot, op, on = o
    assert (gi == oi).all()
    assert (bits(gt) == bits(ot)).all()
    assert (bits(gp) == bits(op)).all() and (bits(gn) == bits(on)).all()


def assert_pixels_equal(g, o):
    for f in ("sum", "avg_depth"):
        assert (bits(np.ascontiguousarray(g[f])) == bits(np.ascontiguousarray(o[f]))).all(), f
    for f in ("n", "bad_avgs", "bloom", "color"):
        assert (g[f] == o[f]).all(), f


@pytest.fixture(scope="module")
def rtiow(api, O):
    fz = api.FrozenHittableList.generate("random_scene", 1)
    return fz, oracle_scene_from(O, fz)


# ---- FrozenHittableList::hit ---------------------------------------------------------------------
@pytest.mark.parametrize("flags_name", ["bvh_smem", "brute", "bvh_global"])
def test_hit_bit_exact_rtiow(api, A, rtiow, flags_name):
    fz, osc = rtiow
    extra = {"bvh_smem": 0, "brute": A.RENDER_BRUTE_FORCE, "bvh_global": A.RENDER_NO_SMEM_SCENE}[flags_name]
    o, d = rays_at_scene(30000, 1)
    assert_hits_equal(fz.hit(o, d, flags=extra), osc.trace_rays(o, d))
    # secondary-like rays: origins on the ground / inside the scene, all directions
    rng = np.random.default_rng(2)
    o2 = rng.uniform(-8, 8, (20000, 3)).astype(np.float32); o2[:, 1] = rng.uniform(0.0, 2.0, 20000)
    d2 = rng.normal(0, 1, (20000, 3)).astype(np.float32)
    assert_hits_equal(fz.hit(o2, d2, flags=extra), osc.trace_rays(o2, d2))


def build_mixed_scene(api):
    """Every primitive type, including the ones random_scene lacks (infinite plane, marched sphere,
    marched box, scaled torus), plus exact duplicates for the tie-break rule."""
    M, w = api.Material, api.HittableList()
    w += api.InfinitePlane.new((0, -0.5, 0), (0, 1, 0), M.new_lambertian((0.5, 0.5, 0.5)))
    w += api.InfinitePlane.new((0, 0, -9), (0.2, 0, 1), M.new_metal((0.9, 0.9, 0.9)))
    rng = np.random.default_rng(7)
    for i in range(40):
        t, r, s = rng.uniform(-4, 4, 3), rng.uniform(0, 6.28, 3), rng.uniform(0.2, 0.8, 3)
        t[1] = rng.uniform(0, 2)
        m = api.Mat4x4.new_translate(t) ^ api.Mat4x4.new_rotate_x(r[0]) ^ api.Mat4x4.new_rotate_y(r[1]) ^ api.Mat4x4.new_scale(s)
        mat = [M.new_lambertian((0.8, 0.3, 0.3)), M.new_metal_fuzz((0.8, 0.8, 0.8), 0.3), M.new_dielectric(1.5)][i % 3]
        if i % 4 == 0:
            w += api.Cube.new(m, mat)
        else:
            w += api.Sphere.new(m, mat)
        if i % 10 == 0:
            w += api.Sphere.new(m, mat)  # exact duplicate: the later one must win
    for i in range(10):
        p = [rng.uniform(-4, 4, 3) for _ in range(3)]
        w += api.Triangle.new3points(p[0], p[1], p[2], M.new_lambertian((1, 1, 0)))
        w += api.Parallelogram.new3points(p[2], p[0], p[1] + 1.0, M.new_metal((1, 0.5, 1)))
    w += api.MarchedSphere((2, 0.5, 2), 0.6, M.new_dielectric(1.5))
    w += api.MarchedBox((-2, 0.4, 1), (0.5, 0.4, 0.3), M.new_lambertian((0.2, 0.8, 0.2)))
    w += api.MarchedTorus.new(api.Mat4x4.new_translate((0, 1, -2)) ^ api.Mat4x4.new_rotate_x(0.6), (0.5, 0.1, 0.1), M.new_dielectric(1.5))
    w += api.MarchedTorus.new(api.Mat4x4.new_translate((1, 1, 1)) ^ api.Mat4x4.new_rotate_z(1.0) ^ api.Mat4x4.new_scale((1.5, 1.5, 1.5)), (0.4, 0.12, 0), M.new_metal((0.7, 0.6, 0.5)))
    return w


@pytest.fixture(scope="module")
def mixed(api, O):
    fz = build_mixed_scene(api).freeze()
    return fz, oracle_scene_from(O, fz)


@pytest.mark.parametrize("extra", [0, 4, 16])  # bvh+smem, brute force, bvh in global memory
def test_hit_bit_exact_all_primitive_types(api, A, mixed, extra):
    fz, osc = mixed
    rng = np.random.default_rng(3)
    o = rng.uniform(-6, 6, (40000, 3)).astype(np.float32); o[:, 1] = rng.uniform(-0.4, 3, 40000)
    d = rng.normal(0, 1, (40000, 3)).astype(np.float32)
    g, ref = fz.hit(o, d, flags=extra), osc.trace_rays(o, d)
    assert_hits_equal(g, ref)
    types = set((ref[0] >> np.uint64(32)).tolist())
    assert types >= {0, 1, 2, 3, 4, 5, 6, 7, 8}, types  # every type (and the sky) was actually hit


def test_hit_and_render_bit_exact_with_forced_ray_padding(api, A, O):
    """OM_FREEZE_RAY_PADDING on a scene with every primitive type: the scene stays in global memory and every node visit widens
    the child boxes by the margin of the node's bounding ball (SceneHeader::off_node_far).  Padding only decides what gets
    tested, so hits and a render must still equal the oracle's, bit for bit, from near and from far origins."""
    fz = build_mixed_scene(api).freeze(flags=A.FREEZE_RAY_PADDING)
    st = fz.stats()
    assert st["ray_padding"] == 1
    osc = oracle_scene_from(O, fz)
    rng = np.random.default_rng(13)
    o = rng.uniform(-6, 6, (30000, 3)).astype(np.float32); o[:, 1] = rng.uniform(-0.4, 3, 30000)
    d = rng.normal(0, 1, (30000, 3)).astype(np.float32)
    ref = osc.trace_rays(o, d)
    assert_hits_equal(fz.hit(o, d, flags=0), ref)
    assert set((ref[0] >> np.uint64(32)).tolist()) >= {0, 1, 2, 3, 4, 5}
    far = (rng.normal(0, 1, (8000, 3)) * 400).astype(np.float32)  # origins hundreds of units away, aimed at the scene
    tgt = rng.uniform(-5, 5, (8000, 3)).astype(np.float32); tgt[:, 1] = rng.uniform(0, 2, 8000)
    assert_hits_equal(fz.hit(far, tgt - far, 0.001, 5000.0, flags=0), osc.trace_rays(far, tgt - far, 0.001, 5000.0))
    cam = api.Camera.new((6, 3, 7), (0, 0.5, 0), (0, 1, 0), 40.0, 1.5, 0.05, 9.0)
    p = api.render_params(48, 32, 3, seed=5, flags=0)
    r = api.Renderer(fz, cam, p); r.launch(0); r.sync()
    assert_pixels_equal(r.read_pixels(), osc.render(cam, p)["pixels"])
    assert r.perf_counters()["scene_in_smem"] == 0  # a padded scene is never staged: the shared-memory variant has no padding code
    r.close()


def test_hit_edge_cases(api, A, mixed, O):
    fz, osc = mixed
    ids, t, p, n = fz.hit(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32), flags=0)
    assert ids.shape == (0,)
    o, d = rays_at_scene(2000, 5, 4.0)
    for tmin, tmax in [(0.001, 100.0), (0.001, 9.0), (5.0, 12.0), (0.0, 1e-3)]:
        assert_hits_equal(fz.hit(o, d, tmin, tmax, flags=0), osc.trace_rays(o, d, tmin, tmax))
    # axis-parallel directions (zero components -> infinite slab reciprocals) and a ray inside a box
    o2 = np.array([[0, 0.3, 8], [0, 5, 0], [-8, 0.4, 1], [-2, 0.4, 1]], np.float32)
    d2 = np.array([[0, 0, -1], [0, -1, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    assert_hits_equal(fz.hit(o2, d2, flags=0), osc.trace_rays(o2, d2))
    # empty scene: everything is sky
    empty = api.HittableList().freeze()
    ids, t, _, _ = empty.hit(o, d, flags=0)
    assert (ids == 0).all() and np.isinf(t).all()


def test_hit_bit_exact_many_spheres_global_bvh(api, A, O):
    fz = api.FrozenHittableList.generate("many_spheres", 5, 20000)
    assert fz.stats()["fits_smem"] == 0
    osc = oracle_scene_from(O, fz)
    rng = np.random.default_rng(4)
    o = np.tile(np.array([60, 12, 14], np.float32), (4000, 1))
    tgt = rng.uniform(-40, 40, (4000, 3)).astype(np.float32); tgt[:, 1] = rng.uniform(0, 4, 4000)
    assert_hits_equal(fz.hit(o, tgt - o, 0.001, 1000.0, flags=0), osc.trace_rays(o, tgt - o, 0.001, 1000.0))


@pytest.mark.parametrize("freeze_flags", [0, 2, 4])  # chosen from the scene / per-ray box padding / worst-case static padding
def test_hit_bit_exact_ill_conditioned_tiny_spheres(api, A, O, freeze_flags):
    """Sphere::hit (traced.rs:38-56) from an origin thousands of radii away rounds `|o_l|^2 - 1` and the discriminant so
    coarsely that rays passing well outside a sphere are reported as hits; the culling has to keep every one of them."""
    fz = api.FrozenHittableList.generate("many_spheres", 5, 20000, flags=freeze_flags)
    st = fz.stats()
    assert st["ray_padding"] == (0 if freeze_flags == 4 else 1)
    osc = oracle_scene_from(O, fz)
    sph = np.frombuffer(fz.lists()["spheres"], dtype=np.float32).reshape(-1, ctypes.sizeof(A.Sphere) // 4)
    ctr = sph[1:, [3, 7, 11]]  # (sphere 0 is the ground)
    rad = np.abs(sph[1:, 0])
    rng = np.random.default_rng(9)
    n = 6000
    pick = rng.integers(0, ctr.shape[0], n)
    u = rng.normal(size=(n, 3)); u /= np.linalg.norm(u, axis=1, keepdims=True); u[:, 1] = np.abs(u[:, 1]) + 0.05
    v = np.cross(u, rng.normal(size=(n, 3))); v /= np.linalg.norm(v, axis=1, keepdims=True)
    dist = rng.uniform(150, 420, (n, 1))
    miss = rng.uniform(0.2, 5.0, (n, 1))  # impact parameter in radii
    o = (ctr[pick] + u * dist).astype(np.float32)
    tgt = ctr[pick] + v * miss * rad[pick, None]
    d = ((tgt - o) / dist * rng.uniform(0.5, 2.0, (n, 1))).astype(np.float32)
    want = osc.trace_rays(o, d, 0.001, 1000.0)
    assert_hits_equal(fz.hit(o, d, 0.001, 1000.0, flags=0), want)
    assert_hits_equal(fz.hit(o, d, 0.001, 1000.0, flags=A.RENDER_BRUTE_FORCE), want)
    # the case is really exercised: the aimed-at sphere is reported hit by rays that geometrically pass outside it
    aimed = want[0] == ((np.uint64(1) << np.uint64(32)) | (pick + 2).astype(np.uint64))  # OM_OBJ_ID(sphere, 1 + pick)
    assert (aimed & (miss[:, 0] > 1.5)).sum() > 20


def test_fast_math_build_hit_tolerance(api, A, rtiow, O):
    """OM_RENDER_FAST_MATH (FMA contraction everywhere) is NOT the parity path; this test states how far
    it is from the oracle.  Cube / triangle / parallelogram: t within 1e-5 relative.  Sphere::hit
    (traced.rs:38-56) is ill-conditioned by construction (`|o_l|^2 - 1` with the origin ~50 radii
    away), so ANY change of rounding moves a grazing root by up to ~1e-3 relative: median and p99
    are bounded here, and the 1e-5 bound of north_star is met by the default build (bit-exact).
    Marched hits stop anywhere inside the HIT_SIZE band (hits.rs:113,325), so their bound is the band."""
    fz, osc = rtiow
    o, d = rays_at_scene(60000, 6)
    gi, gt, gp, gn = fz.hit(o, d, flags=A.RENDER_FAST_MATH)
    oi, ot, op, on = osc.trace_rays(o, d)
    assert (gi == oi).mean() > 0.9995
    same = (gi == oi) & (oi != 0)
    typ = (oi >> np.uint64(32)).astype(np.int64)
    flat = same & (typ >= 2) & (typ <= 5)
    rel = np.abs(gt[flat] - ot[flat]) / np.abs(ot[flat])
    assert rel.max() <= 1e-5
    sph = same & (typ == 1)
    rel = np.abs(gt[sph] - ot[sph]) / np.abs(ot[sph])
    print("fast-math build: spheres rel t err median %.3g p99 %.3g max %.3g, ids equal %.5f" % (np.median(rel), np.quantile(rel, 0.99), rel.max(), (gi == oi).mean()))
    assert np.median(rel) <= 1e-5 and np.quantile(rel, 0.99) <= 2e-4 and rel.max() <= 5e-3
    marched = same & (typ >= 6)
    assert marched.sum() > 100
    # the GPU's marched hit point lies within the reference's hit band of the same object
    L = O.lib()
    torus = fz.lists()["marched_tori"][0]
    import ctypes as C
    sdf = np.array([abs(L.omo_marched_sdf(2, C.byref(torus), (C.c_float * 3)(*gp[i]))) for i in np.nonzero(marched)[0][:2000]])
    assert sdf.max() < 1.05e-3


# ---- render(): full Stats ------------------------------------------------------------------------
@pytest.mark.parametrize("W,H,spp,depth", [(64, 48, 6, 50), (37, 23, 5, 50), (2, 2, 3, 50), (48, 32, 4, 1), (48, 32, 4, 0), (40, 30, 3, 3)])
def test_render_bit_exact_fixed_spp(api, A, rtiow, W, H, spp, depth):
    fz, osc = rtiow
    cam = cam_rtiow(api, W / H)
    p = api.render_params(W, H, spp, max_depth=depth, seed=21, flags=0)
    r = api.Renderer(fz, cam, p)
    r.launch(0); r.sync()
    g = r.read_pixels()
    ref = osc.render(cam, p)
    assert_pixels_equal(g, ref["pixels"])
    assert r.samples_done() == ref["credited"] == W * H * spp and r.samples_traced() == ref["traced"]
    assert (r.read_rgb().reshape(-1, 3) == g["color"]).all()
    r.close()


@pytest.mark.parametrize("W,H", [(1, 1), (1, 6), (7, 1), (0, 4)])
def test_render_rejects_one_pixel_wide_images(api, A, rtiow, W, H):
    """u = (col + jitter) / (image_width - 1) (render_thread.rs:183-186) divides by zero on a 1-pixel-wide image: the
    reference renders NaN rays there; the library refuses the size instead of reproducing NaN payloads."""
    fz, _ = rtiow
    with pytest.raises(Exception) as e:
        api.Renderer(fz, cam_rtiow(api, 1.5), api.render_params(W, H, 4, seed=8))
    assert "at least 2x2" in str(e.value)


def test_render_bit_exact_adaptive_multipass(api, A, rtiow):
    """The reference's only mode: retire a pixel after 5 unchanged 8-bit colours; sample-major passes."""
    fz, osc = rtiow
    cam = cam_rtiow(api, 4 / 3)
    p = api.render_params(64, 48, 40, seed=8, flags=A.RENDER_ADAPTIVE)
    ref = osc.render(cam, p)
    r = api.Renderer(fz, cam, p)
    for k in (1, 7, 12, 0):  # uneven passes; 0 = all that is left
        r.launch(k)
    r.sync()
    g = r.read_pixels()
    assert_pixels_equal(g, ref["pixels"])
    assert r.samples_done() == ref["credited"] == 64 * 48 * 40
    assert r.samples_traced() == ref["traced"] < 64 * 48 * 40
    assert ((g["bad_avgs"] >= 5) | (g["n"] == 40)).all()
    r.launch(5); r.sync()  # nothing left to do: a further pass changes nothing
    assert r.read_pixels().tobytes() == g.tobytes()
    r.close()


@pytest.mark.parametrize("kind,seed,pass_spp", [("glass_heavy", 4, 16), ("random_scene", 1, 3), ("random_scene", 1, 40), ("marched", 3, 8)])
def test_adaptive_spread_passes_bit_exact(api, A, O, kind, seed, pass_spp, monkeypatch):
    """Adaptive passes after the first walk a compacted live list; with few live pixels they run on k_render_spread (2..32 lanes
    per pixel, one sample each, folded in sample order).  Pixels, credited and traced samples must equal the oracle's sequential
    adaptive render - also when a pass is not a whole number of rounds (pass_spp 3 and 40) - and the k_render-only path's."""
    fz = api.FrozenHittableList.generate(kind, seed)
    osc = oracle_scene_from(O, fz)
    cam = cam_rtiow(api, 4 / 3)
    W, H, spp = 72, 54, 50
    p = api.render_params(W, H, spp, seed=31, flags=A.RENDER_ADAPTIVE)
    ref = osc.render(cam, p)

    def render():
        r = api.Renderer(fz, cam, p)
        for _ in range((spp + pass_spp - 1) // pass_spp):
            r.launch(pass_spp)
        r.sync()
        out = (r.read_pixels().copy(), r.samples_done(), r.samples_traced(), r.perf_counters()["kernels"])
        r.close()
        return out

    g, credited, traced, kernels = render()
    assert_pixels_equal(g, ref["pixels"])
    assert credited == ref["credited"] == W * H * spp and traced == ref["traced"] < W * H * spp
    monkeypatch.setenv("OM_NO_SPREAD", "1")
    g2, credited2, traced2, kernels2 = render()
    assert g2.tobytes() == g.tobytes() and (credited2, traced2) == (credited, traced)
    assert kernels > kernels2  # the spread kernel really was launched beside k_render


def test_render_bit_exact_marched_and_mixed(api, A, mixed, O):
    fz, osc = mixed
    cam = api.Camera.new((6, 3, 7), (0, 0.5, 0), (0, 1, 0), 40.0, 1.5, 0.05, 9.0)
    p = api.render_params(48, 32, 4, seed=2, flags=0)
    r = api.Renderer(fz, cam, p); r.launch(0); r.sync()
    assert_pixels_equal(r.read_pixels(), osc.render(cam, p)["pixels"])
    r.close()
    fzm = api.FrozenHittableList.generate("marched", 3)
    oscm = oracle_scene_from(O, fzm)
    p = api.render_params(32, 24, 2, seed=2, flags=0)
    r = api.Renderer(fzm, cam_rtiow(api, 4 / 3), p); r.launch(0); r.sync()
    assert_pixels_equal(r.read_pixels(), oscm.render(cam_rtiow(api, 4 / 3), p)["pixels"])
    r.close()


def test_render_partitions_bit_exact(api, A, rtiow):
    """spp slices and interleaved 2730-pixel chunks (main.rs:172-189) against the oracle's same split."""
    fz, osc = rtiow
    W, H, spp = 120, 80, 9  # 9600 pixels = 4 chunks
    cam = cam_rtiow(api, 1.5)
    for kw in [dict(sample_base=3, sample_count=4), dict(sample_base=7), dict(part_rank=1, part_world=3), dict(part_rank=0, part_world=2)]:
        p = api.render_params(W, H, spp, seed=9, flags=0, **kw)
        r = api.Renderer(fz, cam, p); r.launch(0); r.sync()
        g = r.read_pixels()
        assert_pixels_equal(g, osc.render(cam, p)["pixels"])
        if "part_world" in kw:
            from raytracingoneweekend_b200 import dist as D
            own = D.owned_pixel_mask(W * H, kw["part_rank"], kw["part_world"])
            assert (g["n"][own] == spp).all() and (g["n"][~own] == 0).all()
        r.close()


def test_golden_fixtures(api, A, rtiow):
    """Committed oracle outputs (tools/make_golden.py) — guards oracle and kernels against drifting together."""
    fz, _ = rtiow
    gold = np.load(os.path.join(GOLD, "rtiow_exact_64x48.npz"))
    cam = cam_rtiow(api, 64 / 48)
    r = api.Renderer(fz, cam, api.render_params(64, 48, 12, seed=11, flags=0)); r.launch(0); r.sync()
    assert_pixels_equal(r.read_pixels(), gold["fixed"]); r.close()
    r = api.Renderer(fz, cam, api.render_params(64, 48, 48, seed=11, flags=A.RENDER_ADAPTIVE)); r.launch(0); r.sync()
    assert_pixels_equal(r.read_pixels(), gold["adaptive"])
    assert r.samples_done() == int(gold["adaptive_credited"]) and r.samples_traced() == int(gold["adaptive_traced"]); r.close()
    gp = np.load(os.path.join(GOLD, "rtiow_primary_240x160.npz"))
    r = api.Renderer(fz, cam_rtiow(api, 1.5), api.render_params(240, 160, 8, seed=5, flags=0))
    ids, t = r.primary(0, 0)
    assert (ids == gp["ids_center"]).all() and (bits(t) == bits(gp["t_center"])).all()
    ids, t = r.primary(1, 3)
    assert (ids == gp["ids_sample3"]).all() and (bits(t) == bits(gp["t_sample3"])).all()
    r.close()
    # fast-math build on the same maps: same winner except at silhouettes; t as in test_fast_math_build_hit_tolerance
    r = api.Renderer(fz, cam_rtiow(api, 1.5), api.render_params(240, 160, 8, seed=5, flags=A.RENDER_FAST_MATH))
    ids, t = r.primary(0, 0)
    same = ids == gp["ids_center"]
    assert same.mean() >= 0.9995
    tr = same & (ids != 0) & ((ids >> np.uint64(32)) <= 5)
    rel = np.abs(t[tr] - gp["t_center"][tr]) / gp["t_center"][tr]
    assert np.median(rel) <= 1e-5 and rel.max() <= 5e-3
    r.close()


# ---- wavefront pipeline (OM_RENDER_WAVEFRONT): same bits as the oracle and as the persistent kernel ----
@pytest.mark.parametrize("W,H,spp,depth,flags_extra", [(64, 48, 6, 50, 0), (37, 23, 5, 50, 0), (2, 2, 3, 50, 0), (48, 32, 4, 1, 0), (40, 30, 3, 3, 0),
                                                       (64, 48, 6, 50, 4), (64, 48, 6, 50, 16)])  # 4 = brute force, 16 = scene in global memory
def test_wavefront_bit_exact_fixed_spp(api, A, rtiow, W, H, spp, depth, flags_extra):
    fz, osc = rtiow
    cam = cam_rtiow(api, W / H)
    p = api.render_params(W, H, spp, max_depth=depth, seed=21, flags=A.RENDER_WAVEFRONT | flags_extra)
    r = api.Renderer(fz, cam, p)
    r.launch(0); r.sync()
    g = r.read_pixels()
    ref = osc.render(cam, api.render_params(W, H, spp, max_depth=depth, seed=21, flags=0))
    assert_pixels_equal(g, ref["pixels"])
    assert r.samples_done() == ref["credited"] == W * H * spp and r.samples_traced() == ref["traced"]
    r.close()


def test_wavefront_bit_exact_adaptive_multipass_and_marched(api, A, rtiow, mixed, O):
    fz, osc = rtiow
    cam = cam_rtiow(api, 1.5)
    p = api.render_params(60, 40, 40, seed=4, flags=A.RENDER_ADAPTIVE | A.RENDER_WAVEFRONT)
    r = api.Renderer(fz, cam, p)
    for _ in range(5):
        r.launch(8)
    r.sync()
    ref = osc.render(cam, api.render_params(60, 40, 40, seed=4, flags=A.RENDER_ADAPTIVE))
    assert_pixels_equal(r.read_pixels(), ref["pixels"])
    assert r.samples_done() == ref["credited"] and r.samples_traced() == ref["traced"]
    r.close()
    fzx, oscx = mixed
    camx = api.Camera.new((6, 3, 7), (0, 0.5, 0), (0, 1, 0), 40.0, 1.5, 0.05, 9.0)
    p = api.render_params(48, 32, 4, seed=2, flags=A.RENDER_WAVEFRONT)
    r = api.Renderer(fzx, camx, p); r.launch(0); r.sync()
    assert_pixels_equal(r.read_pixels(), oscx.render(camx, api.render_params(48, 32, 4, seed=2))["pixels"])
    r.close()
    fzm = api.FrozenHittableList.generate("marched", 3)
    oscm = oracle_scene_from(O, fzm)
    p = api.render_params(32, 24, 2, seed=2, flags=A.RENDER_WAVEFRONT)
    r = api.Renderer(fzm, cam_rtiow(api, 4 / 3), p); r.launch(0); r.sync()
    assert_pixels_equal(r.read_pixels(), oscm.render(cam_rtiow(api, 4 / 3), api.render_params(32, 24, 2, seed=2))["pixels"])
    r.close()


# ---- CTA-resident wavefront (OM_RENDER_POOL, csrc/om_pool.cuh): same bits as the oracle ----
@pytest.mark.parametrize("W,H,spp,depth,flags_extra", [(64, 48, 6, 50, 0), (37, 23, 5, 50, 0), (2, 2, 3, 50, 0), (48, 32, 4, 1, 0), (40, 30, 3, 3, 0), (64, 48, 6, 50, 4)])
def test_pool_bit_exact_fixed_spp(api, A, rtiow, W, H, spp, depth, flags_extra):
    fz, osc = rtiow
    cam = cam_rtiow(api, W / H)
    r = api.Renderer(fz, cam, api.render_params(W, H, spp, max_depth=depth, seed=21, flags=A.RENDER_POOL | flags_extra))
    r.launch(0); r.sync()
    ref = osc.render(cam, api.render_params(W, H, spp, max_depth=depth, seed=21, flags=0))
    assert_pixels_equal(r.read_pixels(), ref["pixels"])
    assert r.samples_done() == ref["credited"] == W * H * spp and r.samples_traced() == ref["traced"]
    r.close()


def test_pool_bit_exact_adaptive_multipass_marched_and_limits(api, A, rtiow, mixed, O):
    fz, osc = rtiow
    cam = cam_rtiow(api, 1.5)
    r = api.Renderer(fz, cam, api.render_params(60, 40, 40, seed=4, flags=A.RENDER_ADAPTIVE | A.RENDER_POOL))
    for _ in range(5):
        r.launch(8)
    r.sync()
    ref = osc.render(cam, api.render_params(60, 40, 40, seed=4, flags=A.RENDER_ADAPTIVE))
    assert_pixels_equal(r.read_pixels(), ref["pixels"])
    assert r.samples_done() == ref["credited"] and r.samples_traced() == ref["traced"]
    r.close()
    fzx, oscx = mixed   # every primitive type, marched objects in front of traced ones
    camx = api.Camera.new((6, 3, 7), (0, 0.5, 0), (0, 1, 0), 40.0, 1.5, 0.05, 9.0)
    r = api.Renderer(fzx, camx, api.render_params(48, 32, 4, seed=2, flags=A.RENDER_POOL)); r.launch(0); r.sync()
    assert_pixels_equal(r.read_pixels(), oscx.render(camx, api.render_params(48, 32, 4, seed=2))["pixels"])
    r.close()
    fzm = api.FrozenHittableList.generate("marched", 3)   # 65 marched objects: marches span many phases
    oscm = oracle_scene_from(O, fzm)
    r = api.Renderer(fzm, cam_rtiow(api, 4 / 3), api.render_params(32, 24, 2, seed=2, flags=A.RENDER_POOL)); r.launch(0); r.sync()
    assert_pixels_equal(r.read_pixels(), oscm.render(cam_rtiow(api, 4 / 3), api.render_params(32, 24, 2, seed=2))["pixels"])
    r.close()
    # what does not fit shared memory is refused, not silently rendered by something else
    big = api.FrozenHittableList.generate("many_spheres", 5, 20000)
    for bad in (dict(flags=A.RENDER_POOL), ):
        with pytest.raises(api.OmError) as e:
            api.Renderer(big, cam, api.render_params(16, 16, 1, **bad))
        assert e.value.status == A.OM_ERR_INVALID
    with pytest.raises(api.OmError) as e:
        api.Renderer(fz, cam, api.render_params(16, 16, 1, flags=A.RENDER_POOL | A.RENDER_WAVEFRONT))
    assert e.value.status == A.OM_ERR_INVALID


@pytest.mark.parametrize("kind,seed", [("random_scene", 1), ("glass_heavy", 4), ("marched", 3)])
def test_wavefront_equals_persistent_kernel_large(api, A, kind, seed):
    """At a size the oracle cannot reach in seconds (more pixels than slots are exercised by the
    pixel hand-over and the tail): both device paths must agree bit for bit."""
    fz = api.FrozenHittableList.generate(kind, seed)
    cam = cam_rtiow(api, 16 / 9)
    out = []
    for fl in (A.RENDER_PERSISTENT, A.RENDER_WAVEFRONT):
        r = api.Renderer(fz, cam, api.render_params(640, 360, 6, seed=9, flags=fl))
        r.launch(4); r.launch(0); r.sync()
        out.append((r.read_pixels(), r.samples_done(), r.samples_traced()))
        r.close()
    assert_pixels_equal(out[0][0], out[1][0])
    assert out[0][1:] == out[1][1:]


@pytest.mark.parametrize("kind,seed,adaptive", [("random_scene", 1, False), ("random_scene", 1, True), ("marched", 3, False), ("rtiow_classic", 1, False)])
def test_phase_alignment_changes_nothing(api, A, kind, seed, adaptive):
    """OM_RENDER_PHASE_ALIGN only makes the warps of an SM sub-partition wait for each other (named barriers in
    k_render); forced on, forced off and the library's own choice must give the same pixels and counters, also when
    warps run out of tickets at different times (small image, adaptive retirement, a partition)."""
    fz = api.FrozenHittableList.generate(kind, seed)
    cam = cam_rtiow(api, 16 / 9)
    base = A.RENDER_PERSISTENT | (A.RENDER_ADAPTIVE if adaptive else 0)
    out = []
    for fl in (A.RENDER_NO_PHASE_ALIGN, A.RENDER_PHASE_ALIGN, 0):
        r = api.Renderer(fz, cam, api.render_params(320, 180, 12 if not adaptive else 40, seed=5, flags=base | fl))
        r.launch(5); r.launch(0); r.sync()
        out.append((r.read_pixels(), r.samples_done(), r.samples_traced()))
        r.close()
    for o in out[1:]:
        assert_pixels_equal(out[0][0], o[0])
        assert out[0][1:] == o[1:]


@pytest.mark.parametrize("kind,seed", [("random_scene", 1), ("rtiow_classic", 1)])
def test_cost_ordered_tile_scheduling_changes_nothing(api, A, kind, seed, monkeypatch):
    """A fixed-spp launch of 32 samples or more is split into a pilot launch (1/64 of the samples, files every tile's cost)
    and the rest with the 16x16-pixel blocks handed out most expensive first (om_render_launch / refresh_block_order).  Same
    pixels and counters as the plain row order (OM_NO_BLOCK_ORDER), as a later launch of the same handle, and as passes."""
    fz = api.FrozenHittableList.generate(kind, seed)
    cam = cam_rtiow(api, 4 / 3)
    W, H, spp = 800, 600, 40  # 50 x 38 blocks >= 8 per SM
    out = []
    for mode in ("ordered", "rows", "passes"):
        if mode == "rows":
            monkeypatch.setenv("OM_NO_BLOCK_ORDER", "1")
        else:
            monkeypatch.delenv("OM_NO_BLOCK_ORDER", raising=False)
        r = api.Renderer(fz, cam, api.render_params(W, H, spp, seed=4, flags=A.RENDER_PERSISTENT))
        if mode == "passes":
            r.launch(33); r.launch(0)  # pilot inside the first launch, the order reused by the second
        else:
            r.launch(0)
        r.sync()
        px, kernels = r.read_pixels(), r.perf_counters()["kernels"]
        assert r.samples_traced() == W * H * spp and r.samples_done() == W * H * spp
        assert kernels == {"ordered": 2, "rows": 1, "passes": 3}[mode]
        if mode == "ordered":  # the same handle again, now without a pilot
            r.reset(); r.launch(0); r.sync()
            assert_pixels_equal(r.read_pixels(), px)
            assert r.perf_counters()["kernels"] == 3
        r.close()
        out.append(px)
    assert_pixels_equal(out[0], out[1])
    assert_pixels_equal(out[0], out[2])


def test_pipeline_flags_exclusive(api, A, rtiow):
    fz, _ = rtiow
    with pytest.raises(Exception):
        api.Renderer(fz, cam_rtiow(api), api.render_params(16, 16, 1, flags=A.RENDER_PERSISTENT | A.RENDER_WAVEFRONT))


# ---- the viewer's display modes and 3x3 weighted filters (main.rs:219-343, 361-437) on the device ----
@pytest.mark.parametrize("W,H,adaptive", [(64, 48, False), (37, 23, True), (3, 3, False)])
def test_view_modes_bit_exact(api, A, O, rtiow, W, H, adaptive, tmp_path):
    fz, _ = rtiow
    cam = cam_rtiow(api, W / H)
    r = api.Renderer(fz, cam, api.render_params(W, H, 24, seed=3, flags=A.RENDER_ADAPTIVE if adaptive else 0))
    r.launch(0); r.sync()
    px = r.read_pixels()
    for mode in range(7):
        g = r.read_view(mode)
        assert (g == O.view(px, W, H, mode)).all(), "display mode %d" % mode
    assert (r.read_view(0) == r.read_rgb()).all()
    api.write_bmp(tmp_path / "f12.bmp", r.read_view(0))
    raw = (tmp_path / "f12.bmp").read_bytes()
    assert raw[:2] == b"BM" and len(raw) == 54 + ((3 * W + 3) // 4 * 4) * H
    # bottom-up BGR rows: the first stored row is the last image row
    assert raw[54:57] == bytes(r.read_view(0)[H - 1, 0, ::-1])
    r.close()


def test_view_rejects_bad_arguments(api, A, rtiow):
    fz, _ = rtiow
    r = api.Renderer(fz, cam_rtiow(api, 1.0), api.render_params(2, 2, 1))
    r.launch(0); r.sync()
    with pytest.raises(Exception):
        r.read_view(7)
    with pytest.raises(Exception):
        r.read_view(2)  # 3x3 filter on a 2x2 image
    assert r.read_view(0).shape == (2, 2, 3)
    r.close()


def test_marched_candidate_grid_hit_bit_exact(api, A, O):
    """65 marched objects: min_j |sdf_j| is taken over the candidate list of the point's grid cell
    (om_scene.cpp build_marched_grid); ids, t, points and normals must stay those of the full loop."""
    fz = api.FrozenHittableList.generate("marched", 3)
    osc = oracle_scene_from(O, fz)
    o, d = rays_at_scene(6000, 5, spread=9.0)
    assert_hits_equal(fz.hit(o, d), osc.trace_rays(o, d))
    rng = np.random.default_rng(6)   # rays born between the objects, every direction (also leaving the grid)
    o2 = rng.uniform(-9, 9, (6000, 3)).astype(np.float32); o2[:, 1] = rng.uniform(0.0, 1.6, 6000)
    d2 = rng.normal(0, 1, (6000, 3)).astype(np.float32)
    assert_hits_equal(fz.hit(o2, d2), osc.trace_rays(o2, d2))
    assert_hits_equal(fz.hit(o2, d2, flags=A.RENDER_NO_SMEM_SCENE), osc.trace_rays(o2, d2))


@pytest.mark.parametrize("freeze_flags", [0, 8])  # with / without the distance BVH over the marched objects
def test_marched_distance_bvh_hit_bit_exact(api, A, O, freeze_flags):
    """Outside the candidate grid min_j |sdf_j| comes from a BVH walk bounded by the exact minimum so far
    (nearest_marched); rays that start or travel far from the objects must give the full loop's ids, t, points, normals."""
    fz = api.FrozenHittableList.generate("marched", 3, flags=freeze_flags)
    osc = oracle_scene_from(O, fz)
    rng = np.random.default_rng(16)
    n = 6000
    o = rng.uniform(-60, 60, (n, 3)).astype(np.float32); o[:, 1] = rng.uniform(0.05, 40.0, n)   # mostly outside the grid
    tgt = rng.uniform(-9, 9, (n, 3)).astype(np.float32); tgt[:, 1] = rng.uniform(-0.5, 1.5, n)
    d = (tgt - o).astype(np.float32)
    d[n // 2:] = rng.normal(0, 1, (n - n // 2, 3)).astype(np.float32)  # half of them aimed anywhere, sky included
    want = osc.trace_rays(o, d, 0.001, 100.0)
    assert_hits_equal(fz.hit(o, d, 0.001, 100.0), want)
    assert (want[0] != 0).sum() > n // 4
    # a scene whose objects tie exactly: two identical marched spheres, the first of the scan order must win
    w = api.HittableList()
    m = api.Material.new_lambertian((0.5, 0.5, 0.5))
    for k in range(8):
        w += api.MarchedSphere((float(k % 4) * 3.0, 1.0, float(k // 4) * 3.0), 1.0, m)
        w += api.MarchedSphere((float(k % 4) * 3.0, 1.0, float(k // 4) * 3.0), 1.0, m)
    fz2 = w.freeze(flags=freeze_flags)
    osc2 = oracle_scene_from(O, fz2)
    o3 = rng.uniform(-30, 30, (2000, 3)).astype(np.float32); o3[:, 1] = rng.uniform(3, 30, 2000)
    t3 = rng.uniform(0, 9, (2000, 3)).astype(np.float32); t3[:, 1] = 1.0
    want3 = osc2.trace_rays(o3, t3 - o3, 0.001, 100.0)
    assert_hits_equal(fz2.hit(o3, t3 - o3, 0.001, 100.0), want3)
    hit_ids = want3[0][want3[0] != 0]
    assert hit_ids.size > 500 and ((hit_ids & np.uint64(0xFFFFFFFF)) % np.uint64(2) == 1).all()  # index+1 odd = the first of each pair


def test_gpu_bvh_build_same_results(api, A, O, rtiow):
    """OM_FREEZE_GPU_BVH: Morton/Karras tree built on the device instead of the host's binned SAH; the tree only culls,
    so hits and images must not change by a bit."""
    fz = api.FrozenHittableList.generate("many_spheres", 5, 20000, flags=A.FREEZE_GPU_BVH)
    st = fz.stats()
    assert (st["bvh_prims"] - 1) / 3 <= st["bvh_nodes"] <= st["bvh_prims"] - 1 and 5 < st["bvh_depth"] <= 60  # four children per node
    osc = oracle_scene_from(O, fz)
    rng = np.random.default_rng(4)
    o = np.tile(np.array([60, 12, 14], np.float32), (4000, 1))
    tgt = rng.uniform(-40, 40, (4000, 3)).astype(np.float32); tgt[:, 1] = rng.uniform(0, 4, 4000)
    assert_hits_equal(fz.hit(o, tgt - o, 0.001, 1000.0, flags=0), osc.trace_rays(o, tgt - o, 0.001, 1000.0))
    # every primitive type, and both pipelines, on the RTIOW scene
    g = api.FrozenHittableList.generate("random_scene", 1, flags=A.FREEZE_GPU_BVH)
    cam = cam_rtiow(api)
    want = None
    for scene in (rtiow[0], g):
        for flags in (A.RENDER_PERSISTENT, A.RENDER_WAVEFRONT):
            r = api.Renderer(scene, cam, api.render_params(240, 160, 8, seed=21, flags=flags))
            r.launch(0); px = r.read_pixels(); r.close()
            if want is None:
                want = px
            else:
                assert_pixels_equal(px, want)
    # duplicate centres (equal Morton keys are told apart by position)
    w = api.HittableList()
    m = api.Material.new_lambertian((0.5, 0.5, 0.5))
    for k in range(40):
        w += api.Sphere.new_with_radius((float(k % 5), 0.0, -3.0), 0.3 + 0.01 * (k // 5), m)
    d1, d0 = w.freeze(flags=A.FREEZE_GPU_BVH), w.freeze()
    o2, dd2 = rays_at_scene(3000, 8)
    o2[:] = (2.0, 0.5, 4.0)
    assert_hits_equal(d1.hit(o2, dd2), d0.hit(o2, dd2))


@pytest.mark.parametrize("freeze_flags", [0, 16])  # host build (threaded above 65536 objects) / device build
def test_large_bvh_equals_linear_scan(api, A, freeze_flags):
    fz = api.FrozenHittableList.generate("many_spheres", 7, 100000, flags=freeze_flags)
    st = fz.stats()
    assert (st["bvh_prims"] - 1) / 3 <= st["bvh_nodes"] <= st["bvh_prims"] - 1 and st["bvh_depth"] <= 60  # four children per node
    rng = np.random.default_rng(14)
    n = 3000
    o = rng.uniform(-90, 90, (n, 3)).astype(np.float32); o[:, 1] = rng.uniform(0.2, 12.0, n)
    tgt = rng.uniform(-90, 90, (n, 3)).astype(np.float32); tgt[:, 1] = rng.uniform(0, 4, n)
    got = fz.hit(o, tgt - o, 0.001, 100.0)
    assert_hits_equal(got, fz.hit(o, tgt - o, 0.001, 100.0, flags=A.RENDER_BRUTE_FORCE))
    assert (got[0] != 0).sum() > n // 5


@pytest.mark.parametrize("world,rebalance", [(2, True), (3, True), (2, False)])
def test_adaptive_pixels_redealt_between_handles_bit_exact(api, A, rtiow, world, rebalance):
    """dist.AdaptiveScheduler's protocol with the ranks emulated by `world` handles on one GPU: chunk owner table ->
    pass -> export owned states -> integer sum -> import everywhere -> live pixels per chunk -> next deal.  The frame and
    the credited total must be those of the oracle's adaptive render of the same frame."""
    import torch
    from raytracingoneweekend_b200 import dist as D
    fz, osc = rtiow
    W, H, spp, pass_spp = 200, 130, 40, 8
    cam = cam_rtiow(api, W / H)
    p = api.render_params(W, H, spp, seed=31, flags=A.RENDER_ADAPTIVE)
    ref = osc.render(cam, p)   # the ORACLE's adaptive render (render_thread.rs:176-200), not another GPU handle
    want, want_credited = ref["pixels"], ref["credited"]
    assert (want["n"] < spp).any() and (want["n"] == spp).any()  # some pixels retired early, some did not
    rs = [api.Renderer(fz, cam, p) for _ in range(world)]
    states = [torch.empty((W * H, 8), dtype=torch.int32, device="cuda") for _ in range(world)]
    owners = (np.arange(rs[0].n_chunks) % world).astype(np.uint8)
    live = rs[0].live_per_chunk()
    assert live.sum() == W * H
    passes = 0
    while passes * pass_spp < spp and live.sum() > 0:
        if rebalance:
            owners, load = D.deal_chunks(live, world)
            assert load.max() - load.min() <= A.CHUNK_PIXELS
        for k, r in enumerate(rs):
            r.set_chunk_owners(owners, k)
            r.launch(pass_spp)
            r.export_state(states[k].data_ptr(), only_owned=True)
        merged = states[0].clone()
        for k in range(1, world):
            merged += states[k]
        torch.cuda.synchronize()
        for r in rs:
            r.import_state(merged.data_ptr())
        live = rs[0].live_per_chunk()
        assert all((r.live_per_chunk() == live).all() for r in rs[1:])
        passes += 1
    credited = sum(r.samples_done() for r in rs)
    for r in rs:
        assert_pixels_equal(r.read_pixels(), want)
    assert credited == want_credited == W * H * spp
    # leaving scheduled mode: the static rule is back and more passes change nothing on a finished frame
    rs[0].set_chunk_owners(None, 0)
    rs[0].launch(pass_spp)
    assert_pixels_equal(rs[0].read_pixels(), want)
    with pytest.raises(Exception):
        rs[1].set_chunk_owners(np.zeros(3, np.uint8), 0)  # wrong table length
    for r in rs:
        r.close()


def test_adaptive_scheduler_single_rank_equals_plain_render(api, A, rtiow):
    from raytracingoneweekend_b200 import dist as D
    fz, _ = rtiow
    W, H, spp = 160, 100, 24
    cam = cam_rtiow(api, W / H)
    s = D.AdaptiveScheduler(fz, cam, W, H, spp, 0, 1, pass_spp=8, seed=77)
    seen = []
    credited = s.run(on_pass=lambda sch: seen.append(int(sch.r.read_pixels()["n"].sum())))
    got = s.r.read_pixels()
    s.close()
    r = api.Renderer(fz, cam, api.render_params(W, H, spp, seed=77, flags=A.RENDER_ADAPTIVE))
    r.launch(0)
    want = r.read_pixels()
    r.close()
    assert_pixels_equal(got, want)
    assert credited == W * H * spp and len(seen) == 3 and seen == sorted(seen)
