"""`Arc<dyn Marched>` as data (hits.rs:40,96-100,312-320; marched.rs:8-45): om_scene_add_dyn_marched.
CPU: the program checker, the C ABI round trip, and the oracle's interpreter pinned against the oracle's own restatement
of the reference's three SDFs.  GPU: closest hits and rendered pixels bit-identical to the oracle."""
import ctypes as C

import numpy as np
import pytest

from conftest import bits, oracle_scene_from
from dyn_shapes import box_program, build_dyn_scene, dyn_from_matrix, sphere_program, torus_program


def cam(api, aspect=4 / 3):
    return api.Camera.new((7, 3, 9), (0, 0.7, 1.5), (0, 1, 0), 30.0, aspect, 0.05, 10.0)


def rays(n, seed):
    rng = np.random.default_rng(seed)
    o = np.tile(np.array([7, 3, 9], np.float32), (n, 1)) + rng.normal(0, 0.3, (n, 3)).astype(np.float32)
    tgt = rng.uniform(-3, 3, (n, 3)).astype(np.float32)
    tgt[:, 1] = rng.uniform(0, 1.6, n)
    tgt[:, 2] = rng.uniform(-0.5, 3.8, n)
    return o, tgt - o


def assert_hits_equal(g, o, ids=True):
    if ids:
        assert (g[0] == o[0]).all()
    for k in (1, 2, 3):
        assert (bits(np.ascontiguousarray(g[k])) == bits(np.ascontiguousarray(o[k]))).all(), k


# ---- CPU -----------------------------------------------------------------------------------------
def test_program_checker(api, A):
    import struct
    from raytracingoneweekend_b200 import _lib
    I = api.Mat4x4()
    mat = api.Material.new_lambertian((0.5, 0.5, 0.5))

    def add(words):
        h = C.c_void_p()
        _lib.call("om_scene_create", C.byref(h))
        try:
            pod = api.DynMarched(I, I, 1.0, [], mat).pod
            code = (C.c_uint32 * max(1, len(words)))(*words)
            _lib.call("om_scene_add_dyn_marched", h, C.byref(pod), code, len(words))
        finally:
            _lib.load().om_scene_destroy(h)

    ops = A.SDF_OPS
    add(api.sdf_program(torus_program(0.5, 0.1)))
    add(api.sdf_program(box_program(0.5, 0.4, 0.3)))
    one = struct.unpack("<I", struct.pack("<f", 1.0))[0]
    bad = {
        "empty": [],
        "underflow": [ops["x"], ops["add"]],
        "two results": [ops["x"], ops["y"]],
        "no result": [ops["x"], ops["store"]],
        "const without value": [ops["x"], ops["const"]],
        "unknown opcode": [ops["x"], 99],
        "register out of range": [ops["x"], ops["store"] | (4 << 8), ops["load"] | (4 << 8)],
        "read before write": [ops["load"] | (1 << 8)],
        "argument on plain opcode": [ops["x"] | (1 << 8)],
        "overflow": [ops["x"]] * 9 + [ops["add"]] * 8,
        "too long": [ops["x"]] + [ops["abs"]] * 256,
    }
    for why, words in bad.items():
        with pytest.raises(Exception):
            add(words)
    add([ops["x"]] * 8 + [ops["add"]] * 7)  # exactly OM_SDF_STACK deep
    add([ops["const"], one])


def test_abi_round_trip_and_scan_order(api, A, O):
    w = build_dyn_scene(api)
    from raytracingoneweekend_b200 import _lib
    h = C.c_void_p()
    _lib.call("om_scene_create", C.byref(h))
    for d in w.dyn_marched:
        code = (C.c_uint32 * len(d.code))(*d.code)
        _lib.call("om_scene_add_dyn_marched", h, C.byref(d.pod), code, len(d.code))
    n = C.c_uint64()
    _lib.call("om_scene_dyn_marched_count", h, C.byref(n))
    assert n.value == len(w.dyn_marched) == 4
    for i, d in enumerate(w.dyn_marched):
        nw = C.c_size_t(0)
        _lib.call("om_scene_get_dyn_marched", h, i, None, None, C.byref(nw))
        assert nw.value == len(d.code)
        pod, code = A.DynMarched(), (C.c_uint32 * nw.value)()
        _lib.call("om_scene_get_dyn_marched", h, i, C.byref(pod), code, C.byref(nw))
        assert list(code) == d.code and bytes(pod) == bytes(d.pod)
    with pytest.raises(Exception):
        _lib.call("om_scene_get_dyn_marched", h, 4, None, None, C.byref(nw))
    with pytest.raises(Exception):  # native trait objects still cannot cross
        _lib.call("om_scene_add_dyn_object", h, None)
    _lib.load().om_scene_destroy(h)


def test_oracle_interpreter_equals_the_typed_sdfs(api, A, O):
    """The oracle's DynMarched running torus / box programs must give, bit for bit, what the oracle's MarchedTorus /
    MarchedBox (restated from marched.rs:78-151) give: t, point and the numeric normal with its sign fix.  Unit scale, so
    that `p * scale` and `f * min(scale)` are exact in both.  Only the object id differs (dyn objects have their own tag)."""
    M = api.Material
    m = api.Mat4x4.new_translate((0, 1, 0)) ^ api.Mat4x4.new_rotate_x(0.6) ^ api.Mat4x4.new_rotate_z(1.33)
    typed, dyn = O.OracleScene(), O.OracleScene()
    tor = api.MarchedTorus.new(m, (0.5, 0.1, 0.1), M.new_dielectric(1.5))
    typed.add("marched_tori", (A.MarchedTorus * 1)(tor))
    # the torus keeps translate*rotate and its inverse; hand the same matrices to the dyn object
    w2l = api.Mat4x4(tor.m_world_to_local_translate_rotate[:])
    l2w = api.Mat4x4(tor.m_local_to_world_translate_rotate[:])
    d = api.DynMarched(w2l, l2w, 1.0, api.sdf_program(torus_program(0.5, 0.1)), M.new_dielectric(1.5))
    dyn.add_dyn_marched(d.pod, d.code)
    box = api.MarchedBox((-2, 0.5, 2), (0.5, 0.4, 0.3), M.new_lambertian((0.2, 0.8, 0.2)))
    typed.add("marched_boxes", (A.MarchedBox * 1)(box))
    tb = api.Mat4x4.new_translate((-2, 0.5, 2))
    db = api.DynMarched(api.Mat4x4.new_translate((2, -0.5, -2)), tb, 1.0, api.sdf_program(box_program(0.5, 0.4, 0.3)), M.new_lambertian((0.2, 0.8, 0.2)))
    dyn.add_dyn_marched(db.pod, db.code)
    o, dd = rays(6000, 3)
    a, b = typed.trace_rays(o, dd), dyn.trace_rays(o, dd)
    assert ((a[0] != 0) == (b[0] != 0)).all() and (a[0] != 0).sum() > 300
    assert_hits_equal(a, b, ids=False)
    # typed box is scanned before the typed torus, the dyn torus was added before the dyn box: both orders give the
    # nearest object, so the type tags map onto each other
    tag = (b[0] >> np.uint64(32))
    assert set(np.unique(tag).tolist()) <= {0, A.TYPE_DYN_MARCHED}
    # marched sphere: same distance function (marched.rs:57); its normal is overridden analytically (marched.rs:60-63),
    # so only t and the point are comparable
    s1, s2 = O.OracleScene(), O.OracleScene()
    s1.add("marched_spheres", (A.MarchedSphere * 1)(api.MarchedSphere((0.5, 0.8, 1.0), 0.7, M.new_metal((0.8, 0.8, 0.8)))))
    ds = api.DynMarched(api.Mat4x4.new_translate((-0.5, -0.8, -1.0)), api.Mat4x4.new_translate((0.5, 0.8, 1.0)), 1.0, api.sdf_program(sphere_program(0.7)), M.new_metal((0.8, 0.8, 0.8)))
    s2.add_dyn_marched(ds.pod, ds.code)
    a, b = s1.trace_rays(o, dd), s2.trace_rays(o, dd)
    assert (a[0] != 0).sum() > 300
    for k in (1, 2):
        assert (bits(np.ascontiguousarray(a[k])) == bits(np.ascontiguousarray(b[k]))).all()


# ---- GPU -----------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("with_bounds", [True, False])
def test_dyn_marched_hits_bit_exact(api, A, O, with_bounds):
    fz = build_dyn_scene(api, with_bounds).freeze()
    osc = oracle_scene_from(O, fz)
    o, dd = rays(8000, 5)
    want = osc.trace_rays(o, dd)
    tags = (want[0] >> np.uint64(32))
    assert (tags == A.TYPE_DYN_MARCHED).sum() > 500 and len(np.unique(want[0][tags == A.TYPE_DYN_MARCHED])) == 4
    for fl in (0, A.RENDER_NO_SMEM_SCENE, A.RENDER_BRUTE_FORCE):
        assert_hits_equal(fz.hit(o, dd, flags=fl), want)


@pytest.mark.gpu
@pytest.mark.parametrize("flags_name", ["persistent", "wavefront", "global", "adaptive"])
def test_dyn_marched_render_bit_exact(api, A, O, flags_name):
    from test_gpu_parity import assert_pixels_equal
    fz = build_dyn_scene(api).freeze()
    osc = oracle_scene_from(O, fz)
    fl = {"persistent": A.RENDER_PERSISTENT, "wavefront": A.RENDER_WAVEFRONT, "global": A.RENDER_PERSISTENT | A.RENDER_NO_SMEM_SCENE,
          "adaptive": A.RENDER_ADAPTIVE}[flags_name]
    spp = 24 if flags_name == "adaptive" else 4
    p = api.render_params(64, 48, spp, seed=21, flags=fl)
    r = api.Renderer(fz, cam(api), p)
    r.launch(0); r.sync()
    got, traced = r.read_pixels(), r.samples_traced()
    r.close()
    want = osc.render(cam(api), p, nthreads=4)
    assert_pixels_equal(got, want["pixels"])
    if flags_name != "adaptive":
        assert traced == 64 * 48 * spp


@pytest.mark.gpu
def test_dyn_bounds_do_not_change_the_image(api, A):
    """Finite user bounds only let rays skip the march; with and without them the pixels are the same."""
    from test_gpu_parity import assert_pixels_equal
    out = []
    for with_bounds in (True, False):
        fz = build_dyn_scene(api, with_bounds).freeze()
        r = api.Renderer(fz, cam(api, 16 / 9), api.render_params(320, 180, 6, seed=2))
        r.launch(0); r.sync()
        out.append(r.read_pixels())
        r.close()
    assert_pixels_equal(out[0], out[1])
