"""Known-answer tests that pin the ORACLE to the reference's source formulas.

The reference has no tests or golden vectors, so these values are derived by hand
from the cited Rust lines; Philox is pinned to the Random123 known-answer vectors.
"""
import ctypes as C
import math

import numpy as np
import pytest


def _f3(*v):
    return (C.c_float * 3)(*v)


def test_philox_random123_kat(O):
    L = O.lib()

    def ph(ctr, key):
        out = (C.c_uint32 * 4)()
        L.omo_philox((C.c_uint32 * 4)(*ctr), (C.c_uint32 * 2)(*key), out)
        return list(out)

    assert ph([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert ph([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert ph([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]) == [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_rand_f32_is_24_bit_uniform(O):
    out = np.zeros(4096, np.float32)
    O.lib().omo_rand_stream(5, 1, 2, 0, 4096, out.ctypes.data)
    assert out.min() >= 0.0 and out.max() < 1.0
    assert np.all(out * 16777216.0 == np.floor(out * 16777216.0))  # multiples of 2^-24 (utils.rs:25)
    assert abs(out.mean() - 0.5) < 0.02


def test_scramble_and_bloom(O):
    L = O.lib()
    assert L.omo_scramble(0) == 0                      # "maps 0 to 0", utils.rs:46
    assert L.omo_bloom_hash(0) == 0                    # bit = (id != 0), utils.rs:96
    h = L.omo_bloom_hash((1 << 32) | 1)
    assert 1 <= bin(h).count("1") <= 9                 # nine hashes into 64 bits
    # scramble by hand for obj_id = 1: id1 = 1 -> xorshift 13,7,17; id2 = 0
    id1 = 1
    id1 ^= (id1 << 13) & (2**64 - 1); id1 ^= id1 >> 7; id1 ^= (id1 << 17) & (2**64 - 1)
    assert L.omo_scramble(1) == id1


def test_u64_to_color(O):
    out = (C.c_uint8 * 3)()
    O.lib().omo_u64_to_color(0x0807060504030201, out)  # utils.rs:59-69
    assert list(out) == [0x01 ^ 0x08 ^ 0x04, 0x02 ^ 0x05 ^ 0x06, 0x03 ^ 0x07]


@pytest.mark.parametrize("v,expect", [(0.0, 0), (0.25, 128), (1.0, 255), (4.0, 255), (-1.0, 0), (float("nan"), 0)])
def test_tone_map(O, v, expect):
    # normalize_color: clamp(sqrt(c),0,0.999) then (c*256) as u8 (utils.rs:12-17, vec3.rs:73-76)
    out = (C.c_uint8 * 3)()
    O.lib().omo_tone_map(_f3(v, v, v), out)
    assert list(out) == [expect] * 3


def test_reflect_refract_schlick(O):
    L = O.lib()
    out = _f3(0, 0, 0)
    L.omo_reflect(_f3(1, -1, 0), _f3(0, 1, 0), out)          # v - 2 v.n n
    assert list(out) == [1.0, 1.0, 0.0]
    L.omo_refract(_f3(0, -1, 0), _f3(0, 1, 0), 1.0 / 1.5, out)  # normal incidence goes straight through
    assert np.allclose(list(out), [0, -1, 0], atol=1e-7)
    r0 = ((1 - 1.5) / (1 + 1.5)) ** 2
    assert abs(L.omo_reflectance(1.0, 1.5) - r0) < 1e-7       # cos = 1 -> r0^2
    assert abs(L.omo_reflectance(0.0, 1.5) - 1.0) < 1e-6      # grazing -> 1


def _one_sphere_scene(api, O, A, center=(0, 0, 0), r=1.0):
    s = O.OracleScene()
    sp = api.Sphere.new_with_radius(center, r, api.Material.new_lambertian((0.5, 0.5, 0.5)))
    s.add("spheres", (A.Sphere * 1)(sp))
    return s


def test_sphere_kat(api, O, A):
    s = _one_sphere_scene(api, O, A)
    ids, t, p, n = s.trace_rays([[0, 0, -3]], [[0, 0, 1]])
    assert ids[0] == A.obj_id(A.TYPE_SPHERE, 0) and t[0] == 2.0            # near root
    assert list(p[0]) == [0, 0, -1] and list(n[0]) == [0, 0, -1]
    ids, t, p, n = s.trace_rays([[0, 0, 0]], [[0, 0, 1]])                   # origin inside -> far root
    assert t[0] == 1.0 and list(n[0]) == [0, 0, 1]                          # normal stays outward
    ids, t, _, _ = s.trace_rays([[0, 0, -3]], [[0, 0, 1]], tmin=0.001, tmax=1.5)  # both roots out of range
    assert ids[0] == 0 and np.isinf(t[0])
    ids, t, _, _ = s.trace_rays([[0, 2, -3]], [[0, 0, 1]])                  # discriminant < 0
    assert ids[0] == 0
    ids, t, _, _ = s.trace_rays([[0, 0, -3]], [[0, 0, 5]])                  # Ray::new normalises the direction
    assert t[0] == 2.0


def test_plane_kat(api, O, A):
    s = O.OracleScene()
    pl = api.InfinitePlane.new((0, -1, 0), (0, 2, 0), api.Material.new_metal((1, 1, 1)))
    assert list(pl.normal) == [0, 1, 0]                                      # InfinitePlane::new normalises
    s.add("infinite_planes", (A.InfinitePlane * 1)(pl))
    ids, t, p, n = s.trace_rays([[0, 1, 0]], [[0, -1, 0]])
    assert ids[0] == A.obj_id(A.TYPE_INFINITE_PLANE, 0) and t[0] == 2.0 and list(n[0]) == [0, 1, 0]
    ids, t, p, n = s.trace_rays([[0, -3, 0]], [[0, 1, 0]])                   # from below: normal flips against the ray
    assert t[0] == 2.0 and list(n[0]) == [0, -1, 0]
    ids, _, _, _ = s.trace_rays([[0, 1, 0]], [[1, 1e-7, 0]])                 # |n.d| < 1e-6 -> parallel -> miss
    assert ids[0] == 0
    ids, _, _, _ = s.trace_rays([[0, 1, 0]], [[0, 1, 0]])                    # negative root
    assert ids[0] == 0


def test_triangle_vs_parallelogram_kat(api, O, A):
    m = api.Material.new_lambertian((1, 1, 0))
    tri = api.Triangle.new3points((0, 0, 0), (1, 0, 0), (0, 1, 0), m).pod
    par = api.Parallelogram.new3points((0, 0, 0), (1, 0, 0), (0, 1, 0), m).pod
    st = O.OracleScene().add("triangles", (A.Barycentric * 1)(tri))
    sp = O.OracleScene().add("parallelograms", (A.Barycentric * 1)(par))
    inside_tri, inside_par_only, outside, edge = [0.25, 0.25, 1], [0.75, 0.75, 1], [1.5, 0.5, 1], [0.5, 0.0, 1]
    d = [0, 0, -1]
    assert st.trace_rays([inside_tri], [d])[0][0] == A.obj_id(A.TYPE_TRIANGLE, 0)
    assert st.trace_rays([inside_par_only], [d])[0][0] == 0                  # lambda3 < 0 (traced.rs:169-171)
    assert sp.trace_rays([inside_par_only], [d])[0][0] == A.obj_id(A.TYPE_PARALLELOGRAM, 0)
    assert sp.trace_rays([outside], [d])[0][0] == 0
    assert st.trace_rays([edge], [d])[0][0] == 0 and sp.trace_rays([edge], [d])[0][0] == 0  # strict inequalities
    ids, t, p, n = st.trace_rays([inside_tri], [d])
    assert t[0] == 1.0 and list(n[0]) == [0, 0, 1]


def test_cube_kat(api, O, A):
    cb = api.Cube.new_with_length((0, 0, 0), 2.0, api.Material.new_metal((0.7, 0.6, 0.5)))
    s = O.OracleScene().add("cubes", (A.Cube * 1)(cb))
    ids, t, p, n = s.trace_rays([[0.3, 0.2, -5]], [[0, 0, 1]])
    assert ids[0] == A.obj_id(A.TYPE_CUBE, 0) and t[0] == 4.0
    assert list(n[0]) == [0, 0, -2]                                          # M . axis: NOT normalised (traced.rs:296)
    ids, t, p, n = s.trace_rays([[0, 0, 0]], [[1, 0, 0]])                     # from inside: exits through +x
    assert t[0] == 1.0 and list(n[0]) == [2, 0, 0]
    assert s.trace_rays([[3, 3, -5]], [[0, 0, 1]])[0][0] == 0


def test_marched_sdf_kat(api, O, A):
    L = O.lib()
    m = api.Material.new_dielectric(1.5)
    ms = api.MarchedSphere((1, 0, 0), 0.5, m)
    assert L.omo_marched_sdf(0, C.byref(ms), _f3(3, 0, 0)) == 1.5           # |p-c| - r
    mb = api.MarchedBox((0, 0, 0), (1, 1, 1), m)
    assert L.omo_marched_sdf(1, C.byref(mb), _f3(3, 0, 0)) == 2.0
    assert L.omo_marched_sdf(1, C.byref(mb), _f3(0, 0, 0)) == -1.0          # inside: max component
    mt = api.MarchedTorus.new(api.Mat4x4(), (0.5, 0.1, 0.1), m)
    assert abs(L.omo_marched_sdf(2, C.byref(mt), _f3(0, 0, 0)) - 0.4) < 1e-7  # R - r at the centre
    assert abs(L.omo_marched_sdf(2, C.byref(mt), _f3(0.5, 1.0, 0)) - 0.9) < 1e-7
    n = _f3(0, 0, 0)
    L.omo_marched_normal(0, C.byref(ms), _f3(3, 0, 0), n)                   # analytic override marched.rs:60-63
    assert list(n) == [1, 0, 0]


def test_marched_hit_and_unstuck(api, O, A):
    m = api.Material.new_lambertian((0.5, 0.5, 0.5))
    s = O.OracleScene().add("marched_spheres", (A.MarchedSphere * 1)(api.MarchedSphere((0, 0, 0), 1.0, m)))
    ids, t, p, n = s.trace_rays([[0, 0, -3]], [[0, 0, 1]])
    assert ids[0] == A.obj_id(A.TYPE_MARCHED_SPHERE, 0)
    assert 2.0 - 0.0011 < t[0] <= 2.0                                        # stops within HIT_SIZE of the surface
    # starting on the surface: unstuck steps 0.0005 at a time until |sdf| >= 0.001, then marches |sdf| (inside too)
    ids, t, _, _ = s.trace_rays([[0, 0, -1]], [[0, 0, 1]])
    assert ids[0] == A.obj_id(A.TYPE_MARCHED_SPHERE, 0) and 1.99 < t[0] < 2.01


def test_scan_order_tie_break(api, O, A):
    """Two identical spheres: `root > t_max` rejects only strictly larger roots, so the LATER one wins."""
    m = api.Material.new_lambertian((0.5, 0.5, 0.5))
    sp = api.Sphere.new_with_radius((0, 0, 0), 1.0, m)
    s = O.OracleScene().add("spheres", (A.Sphere * 2)(sp, sp))
    assert s.trace_rays([[0, 0, -3]], [[0, 0, 1]])[0][0] == A.obj_id(A.TYPE_SPHERE, 1)


def test_stats_add(O, A):
    px = A.Pixel()
    L = O.lib()
    assert L.omo_stats_add(C.byref(px), _f3(0.25, 0.25, 0.25), 2.0, 0) == 0
    assert px.n == 1 and list(px.color) == [128, 128, 128] and px.bad_avgs == 0 and px.avg_depth == 2.0 and px.bloom == 0
    for k in range(4):
        assert L.omo_stats_add(C.byref(px), _f3(0.25, 0.25, 0.25), 4.0, 7) == 0
    assert px.bad_avgs == 4
    assert L.omo_stats_add(C.byref(px), _f3(0.25, 0.25, 0.25), math.inf, 7) == 1   # 5 unchanged colours -> done
    assert px.n == 6 and math.isinf(px.avg_depth) and px.bloom == L.omo_bloom_hash(7)
    assert L.omo_stats_add(C.byref(px), _f3(50, 0, 0), 1.0, 0) == 0               # colour changes -> run resets
    assert px.bad_avgs == 0 and math.isinf(px.avg_depth)                           # INF is absorbing
    L.omo_stats_add(C.byref(px), _f3(-0.0, -0.0, -0.0), 1.0, 0)                    # -black adds nothing
    assert px.sum[1] == 1.5


def test_jitter_table_is_shuffled_2x2(O):
    t = np.zeros(64, np.uint8)
    O.lib().omo_jitter_table(64, 9, t.ctypes.data)
    assert sorted(t.tolist()) == sorted([0, 1, 2, 3] * 16)   # multiset of render_thread.rs:166-171
    assert t.tolist() != [(((s // 2) & 1) << 1) | (s & 1) for s in range(64)]


def test_render_counters_and_determinism(api, O, A):
    cam = O.camera_new((13, 2, 3), (0, 0, 0), (0, 1, 0), 20.0, 1.5, 0.1, 10.0)
    s = O.OracleScene().random_scene(1)
    p = api.render_params(24, 16, 6, seed=3)
    a = s.render(cam, p, nthreads=1)
    b = s.render(cam, p, nthreads=3)
    assert a["traced"] == a["credited"] == 24 * 16 * 6
    assert a["pixels"].tobytes() == b["pixels"].tobytes()    # result independent of the thread split
    # two passes of 3 == one pass of 6 (sample-major outer loop, render_thread.rs:176)
    c = s.render(cam, p, spp_pass=3, nthreads=2)
    c = s.render(cam, p, spp_pass=3, nthreads=2, pixels=c["pixels"])
    assert c["pixels"].tobytes() == a["pixels"].tobytes()
    # adaptive: credited always reaches W*H*spp (left-overs credited at retirement, :196-198)
    pa = api.render_params(24, 16, 64, seed=3, flags=A.RENDER_ADAPTIVE)
    d = s.render(cam, pa, nthreads=2)
    assert d["credited"] == 24 * 16 * 64 and d["traced"] < d["credited"]
    assert ((d["pixels"]["bad_avgs"] >= 5) | (d["pixels"]["n"] == 64)).all()


def test_view_modes_known_answers(O, A):
    """Display modes of main.rs:361-437 on hand-made pixel buffers."""
    import numpy as np
    W, H = 4, 3
    px = np.zeros(W * H, dtype=A.PIXEL_DTYPE)
    px["n"] = 4
    px["sum"] = np.array([1.0, 0.3, 0.0], np.float32) * 4   # average colour (1, 0.3, 0): sqrt(.3) * 256 = 140.2, far from a rounding edge
    px["color"] = np.array([255, 140, 0], np.uint8)         # sqrt, clamp .999, *256
    px["avg_depth"] = 2.0
    px["bloom"] = 0
    assert (O.view(px, W, H, 0) == px["color"].reshape(H, W, 3)).all()
    assert (O.view(px, W, H, 1) == 255).all()                      # n / max n = 1 -> .999 * 256
    assert (O.view(px, W, H, 3) == 255).all()                      # depth / max depth = 1
    assert (O.view(px, W, H, 5) == 0).all()                        # scramble(0) = 0 -> black
    for mode in (2, 4, 6):                                         # any weighted mean of a constant image is that image
        assert (O.view(px, W, H, mode) == px["color"].reshape(H, W, 3)).all(), mode
    px["avg_depth"][5] = np.inf                                    # sky pixel: depth view paints it green, depth blur copies its colour
    v3 = O.view(px, W, H, 3)
    assert tuple(v3[1, 1]) == (0, 255, 0)
    assert tuple(O.view(px, W, H, 4)[1, 1]) == (255, 140, 0)
    px["n"][0] = 1                                                 # quarter of the samples -> sqrt(.25) * 256 = 128
    assert tuple(O.view(px, W, H, 1)[0, 0]) == (128, 128, 128)
