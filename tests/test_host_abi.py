"""CPU-side tests of the product: the C-ABI library loads and exports what the header declares,
host-side constructors / generators equal the oracle's restatement bit for bit, arguments are
validated, and nothing computes without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(api):
    from raytracingoneweekend_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "ottomarcher.h")).read()
    declared = set(re.findall(r"OM_API\s+[\w\s\*]+?\b(om_\w+)\s*\(", hdr))
    assert len(declared) >= 55
    L = C.CDLL(_lib.LIB_PATH)
    missing = [n for n in sorted(declared) if not hasattr(L, n)]
    assert not missing, missing
    assert declared == set(_lib.PROTOTYPES), declared ^ set(_lib.PROTOTYPES)


def test_struct_layout_matches_header(A):
    assert C.sizeof(A.Material) == 24 and C.sizeof(A.Sphere) == 152 and C.sizeof(A.Cube) == 152
    assert C.sizeof(A.InfinitePlane) == 48 and C.sizeof(A.Barycentric) == 140
    assert C.sizeof(A.MarchedSphere) == 40 and C.sizeof(A.MarchedBox) == 48 and C.sizeof(A.MarchedTorus) == 196
    assert C.sizeof(A.Camera) == 104 and C.sizeof(A.Pixel) == 40 and C.sizeof(A.RenderParams) == 56
    assert np.dtype(A.PIXEL_DTYPE).itemsize == 40


def test_constructors_equal_oracle(api, O, A):
    L = O.lib()
    rng = np.random.default_rng(1)
    mat = api.Material.new_metal_fuzz((0.7, 0.6, 0.5), 0.25)
    for _ in range(20):
        t, r, s = rng.uniform(-5, 5, 3), rng.uniform(0, 6.28, 3), rng.uniform(0.2, 3, 3)
        m = (api.Mat4x4.new_translate(t) ^ api.Mat4x4.new_rotate_x(r[0]) ^ api.Mat4x4.new_rotate_y(r[1]) ^ api.Mat4x4.new_rotate_z(r[2])
             ^ api.Mat4x4.new_scale(s))
        so = A.Sphere(); L.omo_sphere_new(C.byref(so), m.e, C.byref(mat))
        assert bytes(api.Sphere.new(m, mat)) == bytes(so)
        co = A.Cube(); L.omo_cube_new(C.byref(co), m.e, C.byref(mat))
        assert bytes(api.Cube.new(m, mat)) == bytes(co)
        to = A.MarchedTorus(); sizes = (C.c_float * 3)(0.5, 0.1, 0.1); L.omo_marched_torus_new(C.byref(to), m.e, sizes, C.byref(mat))
        assert bytes(api.MarchedTorus.new(m, (0.5, 0.1, 0.1), mat)) == bytes(to)
        inv = np.array(m.fast_homogenous_inverse().e[:], np.float64).reshape(4, 4)
        assert np.allclose(inv @ np.array(m.e[:], np.float64).reshape(4, 4), np.eye(4), atol=2e-4)
        p = [rng.uniform(-3, 3, 3) for _ in range(3)]
        bo = A.Barycentric(); L.omo_barycentric_new3points(C.byref(bo), O.f3(p[0]), O.f3(p[1]), O.f3(p[2]), C.byref(mat))
        assert bytes(api.Triangle.new3points(p[0], p[1], p[2], mat).pod) == bytes(bo)
        lf, la = rng.uniform(-10, 10, 3), rng.uniform(-1, 1, 3)
        assert bytes(api.Camera.new(lf, la, (0, 1, 0), 35.0, 1.7, 0.2, 7.0)) == bytes(O.camera_new(lf, la, (0, 1, 0), 35.0, 1.7, 0.2, 7.0))
    so = A.Sphere(); L.omo_sphere_new_with_radius(C.byref(so), O.f3(0, -1000, 0), 1000.0, C.byref(mat))
    assert bytes(api.Sphere.new_with_radius((0, -1000, 0), 1000.0, mat)) == bytes(so)


@pytest.mark.parametrize("seed", [1, 2, 12345678901234])
def test_random_scene_generator_equals_oracle(api, O, seed):
    fz = api.FrozenHittableList.generate("random_scene", seed)
    osc = O.OracleScene().random_scene(seed)
    got = fz.lists()
    assert list(fz.counts().values()) == osc.counts()
    for k in ("spheres", "cubes", "triangles", "parallelograms", "marched_tori"):
        assert bytes(got[k]) == bytes(osc.get(k)), k
    assert 440 <= fz.counts()["spheres"] <= 485 and fz.counts()["marched_tori"] == 1


def test_other_generators(api):
    assert api.FrozenHittableList.generate("basic_scene").counts()["spheres"] == 3
    c = api.FrozenHittableList.generate("rtiow_classic", 1).counts()
    assert c["marched_tori"] == 0 and c["cubes"] == 0 and c["spheres"] > 400
    c = api.FrozenHittableList.generate("marched", 3).counts()
    assert c["marched_spheres"] + c["marched_boxes"] + c["marched_tori"] == 65 and c["spheres"] == 0
    fz = api.FrozenHittableList.generate("glass_heavy", 4)
    mats = np.array([s.material.mat_type for s in fz.lists()["spheres"]][1:])
    assert (mats == 2).mean() >= 0.5
    fz = api.FrozenHittableList.generate("many_spheres", 5, 20000)
    st = fz.stats()
    assert fz.counts()["spheres"] == 20001 and st["bvh_prims"] == 20000 and st["always_test"] == 1
    assert 19999 / 3 <= st["bvh_nodes"] <= 19999  # four children per node (omb::collapse4): between (n - 1) / 3 and n - 1 nodes
    assert st["fits_smem"] == 0 and st["bvh_depth"] < 64


def test_bvh_stats_of_rtiow(api):
    st = api.FrozenHittableList.generate("random_scene", 1).stats()
    assert st["always_test"] == 1                   # the r=1000 ground sphere is tested once per ray
    assert (st["bvh_prims"] - 1) / 3 <= st["bvh_nodes"] <= st["bvh_prims"] - 1   # one primitive per leaf, up to four children per node
    assert st["fits_smem"] == 1 and st["device_bytes"] < 100 * 1024


def test_bloom_and_jitter_equal_oracle(api, O, A):
    L = O.lib()
    ids = [0, 1, A.obj_id(1, 0), A.obj_id(8, 0), A.obj_id(5, 123456), 2**64 - 1] + list(range(2, 500))
    assert all(api.bloom_get_hash(i) == L.omo_bloom_hash(i) for i in ids)
    for spp, seed in [(1, 0), (10, 3), (1024, 7), (4096, 2**40 + 1)]:
        t = np.zeros(spp, np.uint8)
        L.omo_jitter_table(spp, seed, t.ctypes.data)
        assert (api.jitter_table(spp, seed) == t).all()


def test_errors_not_aborts(api, A):
    from raytracingoneweekend_b200 import _lib
    fz = api.FrozenHittableList.generate("basic_scene")
    cam = api.Camera.world_camera(90.0, 1.5)
    bad = [dict(width=1, height=8, spp=4), dict(width=8, height=8, spp=0), dict(width=8, height=8, spp=4, sample_base=4),
           dict(width=8, height=8, spp=4, sample_base=2, sample_count=3), dict(width=8, height=8, spp=4, part_rank=2, part_world=2),
           dict(width=8, height=8, spp=8, flags=A.RENDER_ADAPTIVE, sample_base=1, sample_count=2)]
    for kw in bad:
        with pytest.raises(api.OmError) as e:
            api.Renderer(fz, cam, api.render_params(**kw))
        assert e.value.status == A.OM_ERR_INVALID, kw
    unfrozen = api.FrozenHittableList()
    with pytest.raises(api.OmError) as e:
        api.Renderer(unfrozen, cam, api.render_params(8, 8, 1))
    assert e.value.status == A.OM_ERR_INVALID
    with pytest.raises(api.OmError) as e:     # frozen scenes are immutable, like FrozenHittableList
        _lib.call("om_scene_generate", fz.h, 1, 0, 0)
    assert e.value.status == A.OM_ERR_INVALID
    w = api.HittableList()
    with pytest.raises(api.OmError) as e:     # world += Arc<dyn Traced> (hits.rs:91-95)
        w += object()
    assert e.value.status == A.OM_ERR_UNSUPPORTED
    assert _lib.load().om_render_destroy(None) == 0 and _lib.load().om_scene_destroy(None) == 0


def test_no_cpu_fallback(api, A):
    """Without a CUDA device the compute entry points fail loudly; they never compute on the CPU."""
    from raytracingoneweekend_b200 import _lib
    if _lib.load().om_device_count() > 0:
        pytest.skip("a GPU is visible")
    fz = api.FrozenHittableList.generate("basic_scene")
    cam = api.Camera.world_camera(90.0, 1.5)
    with pytest.raises(api.OmError) as e:
        api.render(cam, fz, 5, 0.001, 100.0, 2, 8, 8)
    assert e.value.status == A.OM_ERR_NO_DEVICE
    with pytest.raises(api.OmError) as e:
        fz.hit([[0, 0, 0]], [[0, 0, -1]])
    assert e.value.status == A.OM_ERR_NO_DEVICE


def test_product_does_not_use_the_oracle():
    """The oracle is test infrastructure: nothing under the package or include/ may import, include,
    link or load it."""
    banned = [r"^\s*(from|import)\s+oracle", r"omo_\w+", r"libom_oracle", r"#include\s+\"[^\"]*oracle", r"ref_(math|scene|render|rng)\.hpp"]
    files = [os.path.join(ROOT, "include", "ottomarcher.h")]
    for dirpath, _, names in os.walk(os.path.join(ROOT, "raytracingoneweekend_b200")):
        files += [os.path.join(dirpath, f) for f in names if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".inl"))]
    for f in files:
        src = open(f, errors="replace").read()
        for pat in banned:
            assert not re.search(pat, src, re.M), (f, pat)


def test_gpu_bvh_build_fails_loudly_without_a_device(api, A):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    with pytest.raises(Exception) as e:
        api.FrozenHittableList.generate("many_spheres", 5, 100, flags=A.FREEZE_GPU_BVH)
    assert "GPU BVH build" in str(e.value)


def test_host_bvh_depth_stays_below_the_traversal_stack(api, A):
    """Exponentially spread spheres make SAH peel one object per level; the builder has to switch to halving."""
    w = api.HittableList()
    m = api.Material.new_lambertian((0.5, 0.5, 0.5))
    for k in range(-100, 101):
        w += api.Sphere.new_with_radius((2.0 ** k, 0.0, 0.0), 0.1 * 2.0 ** k, m)
    st = w.freeze(flags=A.FREEZE_NO_ALWAYS_LIST).stats()
    assert st["bvh_prims"] == 201 and st["bvh_depth"] <= 62


def test_reference_arm_never_loads_the_product_library():
    """bench.py --impl reference builds its scene with the oracle alone: after timing a sample of C1 the process
    maps oracle/libom_oracle.so and not libottomarcher_cuda.so."""
    import subprocess, sys
    code = ("import sys; sys.path.insert(0, %r); import bench\n"
            "rate, threads, desc, dt = bench.cpu_reference_rate(bench.WORKLOADS['C1'], seconds_target=0.5, threads=2)\n"
            "maps = open('/proc/self/maps').read()\n"
            "assert rate > 0 and 'libom_oracle.so' in maps and 'libottomarcher_cuda' not in maps, maps\n"
            "print('ok', desc)\n") % ROOT
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and out.stdout.startswith("ok"), out.stderr[-2000:]
