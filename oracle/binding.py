"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes loader for oracle/libom_oracle.so (the CPU restatement of the reference).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
PARITY UNPINNED: see oracle/ref_math.hpp.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from raytracingoneweekend_b200 import _abi as A

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libom_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile the oracle with oracle/Makefile (g++ only)."""
    if force or not os.path.exists(_SO):
        subprocess.check_call(["make", "-C", _HERE] + (["-B"] if force else []), stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_SO):
        build()
    L = C.CDLL(_SO)
    vp, u32, u64, f32, sz = C.c_void_p, C.c_uint32, C.c_uint64, C.c_float, C.c_size_t
    P = C.POINTER
    L.omo_scene_create.restype = vp
    L.omo_scene_destroy.argtypes = [vp]
    for name, T in [("spheres", A.Sphere), ("cubes", A.Cube), ("triangles", A.Barycentric),
                    ("infinite_planes", A.InfinitePlane), ("parallelograms", A.Barycentric),
                    ("marched_spheres", A.MarchedSphere), ("marched_boxes", A.MarchedBox),
                    ("marched_tori", A.MarchedTorus)]:
        getattr(L, "omo_scene_add_" + name).argtypes = [vp, P(T), sz]
    L.omo_scene_add_dyn_marched.argtypes = [vp, P(A.DynMarched), P(u32), sz]
    L.omo_scene_add_dyn_traced.argtypes = [vp, vp, sz]
    L.omo_scene_counts.argtypes = [vp, P(u64)]
    L.omo_scene_random.argtypes = [vp, u64]
    L.omo_scene_basic.argtypes = [vp]
    L.omo_scene_get_spheres.argtypes = [vp, P(A.Sphere)]
    L.omo_scene_get_cubes.argtypes = [vp, P(A.Cube)]
    L.omo_scene_get_triangles.argtypes = [vp, P(A.Barycentric)]
    L.omo_scene_get_parallelograms.argtypes = [vp, P(A.Barycentric)]
    L.omo_scene_get_marched_tori.argtypes = [vp, P(A.MarchedTorus)]
    L.omo_render.argtypes = [vp, P(A.Camera), P(A.RenderParams), vp, vp, u32, u32, P(u64), P(u64), P(u64)]
    L.omo_render.restype = C.c_int
    L.omo_view.argtypes = [C.c_int, vp, C.c_int, C.c_int, vp]
    L.omo_primary_rays.argtypes = [P(A.Camera), P(A.RenderParams), u32, u32, vp, vp]
    L.omo_primary.argtypes = [vp, P(A.Camera), P(A.RenderParams), u32, u32, vp, vp]
    L.omo_trace_rays.argtypes = [vp, vp, vp, sz, f32, f32, vp, vp, vp, vp]
    L.omo_sample.argtypes = [vp, P(A.Camera), P(A.RenderParams), u32, u32, P(f32), P(f32), P(u64)]
    L.omo_camera_new.argtypes = [P(A.Camera), P(f32), P(f32), P(f32), f32, f32, f32, f32]
    L.omo_sphere_new.argtypes = [P(A.Sphere), P(f32), P(A.Material)]
    L.omo_sphere_new_with_radius.argtypes = [P(A.Sphere), P(f32), f32, P(A.Material)]
    L.omo_cube_new.argtypes = [P(A.Cube), P(f32), P(A.Material)]
    L.omo_cube_new_with_length.argtypes = [P(A.Cube), P(f32), f32, P(A.Material)]
    L.omo_infinite_plane_new.argtypes = [P(A.InfinitePlane), P(f32), P(f32), P(A.Material)]
    L.omo_barycentric_new3points.argtypes = [P(A.Barycentric), P(f32), P(f32), P(f32), P(A.Material)]
    L.omo_marched_torus_new.argtypes = [P(A.MarchedTorus), P(f32), P(f32), P(A.Material)]
    L.omo_mat4_translate.argtypes = [P(f32), f32, f32, f32]
    L.omo_mat4_scale.argtypes = [P(f32), f32, f32, f32]
    for ax in "xyz":
        getattr(L, "omo_mat4_rotate_" + ax).argtypes = [P(f32), f32]
    L.omo_mat4_mul.argtypes = [P(f32), P(f32), P(f32)]
    L.omo_mat4_fast_homogenous_inverse.argtypes = [P(f32), P(f32)]
    L.omo_philox.argtypes = [P(u32), P(u32), P(u32)]
    L.omo_rand_stream.argtypes = [u64, u32, u32, u32, u32, vp]
    L.omo_jitter_table.argtypes = [u32, u64, vp]
    L.omo_scramble.argtypes = [u64]
    L.omo_scramble.restype = u64
    L.omo_bloom_hash.argtypes = [u64]
    L.omo_bloom_hash.restype = u64
    L.omo_u64_to_color.argtypes = [u64, P(C.c_uint8)]
    L.omo_tone_map.argtypes = [P(f32), P(C.c_uint8)]
    L.omo_stats_add.argtypes = [P(A.Pixel), P(f32), f32, u64]
    L.omo_stats_add.restype = C.c_int
    L.omo_reflect.argtypes = [P(f32), P(f32), P(f32)]
    L.omo_refract.argtypes = [P(f32), P(f32), f32, P(f32)]
    L.omo_reflectance.argtypes = [f32, f32]
    L.omo_reflectance.restype = f32
    L.omo_marched_sdf.argtypes = [u32, vp, P(f32)]
    L.omo_marched_sdf.restype = f32
    L.omo_marched_normal.argtypes = [u32, vp, P(f32), P(f32)]
    L.omo_scatter.argtypes = [P(A.Material), P(f32), P(f32), P(f32), P(f32), u64, u32, u32, P(f32), P(f32), P(f32)]
    L.omo_hardware_threads.restype = u32
    L.omo_version.restype = C.c_char_p
    _lib = L
    return L


def f3(x, y=None, z=None):
    if y is None:
        x, y, z = x
    return (C.c_float * 3)(x, y, z)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class OracleScene:
    """FrozenHittableList of the oracle, filled from the same POD arrays the CUDA library gets."""

    _ADD = {A.Sphere: "spheres", A.Cube: "cubes", A.InfinitePlane: "infinite_planes",
            A.MarchedSphere: "marched_spheres", A.MarchedBox: "marched_boxes", A.MarchedTorus: "marched_tori"}

    def __init__(self):
        self.h = lib().omo_scene_create()

    def __del__(self):
        if getattr(self, "h", None):
            lib().omo_scene_destroy(self.h)
            self.h = None

    def add(self, kind: str, arr):
        """kind in spheres,cubes,triangles,infinite_planes,parallelograms,marched_spheres,marched_boxes,marched_tori;
        arr is a ctypes array of the matching struct."""
        n = len(arr)
        if n:
            getattr(lib(), "omo_scene_add_" + kind)(self.h, arr, n)
        return self

    def add_dyn_traced(self, arr):
        """A ctypes array of om_dyn_traced (Arc<dyn Traced> as data: quadrics)."""
        if len(arr):
            lib().omo_scene_add_dyn_traced(self.h, arr, len(arr))
        return self

    def add_dyn_marched(self, pod, code):
        """One om_dyn_marched and its local_sdf program (list of u32 words)."""
        arr = (C.c_uint32 * max(1, len(code)))(*code)
        lib().omo_scene_add_dyn_marched(self.h, C.byref(pod), arr, len(code))
        return self

    def counts(self):
        c = (C.c_uint64 * 8)()
        lib().omo_scene_counts(self.h, c)
        return list(c)

    def random_scene(self, seed):
        lib().omo_scene_random(self.h, seed)
        return self

    def basic_scene(self):
        lib().omo_scene_basic(self.h)
        return self

    def get(self, kind):
        T = {"spheres": A.Sphere, "cubes": A.Cube, "triangles": A.Barycentric, "parallelograms": A.Barycentric,
             "marched_tori": A.MarchedTorus}[kind]
        idx = {"spheres": 0, "cubes": 1, "triangles": 2, "parallelograms": 4, "marched_tori": 7}[kind]
        n = self.counts()[idx]
        arr = (T * n)()
        if n:
            getattr(lib(), "omo_scene_get_" + kind)(self.h, arr)
        return arr

    # -- queries -----------------------------------------------------------------------------
    def trace_rays(self, origins, dirs, tmin=0.001, tmax=100.0):
        o = np.ascontiguousarray(origins, dtype=np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=np.float32).reshape(-1, 3)
        n = o.shape[0]
        ids = np.zeros(n, np.uint64)
        t = np.zeros(n, np.float32)
        pts = np.zeros((n, 3), np.float32)
        nrm = np.zeros((n, 3), np.float32)
        lib().omo_trace_rays(self.h, _ptr(o), _ptr(d), n, tmin, tmax, _ptr(ids), _ptr(t), _ptr(pts), _ptr(nrm))
        return ids, t, pts, nrm

    def primary(self, cam, params, mode=0, sample_index=0):
        n = params.width * params.height
        ids = np.zeros(n, np.uint64)
        t = np.zeros(n, np.float32)
        lib().omo_primary(self.h, C.byref(cam), C.byref(params), mode, sample_index, _ptr(ids), _ptr(t))
        return ids, t

    def sample(self, cam, params, pixel, sample):
        col = (C.c_float * 3)()
        depth = C.c_float()
        oid = C.c_uint64()
        lib().omo_sample(self.h, C.byref(cam), C.byref(params), pixel, sample, col, C.byref(depth), C.byref(oid))
        return np.array(col[:], np.float32), depth.value, oid.value

    def render(self, cam, params, spp_pass=0, nthreads=0, pixels=None, want_sumsq=False, want_counters=False):
        """render_thread::render over all threads.  Returns dict(pixels, credited, traced[, sumsq, counters])."""
        n = params.width * params.height
        if pixels is None:
            pixels = np.zeros(n, dtype=A.PIXEL_DTYPE)
        sumsq = np.zeros((n, 3), np.float64) if want_sumsq else None
        counters = (C.c_uint64 * 5)() if want_counters else None
        cr, tr = C.c_uint64(0), C.c_uint64(0)
        if nthreads == 0:
            nthreads = max(1, lib().omo_hardware_threads() - 1)  # num_cpus - 1, main.rs:170
        rc = lib().omo_render(self.h, C.byref(cam), C.byref(params), _ptr(pixels), _ptr(sumsq), spp_pass, nthreads,
                              C.byref(cr), C.byref(tr), counters)
        if rc != 0:
            raise ValueError("omo_render: invalid parameters")
        out = dict(pixels=pixels, credited=cr.value, traced=tr.value, threads=nthreads)
        if want_sumsq:
            out["sumsq"] = sumsq
        if want_counters:
            out["counters"] = dict(zip(["segments", "traced_tests", "sdf_evals", "normal_evals", "march_steps"], counters))
        return out


def view(pixels, width, height, mode):
    """draw_to_sdl's pixel buffer for one display mode (main.rs:361-437) from a Vec<Pixel>."""
    px = np.ascontiguousarray(pixels)
    rgb = np.zeros((height, width, 3), np.uint8)
    lib().omo_view(mode, px.ctypes.data, width, height, rgb.ctypes.data)
    return rgb


def primary_rays(cam, params, mode=0, sample_index=0):
    n = params.width * params.height
    o = np.zeros((n, 3), np.float32)
    d = np.zeros((n, 3), np.float32)
    lib().omo_primary_rays(C.byref(cam), C.byref(params), mode, sample_index, _ptr(o), _ptr(d))
    return o, d


def camera_new(lookfrom, lookat, vup, vfov, aspect, aperture, focus_dist):
    c = A.Camera()
    lib().omo_camera_new(C.byref(c), f3(lookfrom), f3(lookat), f3(vup), vfov, aspect, aperture, focus_dist)
    return c
