"""ctypes loader of libottomarcher_cuda.so — the C ABI declared in include/ottomarcher.h.

The library is the product; there is no Python or CPU implementation behind it.  If the shared
object is missing this module raises ImportError with the build command instead of degrading.
"""
import ctypes as C
import os

from . import _abi as A

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("OM_LIB_PATH") or os.path.join(HERE, "libottomarcher_cuda.so")  # OM_LIB_PATH: development builds

_lib = None

vp, u8p = C.c_void_p, C.POINTER(C.c_uint8)
u32, i32, u64, f32, sz = C.c_uint32, C.c_int32, C.c_uint64, C.c_float, C.c_size_t
P = C.POINTER
F3, F16 = P(f32), P(f32)

# name -> (restype, argtypes); restype "st" = om_status
PROTOTYPES = {
    "om_scene_create": ("st", [P(vp)]),
    "om_scene_clear": ("st", [vp]),
    "om_scene_destroy": ("st", [vp]),
    "om_trim_device_cache": ("st", []),
    "om_scene_add_spheres": ("st", [vp, P(A.Sphere), sz]),
    "om_scene_add_cubes": ("st", [vp, P(A.Cube), sz]),
    "om_scene_add_triangles": ("st", [vp, P(A.Barycentric), sz]),
    "om_scene_add_infinite_planes": ("st", [vp, P(A.InfinitePlane), sz]),
    "om_scene_add_parallelograms": ("st", [vp, P(A.Barycentric), sz]),
    "om_scene_add_marched_spheres": ("st", [vp, P(A.MarchedSphere), sz]),
    "om_scene_add_marched_boxes": ("st", [vp, P(A.MarchedBox), sz]),
    "om_scene_add_marched_tori": ("st", [vp, P(A.MarchedTorus), sz]),
    "om_scene_add_dyn_object": ("st", [vp, vp]),
    "om_scene_add_dyn_marched": ("st", [vp, P(A.DynMarched), P(u32), sz]),
    "om_scene_add_dyn_traced": ("st", [vp, vp, sz]),
    "om_scene_dyn_traced_count": ("st", [vp, vp]),
    "om_scene_dyn_marched_count": ("st", [vp, P(u64)]),
    "om_scene_get_dyn_marched": ("st", [vp, u64, P(A.DynMarched), P(u32), P(sz)]),
    "om_scene_counts": ("st", [vp, P(u64)]),
    "om_scene_get_list": ("st", [vp, u32, vp, sz]),
    "om_scene_freeze": ("st", [vp, P(A.Camera), u32]),
    "om_scene_stats": ("st", [vp, P(u64)]),
    "om_scene_generate": ("st", [vp, u32, u64, u64]),
    "om_render_create": ("st", [P(vp), vp, P(A.Camera), P(A.RenderParams)]),
    "om_render_destroy": ("st", [vp]),
    "om_render_set_stream": ("st", [vp, vp]),
    "om_render_launch": ("st", [vp, u32]),
    "om_render_sync": ("st", [vp]),
    "om_render_samples_done": ("st", [vp, P(u64)]),
    "om_render_samples_traced": ("st", [vp, P(u64)]),
    "om_render_perf_counters": ("st", [vp, P(u64)]),
    "om_render_elapsed_ms": ("st", [vp, P(f32), P(u32)]),
    "om_render_read_pixels": ("st", [vp, vp, sz]),
    "om_render_read_rgb": ("st", [vp, vp, sz]),
    "om_render_read_view": ("st", [vp, u32, vp, sz]),
    "om_render_read_view_noalloc": ("st", [vp, u32, vp, sz]),
    "om_write_bmp": ("st", [C.c_char_p, vp, u32, u32]),
    "om_render_reset": ("st", [vp]),
    "om_render_primary": ("st", [vp, u32, u32, vp, vp, sz]),
    "om_render_export_accum": ("st", [vp, vp, vp]),
    "om_render_import_accum": ("st", [vp, vp, vp]),
    "om_bloom_or": ("st", [vp, vp, sz, sz, i32]),
    "om_render_bloom_or": ("st", [vp, vp, vp, sz]),
    "om_render_set_chunk_owners": ("st", [vp, vp, sz, u32]),
    "om_render_export_state": ("st", [vp, vp, u32]),
    "om_render_import_state": ("st", [vp, vp]),
    "om_render_live_per_chunk": ("st", [vp, vp, sz]),
    "om_trace_rays": ("st", [vp, vp, vp, sz, f32, f32, u32, i32, vp, vp, vp, vp]),
    "om_camera_new": (None, [P(A.Camera), F3, F3, F3, f32, f32, f32, f32]),
    "om_material_lambertian": (None, [P(A.Material), F3]),
    "om_material_metal": (None, [P(A.Material), F3, f32]),
    "om_material_dielectric": (None, [P(A.Material), f32]),
    "om_sphere_new": (None, [P(A.Sphere), F16, P(A.Material)]),
    "om_sphere_new_with_radius": (None, [P(A.Sphere), F3, f32, P(A.Material)]),
    "om_cube_new": (None, [P(A.Cube), F16, P(A.Material)]),
    "om_cube_new_with_length": (None, [P(A.Cube), F3, f32, P(A.Material)]),
    "om_infinite_plane_new": (None, [P(A.InfinitePlane), F3, F3, P(A.Material)]),
    "om_barycentric_new3points": (None, [P(A.Barycentric), F3, F3, F3, P(A.Material)]),
    "om_marched_torus_new": (None, [P(A.MarchedTorus), F16, F3, P(A.Material)]),
    "om_mat4_identity": (None, [F16]),
    "om_mat4_translate": (None, [F16, f32, f32, f32]),
    "om_mat4_scale": (None, [F16, f32, f32, f32]),
    "om_mat4_rotate_x": (None, [F16, f32]),
    "om_mat4_rotate_y": (None, [F16, f32]),
    "om_mat4_rotate_z": (None, [F16, f32]),
    "om_mat4_mul": (None, [F16, F16, F16]),
    "om_mat4_fast_homogenous_inverse": (None, [F16, F16]),
    "om_bloom_get_hash": (u64, [u64]),
    "om_jitter_table": (None, [u32, u64, vp]),
    "om_last_error": (C.c_char_p, []),
    "om_version": (C.c_char_p, []),
    "om_device_count": (i32, []),
    "om_bench_fp32_peak": ("st", [i32, P(f32), P(f32)]),
}


class OmError(RuntimeError):
    def __init__(self, status, message):
        super().__init__("%s: %s" % (A.STATUS_NAMES.get(status, status), message))
        self.status = status


def load():
    """Load the shared library (once) and bind every prototype."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "%s is missing: the CUDA library is the only implementation of this package (no CPU fallback). "
            "Build it with `python -m raytracingoneweekend_b200.build` or __graft_entry__.build()." % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(L, name)  # AttributeError here means header and library disagree
        fn.argtypes = args
        fn.restype = C.c_int if res == "st" else res
    _lib = L
    return L


def check(status):
    if status != A.OM_OK:
        raise OmError(status, load().om_last_error().decode("utf-8", "replace"))


def call(name, *args):
    """Call a status-returning entry point and raise OmError on failure."""
    check(getattr(load(), name)(*args))
