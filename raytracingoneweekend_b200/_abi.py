"""ctypes mirrors of the POD structs in include/ottomarcher.h.

Field order and types follow the header exactly (and through it the Rust structs it cites:
materials.rs:18-23, traced.rs:13-19/77-82/118-131/229-235, marched.rs:49-54/78-83/104-113,
camera.rs:9-29, render_thread.rs:9-17).  Nothing here touches CUDA.
"""
import ctypes as C

f32 = C.c_float
u8 = C.c_uint8
u32 = C.c_uint32
i32 = C.c_int32
u64 = C.c_uint64

OM_OK, OM_ERR_INVALID, OM_ERR_CUDA, OM_ERR_UNSUPPORTED, OM_ERR_OOM, OM_ERR_NO_DEVICE = range(6)
STATUS_NAMES = {0: "OM_OK", 1: "OM_ERR_INVALID", 2: "OM_ERR_CUDA", 3: "OM_ERR_UNSUPPORTED",
                4: "OM_ERR_OOM", 5: "OM_ERR_NO_DEVICE"}

MAT_LAMBERTIAN, MAT_METAL, MAT_DIELECTRIC = 0, 1, 2

TYPE_SPHERE, TYPE_CUBE, TYPE_TRIANGLE, TYPE_INFINITE_PLANE, TYPE_PARALLELOGRAM = 1, 2, 3, 4, 5
TYPE_MARCHED_SPHERE, TYPE_MARCHED_BOX, TYPE_MARCHED_TORUS = 6, 7, 8

RENDER_ADAPTIVE = 1 << 0
RENDER_FAST_MATH = 1 << 1
RENDER_BRUTE_FORCE = 1 << 2
RENDER_WAVEFRONT = 1 << 3
RENDER_PERSISTENT = 1 << 6
RENDER_PHASE_ALIGN = 1 << 7
RENDER_NO_PHASE_ALIGN = 1 << 8
RENDER_POOL = 1 << 9
FREEZE_NO_ALWAYS_LIST, FREEZE_RAY_PADDING, FREEZE_STATIC_PADDING, FREEZE_NO_MARCHED_BVH = 1, 2, 4, 8
FREEZE_GPU_BVH = 16
RENDER_NO_SMEM_SCENE = 1 << 4
RENDER_COUNTERS = 1 << 5

CHUNK_PIXELS = 2730  # main.rs:174-175
VIEW_NOALLOC_MAX_PIXELS = 2097152  # OM_VIEW_NOALLOC_MAX_PIXELS


def obj_id(type_tag: int, index: int) -> int:
    """(type << 32) | (index + 1); 0 is the sky (render_thread.rs:122)."""
    return (type_tag << 32) | (index + 1)


class Material(C.Structure):
    _fields_ = [("albedo", f32 * 3), ("fuzz", f32), ("ior", f32), ("mat_type", u32)]


class Sphere(C.Structure):
    _fields_ = [("m_local_to_world", f32 * 16), ("m_world_to_local", f32 * 16), ("material", Material)]


class Cube(C.Structure):
    _fields_ = [("m_local_to_world", f32 * 16), ("m_world_to_local", f32 * 16), ("material", Material)]


class InfinitePlane(C.Structure):
    _fields_ = [("center", f32 * 3), ("normal", f32 * 3), ("material", Material)]


class Barycentric(C.Structure):
    _fields_ = [("origin", f32 * 3), ("u", f32 * 3), ("u_length", f32), ("v", f32 * 3), ("v_length", f32),
                ("uxv", f32 * 3), ("uxvxu", f32 * 3), ("base_inv", f32 * 9), ("v_in_base", f32 * 3),
                ("material", Material)]


class MarchedSphere(C.Structure):
    _fields_ = [("center", f32 * 3), ("radius", f32), ("material", Material)]


class MarchedBox(C.Structure):
    _fields_ = [("center", f32 * 3), ("sizes", f32 * 3), ("material", Material)]


class MarchedTorus(C.Structure):
    _fields_ = [("m_local_to_world_translate_rotate", f32 * 16), ("m_world_to_local_translate_rotate", f32 * 16),
                ("m_local_to_world_scale", f32 * 4), ("m_world_to_local_scale", f32 * 4), ("sizes", f32 * 3),
                ("material", Material)]


class DynMarched(C.Structure):  # om_dyn_marched: Arc<dyn Marched> as data
    _fields_ = [("m_world_to_local", f32 * 16), ("m_local_to_world", f32 * 16), ("world_f", f32),
                ("bounds_lo", f32 * 3), ("bounds_hi", f32 * 3), ("material", Material)]


class DynTraced(C.Structure):  # om_dyn_traced: Arc<dyn Traced> as data (a quadric in a local frame, optionally clipped)
    _fields_ = [("m_local_to_world", f32 * 16), ("m_world_to_local", f32 * 16), ("q", f32 * 10), ("clip_lo", f32 * 3), ("clip_hi", f32 * 3),
                ("bounds_lo", f32 * 3), ("bounds_hi", f32 * 3), ("material", Material)]


# local_sdf program opcodes (OM_SDF_*)
SDF_OPS = {"x": 1, "y": 2, "z": 3, "const": 4, "add": 5, "sub": 6, "mul": 7, "div": 8, "min": 9, "max": 10, "neg": 11, "abs": 12,
           "sqrt": 13, "len2": 14, "len3": 15, "store": 16, "load": 17}
SDF_STACK, SDF_REGS, SDF_MAX_WORDS = 8, 4, 256
TYPE_DYN_MARCHED = 9
TYPE_DYN_TRACED = 10


class Camera(C.Structure):
    _fields_ = [("origin", f32 * 3), ("horizontal", f32 * 3), ("vertical", f32 * 3), ("lower_left_corner", f32 * 3),
                ("u_of_plane", f32 * 3), ("v_of_plane", f32 * 3), ("w_of_plane", f32 * 3), ("lens_radius", f32),
                ("aspect_ratio", f32), ("focus_dist", f32), ("viewport_width", f32), ("viewport_height", f32)]


class Pixel(C.Structure):
    _fields_ = [("sum", f32 * 3), ("n", u32), ("avg_depth", f32), ("bad_avgs", u32), ("bloom", u64),
                ("color", u8 * 3), ("_pad", u8 * 5)]


class RenderParams(C.Structure):
    _fields_ = [("width", u32), ("height", u32), ("spp_max", u32), ("max_depth", u32), ("tmin", f32), ("tmax", f32),
                ("seed", u64), ("flags", u32), ("sample_base", u32), ("sample_count", u32), ("part_rank", u32),
                ("part_world", u32), ("device", i32)]


assert C.sizeof(Material) == 24
assert C.sizeof(Sphere) == 152
assert C.sizeof(Pixel) == 40
assert C.sizeof(Camera) == 104
assert C.sizeof(DynMarched) == 180
assert C.sizeof(DynTraced) == 240

# numpy view of om_pixel (same layout) for vectorised test/bench code
PIXEL_DTYPE = [("sum", "<f4", (3,)), ("n", "<u4"), ("avg_depth", "<f4"), ("bad_avgs", "<u4"), ("bloom", "<u8"),
               ("color", "u1", (3,)), ("_pad", "u1", (5,))]
