"""Host-side mirror of the reference's scene / camera / render interface, over the C ABI.

Names and argument meaning follow the Rust reference so that tests read like its own code would:

    world = HittableList()                       # hits.rs:72
    world += Sphere.new_with_radius(...)         # hits.rs:101-105 (`world += &obj`)
    cam = Camera.new(lookfrom, lookat, vup, vfov, aspect, aperture, focus_dist)   # camera.rs:38
    frozen = world.freeze(cam)                   # hits.rs:87-89
    pixels = render(cam, frozen, max_depth, tmin, tmax, spp, W, H)    # render_thread.rs:145

Everything computes in libottomarcher_cuda.so; nothing here has a CPU path.
"""
import ctypes as C

import numpy as np

from . import _abi as A
from . import _lib
from ._lib import OmError  # noqa: F401  (re-export)


def _f3(v):
    return (C.c_float * 3)(*[float(x) for x in v])


def _f16(m):
    flat = [float(x) for row in m for x in (row if hasattr(row, "__iter__") else [row])] if len(m) != 16 else [float(x) for x in m]
    return (C.c_float * 16)(*flat)


# ---- mat4x4.rs:88-123,167-209: the m4x4! helpers; `^` is matrix product ----------------------
class Mat4x4:
    def __init__(self, data=None):
        self.e = (C.c_float * 16)()
        if data is None:
            _lib.load().om_mat4_identity(self.e)
        else:
            self.e[:] = [float(x) for x in data]

    IDENTITY = None

    @staticmethod
    def new_translate(v):
        m = Mat4x4(); _lib.load().om_mat4_translate(m.e, v[0], v[1], v[2]); return m

    @staticmethod
    def new_scale(v):
        m = Mat4x4(); _lib.load().om_mat4_scale(m.e, v[0], v[1], v[2]); return m

    @staticmethod
    def new_rotate_x(f):
        m = Mat4x4(); _lib.load().om_mat4_rotate_x(m.e, f); return m

    @staticmethod
    def new_rotate_y(f):
        m = Mat4x4(); _lib.load().om_mat4_rotate_y(m.e, f); return m

    @staticmethod
    def new_rotate_z(f):
        m = Mat4x4(); _lib.load().om_mat4_rotate_z(m.e, f); return m

    def dot_mat(self, other):
        m = Mat4x4(); _lib.load().om_mat4_mul(m.e, self.e, other.e); return m

    __xor__ = dot_mat  # mat4x4.rs:167-178

    def fast_homogenous_inverse(self):
        m = Mat4x4(); _lib.load().om_mat4_fast_homogenous_inverse(m.e, self.e); return m

    def numpy(self):
        return np.array(self.e[:], np.float32).reshape(4, 4)


# ---- materials.rs:25-38 -----------------------------------------------------------------------
class Material:
    @staticmethod
    def new_lambertian(albedo):
        m = A.Material(); _lib.load().om_material_lambertian(C.byref(m), _f3(albedo)); return m

    @staticmethod
    def new_metal(albedo):
        m = A.Material(); _lib.load().om_material_metal(C.byref(m), _f3(albedo), 0.0); return m

    @staticmethod
    def new_metal_fuzz(albedo, fuzz):
        m = A.Material(); _lib.load().om_material_metal(C.byref(m), _f3(albedo), fuzz); return m

    @staticmethod
    def new_dielectric(ior):
        m = A.Material(); _lib.load().om_material_dielectric(C.byref(m), ior); return m


# ---- traced.rs / marched.rs constructors --------------------------------------------------------
class Sphere:
    @staticmethod
    def new(m_local_to_world, mat):
        s = A.Sphere(); _lib.load().om_sphere_new(C.byref(s), m_local_to_world.e, C.byref(mat)); return s

    @staticmethod
    def new_with_radius(o, r, mat):
        s = A.Sphere(); _lib.load().om_sphere_new_with_radius(C.byref(s), _f3(o), r, C.byref(mat)); return s


class Cube:
    @staticmethod
    def new(m_local_to_world, mat):
        s = A.Cube(); _lib.load().om_cube_new(C.byref(s), m_local_to_world.e, C.byref(mat)); return s

    @staticmethod
    def new_with_length(o, length, mat):
        s = A.Cube(); _lib.load().om_cube_new_with_length(C.byref(s), _f3(o), length, C.byref(mat)); return s


class InfinitePlane:
    @staticmethod
    def new(center, normal, mat):
        s = A.InfinitePlane(); _lib.load().om_infinite_plane_new(C.byref(s), _f3(center), _f3(normal), C.byref(mat)); return s


class _Tagged:
    """A Barycentric<BT> value together with its alias (Triangle / Parallelogram, traced.rs:225-226)."""

    def __init__(self, kind, pod):
        self.kind, self.pod = kind, pod


class Triangle:
    @staticmethod
    def new3points(origin, upoint, vpoint, mat):
        s = A.Barycentric(); _lib.load().om_barycentric_new3points(C.byref(s), _f3(origin), _f3(upoint), _f3(vpoint), C.byref(mat))
        return _Tagged("triangles", s)


class Parallelogram:
    @staticmethod
    def new3points(origin, upoint, vpoint, mat):
        s = A.Barycentric(); _lib.load().om_barycentric_new3points(C.byref(s), _f3(origin), _f3(upoint), _f3(vpoint), C.byref(mat))
        return _Tagged("parallelograms", s)


def MarchedSphere(center, radius, material):
    return A.MarchedSphere(_f3(center), radius, material)


def MarchedBox(center, sizes, material):
    return A.MarchedBox(_f3(center), _f3(sizes), material)


class MarchedTorus:
    @staticmethod
    def new(m_local_to_world, local_sizes, mat):
        s = A.MarchedTorus(); _lib.load().om_marched_torus_new(C.byref(s), m_local_to_world.e, _f3(local_sizes), C.byref(mat)); return s


def sdf_program(tokens):
    """Assemble a local_sdf program (include/ottomarcher.h, OM_SDF_*): tokens are "x" "y" "z", a number (pushed as an
    f32 constant), "add" "sub" "mul" "div" "min" "max" "neg" "abs" "sqrt" "len2" "len3", ("store", r), ("load", r)."""
    import struct
    words = []
    for t in tokens:
        if isinstance(t, (int, float)):
            words += [A.SDF_OPS["const"], struct.unpack("<I", struct.pack("<f", float(t)))[0]]
        elif isinstance(t, tuple):
            words.append(A.SDF_OPS[t[0]] | (int(t[1]) << 8))
        else:
            words.append(A.SDF_OPS[t])
    return words


class DynMarched:
    """A user's `impl Marched` (marched.rs:8-45) as data: `world += Arc::new(obj) as Arc<dyn Marched + Send + Sync>`
    (hits.rs:96-100).  to_local / to_world are matrices, to_world_f a factor, local_sdf a program (sdf_program)."""

    def __init__(self, m_world_to_local, m_local_to_world, world_f, code, material, bounds=None):
        inf = float("inf")
        lo, hi = bounds if bounds is not None else ((-inf,) * 3, (inf,) * 3)
        self.pod = A.DynMarched(m_world_to_local.e, m_local_to_world.e, world_f, _f3(lo), _f3(hi), material)
        self.code = [int(w) for w in code]


class DynTraced:
    """A user's `impl Traced` (traced.rs:33-35) as data: `world += Arc::new(obj) as Arc<dyn Traced + Send + Sync>` (hits.rs:91-95).
    The object is a quadric q = (A, B, C, D, E, F, G, H, I, J) in its own frame (see om_dyn_traced in ottomarcher.h), optionally cut
    by a box of that frame; `bounds` is its world box (None = the trait's "draw always")."""

    def __init__(self, m_local_to_world, q, material, clip=None, bounds=None):
        inf = float("inf")
        clo, chi = clip if clip is not None else ((-inf,) * 3, (inf,) * 3)
        blo, bhi = bounds if bounds is not None else ((-inf,) * 3, (inf,) * 3)
        self.pod = A.DynTraced(m_local_to_world.e, m_local_to_world.fast_homogenous_inverse().e, (C.c_float * 10)(*[float(x) for x in q]),
                               _f3(clo), _f3(chi), _f3(blo), _f3(bhi), material)

    @staticmethod
    def world_box(m_local_to_world, local_lo, local_hi):
        """BoundingBox3D::new(lo, hi).dot(m) (bounding_box.rs:26-47): the box of the 8 transformed corners."""
        m = np.array(m_local_to_world.e[:], np.float64).reshape(4, 4)
        c = np.array([[x, y, z, 1.0] for x in (local_lo[0], local_hi[0]) for y in (local_lo[1], local_hi[1]) for z in (local_lo[2], local_hi[2])])
        w = (m @ c.T).T[:, :3]
        return tuple(w.min(axis=0)), tuple(w.max(axis=0))

    @staticmethod
    def ellipsoid(m_local_to_world, material, bounded=True):
        """The unit sphere under an affine map: the same surface, and the same bits, as Sphere::new(m, material)."""
        b = DynTraced.world_box(m_local_to_world, (-1, -1, -1), (1, 1, 1)) if bounded else None
        return DynTraced(m_local_to_world, (1, 1, 1, 0, 0, 0, 0, 0, 0, -1), material, bounds=b)

    @staticmethod
    def cylinder(m_local_to_world, material, half_height=1.0):
        """x^2 + z^2 = 1 for |y| <= half_height (an open tube)."""
        h = float(half_height)
        return DynTraced(m_local_to_world, (1, 0, 1, 0, 0, 0, 0, 0, 0, -1), material, clip=((-2, -h, -2), (2, h, 2)),
                         bounds=DynTraced.world_box(m_local_to_world, (-1, -h, -1), (1, h, 1)))


class Camera:
    @staticmethod
    def new(lookfrom, lookat, vup, vfov_in_degrees, aspect_ratio, aperture, focus_dist):
        c = A.Camera()
        _lib.load().om_camera_new(C.byref(c), _f3(lookfrom), _f3(lookat), _f3(vup), vfov_in_degrees, aspect_ratio, aperture, focus_dist)
        return c

    @staticmethod
    def world_camera(vfov_in_degrees, aspect_ratio):  # camera.rs:33-35
        return Camera.new((0, 0, 0), (0, 0, -1), (0, 1, 0), vfov_in_degrees, aspect_ratio, 0.0, 1.0)


SCENE_KINDS = {"random_scene": 0, "basic_scene": 1, "rtiow_classic": 2, "marched": 3, "glass_heavy": 4, "many_spheres": 5}

_KIND_OF = {A.Sphere: "spheres", A.Cube: "cubes", A.InfinitePlane: "infinite_planes", A.MarchedSphere: "marched_spheres",
            A.MarchedBox: "marched_boxes", A.MarchedTorus: "marched_tori"}
_LISTS = ["spheres", "cubes", "triangles", "infinite_planes", "parallelograms", "marched_spheres", "marched_boxes", "marched_tori"]


class HittableList:
    """hits.rs:37-110.  Objects are kept host-side per type (so tests can hand the same arrays to the
    oracle) and pushed through om_scene_add_* at freeze()."""

    def __init__(self):
        self.lists = {k: [] for k in _LISTS}
        self.dyn_marched = []  # marched_objects: Vec<Arc<dyn Marched>> (DynMarched)
        self.dyn_traced = []   # traced_objects: Vec<Arc<dyn Traced>> (DynTraced)
        self._generated = []  # (kind, seed, n) requests replayed into the C scene

    def clear(self):
        self.__init__()

    def __iadd__(self, obj):
        if isinstance(obj, _Tagged):
            self.lists[obj.kind].append(obj.pod)
        elif type(obj) in _KIND_OF:
            self.lists[_KIND_OF[type(obj)]].append(obj)
        elif isinstance(obj, DynMarched):
            self.dyn_marched.append(obj)
        elif isinstance(obj, DynTraced):
            self.dyn_traced.append(obj)
        else:
            # native trait objects (compiled Traced / Marched impls that do not describe themselves as data): the ABI reports OM_ERR_UNSUPPORTED
            h = C.c_void_p()
            _lib.call("om_scene_create", C.byref(h))
            try:
                _lib.call("om_scene_add_dyn_object", h, None)
            finally:
                _lib.load().om_scene_destroy(h)
        return self

    def arrays(self):
        """ctypes arrays per list (the bytes that cross the boundary)."""
        out = {}
        for k, v in self.lists.items():
            T = type(v[0]) if v else None
            out[k] = (T * len(v))(*v) if v else None
        return out

    def freeze(self, cam=None, flags=0):
        s = FrozenHittableList()
        for k, arr in self.arrays().items():
            if arr is not None:
                _lib.call("om_scene_add_" + k, s.h, arr, len(arr))
        for d in self.dyn_marched:
            code = (C.c_uint32 * max(1, len(d.code)))(*d.code)
            _lib.call("om_scene_add_dyn_marched", s.h, C.byref(d.pod), code, len(d.code))
        if self.dyn_traced:
            arr = (A.DynTraced * len(self.dyn_traced))(*[d.pod for d in self.dyn_traced])
            _lib.call("om_scene_add_dyn_traced", s.h, arr, len(arr))
        _lib.call("om_scene_freeze", s.h, C.byref(cam) if cam is not None else None, flags)
        return s


class FrozenHittableList:
    """hits.rs:63-69, as an opaque om_scene handle."""

    def __init__(self):
        self.h = C.c_void_p()
        _lib.call("om_scene_create", C.byref(self.h))

    def __del__(self):
        if getattr(self, "h", None) and _lib._lib is not None:
            _lib._lib.om_scene_destroy(self.h)
            self.h = None

    @staticmethod
    def generate(kind, seed=1, n_param=0, flags=0):
        """Seeded generator scenes, frozen."""
        s = FrozenHittableList()
        _lib.call("om_scene_generate", s.h, SCENE_KINDS[kind] if isinstance(kind, str) else kind, seed, n_param)
        _lib.call("om_scene_freeze", s.h, None, flags)
        return s

    def counts(self):
        c = (C.c_uint64 * 8)()
        _lib.call("om_scene_counts", self.h, c)
        return dict(zip(_LISTS, c))

    def lists(self):
        """The scene's objects as ctypes arrays per list (the bytes that crossed the boundary)."""
        types = [A.Sphere, A.Cube, A.Barycentric, A.InfinitePlane, A.Barycentric, A.MarchedSphere, A.MarchedBox, A.MarchedTorus]
        out = {}
        for i, (name, n) in enumerate(self.counts().items()):
            arr = (types[i] * n)()
            _lib.call("om_scene_get_list", self.h, i, arr, C.sizeof(arr))
            out[name] = arr
        return out

    def dyn_traced(self):
        """ctypes array of the om_dyn_traced objects in list order."""
        n = C.c_uint64()
        _lib.call("om_scene_dyn_traced_count", self.h, C.byref(n))
        arr = (A.DynTraced * n.value)()
        _lib.call("om_scene_get_list", self.h, 8, arr, C.sizeof(arr))
        return arr

    def dyn_marched(self):
        """[(om_dyn_marched, program words)] in list order."""
        n = C.c_uint64()
        _lib.call("om_scene_dyn_marched_count", self.h, C.byref(n))
        out = []
        for i in range(n.value):
            pod, code, nw = A.DynMarched(), (C.c_uint32 * A.SDF_MAX_WORDS)(), C.c_size_t(A.SDF_MAX_WORDS)
            _lib.call("om_scene_get_dyn_marched", self.h, i, C.byref(pod), code, C.byref(nw))
            out.append((pod, list(code[:nw.value])))
        return out

    def stats(self):
        s = (C.c_uint64 * 8)()
        _lib.call("om_scene_stats", self.h, s)
        d = dict(zip(["bvh_nodes", "bvh_depth", "bvh_prims", "always_test", "device_bytes", "fits_smem", "mgrid_cells", "mgrid_entries"], s))
        d["ray_padding"] = (d["fits_smem"] >> 1) & 1
        d["fits_smem"] &= 1
        return d

    def hit(self, origins, dirs, t_min=0.001, t_max=100.0, flags=0, device=-1):
        """FrozenHittableList::hit (hits.rs:270-334) for a batch of rays -> (obj_id, t, point, normal)."""
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        n = o.shape[0]
        ids = np.zeros(n, np.uint64); t = np.zeros(n, np.float32)
        pts = np.zeros((n, 3), np.float32); nrm = np.zeros((n, 3), np.float32)
        _lib.call("om_trace_rays", self.h, o.ctypes.data, d.ctypes.data, n, t_min, t_max, flags, device,
                  ids.ctypes.data, t.ctypes.data, pts.ctypes.data, nrm.ctypes.data)
        return ids, t, pts, nrm


def render_params(width, height, spp, max_depth=50, tmin=0.001, tmax=100.0, seed=0, flags=0, sample_base=0, sample_count=0,
                  part_rank=0, part_world=1, device=-1):
    return A.RenderParams(width, height, spp, max_depth, tmin, tmax, seed, flags, sample_base, sample_count, part_rank, part_world, device)


class Renderer:
    """The N render threads + their shared Vec<Pixel> (main.rs:192-214) as one om_render handle."""

    def __init__(self, frozen, cam, params):
        self.frozen, self.cam, self.params = frozen, cam, params
        self.h = C.c_void_p()
        _lib.call("om_render_create", C.byref(self.h), frozen.h, C.byref(cam), C.byref(params))
        self.npx = params.width * params.height

    def close(self):
        if getattr(self, "h", None) and _lib._lib is not None:
            _lib._lib.om_render_destroy(self.h)
            self.h = None

    __del__ = close

    def set_stream(self, cuda_stream_ptr):
        """Issue this handle's work on a caller-owned stream (e.g. torch.cuda.current_stream().cuda_stream)."""
        _lib.call("om_render_set_stream", self.h, cuda_stream_ptr)

    def launch(self, spp_this_pass=0):
        _lib.call("om_render_launch", self.h, spp_this_pass)

    def sync(self):
        _lib.call("om_render_sync", self.h)

    def reset(self):
        _lib.call("om_render_reset", self.h)

    def samples_done(self):
        v = C.c_uint64(); _lib.call("om_render_samples_done", self.h, C.byref(v)); return v.value

    def samples_traced(self):
        v = C.c_uint64(); _lib.call("om_render_samples_traced", self.h, C.byref(v)); return v.value

    def perf_counters(self):
        v = (C.c_uint64 * 16)(); _lib.call("om_render_perf_counters", self.h, v)
        names = ["credited", "traced", "_ticket", "segments", "node_visits", "prim_tests", "sdf_evals", "march_steps", "normal_evals", "march_entered",
                 "kernels", "flags", "scene_in_smem", "warp_iters", "lane_iters", "graph_launches"]
        return dict(zip(names, v))

    def elapsed_ms(self):
        ms, n = C.c_float(), C.c_uint32(); _lib.call("om_render_elapsed_ms", self.h, C.byref(ms), C.byref(n)); return ms.value, n.value

    def read_pixels(self, out=None):
        px = np.zeros(self.npx, dtype=A.PIXEL_DTYPE) if out is None else out
        _lib.call("om_render_read_pixels", self.h, px.ctypes.data, self.npx)
        return px

    def read_rgb(self):
        rgb = np.zeros((self.params.height, self.params.width, 3), np.uint8)
        _lib.call("om_render_read_rgb", self.h, rgb.ctypes.data, rgb.nbytes)
        return rgb

    def read_view(self, mode):
        """One refresh of draw_to_sdl's buffer (main.rs:361-437): mode 0 normal, 1 samples, 2 sample-weighted
        blur, 3 depth, 4 depth-weighted blur, 5 ids, 6 id-weighted blur."""
        rgb = np.zeros((self.params.height, self.params.width, 3), np.uint8)
        _lib.call("om_render_read_view", self.h, mode, rgb.ctypes.data, rgb.nbytes)
        return rgb

    def read_view_noalloc(self, mode, out=None):
        """read_view through om_render_read_view_noalloc: fixed device scratch, no allocation in the library (no
        counterpart in the reference; at most _abi.VIEW_NOALLOC_MAX_PIXELS pixels).  `out` lets the caller reuse a buffer."""
        rgb = np.zeros((self.params.height, self.params.width, 3), np.uint8) if out is None else out
        _lib.call("om_render_read_view_noalloc", self.h, mode, rgb.ctypes.data, rgb.nbytes)
        return rgb

    def primary(self, mode=0, sample_index=0):
        ids = np.zeros(self.npx, np.uint64); t = np.zeros(self.npx, np.float32)
        _lib.call("om_render_primary", self.h, mode, sample_index, ids.ctypes.data, t.ctypes.data, self.npx)
        return ids, t

    def export_accum(self, dev_acc_ptr, dev_bloom_ptr):
        _lib.call("om_render_export_accum", self.h, dev_acc_ptr, dev_bloom_ptr)

    def bloom_or(self, dev_dst_ptr, dev_src_ptr, nsrc):
        """dst[i] |= src[k][i] for k < nsrc over the frame's masks, on the handle's stream (no synchronisation)."""
        _lib.call("om_render_bloom_or", self.h, dev_dst_ptr, dev_src_ptr, nsrc)

    def import_accum(self, dev_acc_ptr, dev_bloom_ptr):
        _lib.call("om_render_import_accum", self.h, dev_acc_ptr, dev_bloom_ptr)

    # ---- adaptive sampling across GPUs (dist.AdaptiveScheduler) ----
    @property
    def n_chunks(self):
        return (self.npx + A.CHUNK_PIXELS - 1) // A.CHUNK_PIXELS

    def set_chunk_owners(self, owners, my_id):
        """owners: uint8 array, one entry per 2730-pixel chunk (None: back to the static partition)."""
        if owners is None:
            _lib.call("om_render_set_chunk_owners", self.h, None, 0, 0)
            return
        o = np.ascontiguousarray(owners, np.uint8)
        _lib.call("om_render_set_chunk_owners", self.h, o.ctypes.data, o.shape[0], my_id)

    def export_state(self, dev_state_ptr, only_owned=True):
        _lib.call("om_render_export_state", self.h, dev_state_ptr, 1 if only_owned else 0)

    def import_state(self, dev_state_ptr):
        _lib.call("om_render_import_state", self.h, dev_state_ptr)

    def live_per_chunk(self):
        c = np.zeros(self.n_chunks, np.uint32)
        _lib.call("om_render_live_per_chunk", self.h, c.ctypes.data, c.shape[0])
        return c


def render(camera, world, max_depth, tmin, tmax, samples_per_pixel, image_width, image_height, seed=0, flags=A.RENDER_ADAPTIVE, device=-1):
    """render_thread::render (render_thread.rs:145-147) for all threads at once; returns the Vec<Pixel>
    as a numpy structured array and the progress counter.  The default flags keep the reference's
    adaptive retirement; pass flags=0 for fixed spp."""
    p = render_params(image_width, image_height, samples_per_pixel, max_depth, tmin, tmax, seed, flags, device=device)
    r = Renderer(world, camera, p)
    try:
        r.launch(0)
        r.sync()
        return r.read_pixels(), r.samples_done()
    finally:
        r.close()


def write_bmp(path, rgb):
    """F12 of the viewer (main.rs:473-476): save an RGB24 image as .bmp."""
    rgb = np.ascontiguousarray(rgb, np.uint8)
    _lib.call("om_write_bmp", str(path).encode(), rgb.ctypes.data, rgb.shape[1], rgb.shape[0])


def bloom_get_hash(obj_id):
    return _lib.load().om_bloom_get_hash(obj_id)


def jitter_table(spp, seed):
    out = np.zeros(spp, np.uint8)
    _lib.load().om_jitter_table(spp, seed, out.ctypes.data)
    return out


def fp32_peak(device=-1):
    tf, ms = C.c_float(), C.c_float()
    _lib.call("om_bench_fp32_peak", device, C.byref(tf), C.byref(ms))
    return tf.value, ms.value
