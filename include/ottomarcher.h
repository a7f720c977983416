/*
 * ottomarcher.h — C ABI of the B200 (sm_100a) renderer that replaces the CPU render threads of
 * octaviogarcia/RaytracingOneWeekend ("ottomarcher").
 *
 * The seam it replaces is the free function
 *     render_thread::render(&cam,&wrld,max_depth,tmin,tmax,samples_per_pixel,image_width,
 *                           image_height,pixels_box,tid,&assigned_thread,&samples_atom)
 * (reference src/render_thread.rs:145-147, called once per CPU thread from src/main.rs:207-213).
 * One om_render handle stands for ALL of those threads: `tid`/`assigned_thread` become the
 * device/partition fields of om_render_params.
 *
 * Every struct is a plain-old-data mirror of the Rust struct named beside it (f32/u32/u64 only,
 * natural alignment, no padding surprises) so that a `#[repr(C)]` conversion on the Rust side is
 * mechanical (see INTEGRATION.md and rust/ottomarcher-cuda-sys).  Matrices are row-major, i.e.
 * m[4*row+col] == Mat4x4.e[row].e[col] (reference src/math/mat4x4.rs:7-10).
 *
 * Nothing in this header needs CUDA, C++ or torch types.
 */
#ifndef OTTOMARCHER_H
#define OTTOMARCHER_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define OM_API __declspec(dllexport)
#else
#define OM_API __attribute__((visibility("default")))
#endif

/* ------------------------------------------------------------------------------------------ */
/* status codes (the reference panics; nothing here aborts or throws across the boundary)      */
typedef enum om_status {
    OM_OK = 0,
    OM_ERR_INVALID = 1,     /* bad argument / bad handle state                                 */
    OM_ERR_CUDA = 2,        /* a CUDA runtime call failed; om_last_error() has the text         */
    OM_ERR_UNSUPPORTED = 3, /* e.g. Arc<dyn Traced>/Arc<dyn Marched> user types (hits.rs:39-40)  */
    OM_ERR_OOM = 4,
    OM_ERR_NO_DEVICE = 5    /* no CUDA device: there is NO CPU fallback in this library          */
} om_status;

/* ------------------------------------------------------------------------------------------ */
/* materials.rs:11-23  (enum MaterialType + struct Material, 24 bytes)                          */
enum { OM_MAT_LAMBERTIAN = 0, OM_MAT_METAL = 1, OM_MAT_DIELECTRIC = 2 };
typedef struct om_material {
    float albedo[3]; /* Lambertian, Metal */
    float fuzz;      /* Metal             */
    float ior;       /* Dielectric        */
    uint32_t mat_type;
} om_material;

/* traced.rs:13-19 Sphere: unit sphere under an affine map (ellipsoids allowed).               */
typedef struct om_sphere {
    float m_local_to_world[16];
    float m_world_to_local[16]; /* Mat4x4::fast_homogenous_inverse of the above (mat4x4.rs:59) */
    om_material material;
} om_sphere;

/* traced.rs:229-235 Cube: cube of half-size 0.5 under an affine map.                          */
typedef struct om_cube {
    float m_local_to_world[16];
    float m_world_to_local[16];
    om_material material;
} om_cube;

/* traced.rs:77-82 InfinitePlane                                                              */
typedef struct om_infinite_plane {
    float center[3];
    float normal[3]; /* unit */
    om_material material;
} om_infinite_plane;

/* traced.rs:118-131 Barycentric<BT>; Parallelogram = BT 0, Triangle = BT 1 (traced.rs:225-226) */
typedef struct om_barycentric {
    float origin[3];
    float u[3];
    float u_length;
    float v[3];
    float v_length;
    float uxv[3];
    float uxvxu[3];
    float base_inv[9]; /* Mat3x3 row-major */
    float v_in_base[3];
    om_material material;
} om_barycentric;

/* marched.rs:49-54 */
typedef struct om_marched_sphere {
    float center[3];
    float radius;
    om_material material;
} om_marched_sphere;

/* marched.rs:78-83 */
typedef struct om_marched_box {
    float center[3];
    float sizes[3];
    om_material material;
} om_marched_box;

/* marched.rs:104-113 */
typedef struct om_marched_torus {
    float m_local_to_world_translate_rotate[16];
    float m_world_to_local_translate_rotate[16];
    float m_local_to_world_scale[4];
    float m_world_to_local_scale[4];
    float sizes[3]; /* (R, r, unused) */
    om_material material;
} om_marched_torus;

/* `Arc<dyn Marched + Send + Sync>` as data (marched.rs:8-45; hits.rs:40,96-100,312-320,353-359): what a user
 * type supplies to the trait — material(), to_local(), to_world(), to_world_f(), local_sdf() — with
 * local_sdf written as a postfix program over f32 (one 32-bit word per instruction, opcode in the low 8
 * bits, argument above; OM_SDF_CONST is followed by one word holding the f32 bits).  Operand stack of
 * OM_SDF_STACK values, OM_SDF_REGS registers, every operation one IEEE f32 operation in the order written,
 * so the device and the CPU oracle evaluate the same bits.  sdf() and get_outward_normal() are the trait's
 * defaults (marched.rs:14-44): sdf(p) = to_world_f(local_sdf(to_local(p))), numeric normal with sign fix.   */
enum {
    OM_SDF_X = 1, OM_SDF_Y = 2, OM_SDF_Z = 3, /* push a component of the local point                     */
    OM_SDF_CONST = 4,                          /* push the f32 whose bits are the next word              */
    OM_SDF_ADD = 5, OM_SDF_SUB = 6, OM_SDF_MUL = 7, OM_SDF_DIV = 8, /* b = pop, a = pop, push a op b      */
    OM_SDF_MIN = 9, OM_SDF_MAX = 10,           /* f32::min / f32::max                                    */
    OM_SDF_NEG = 11, OM_SDF_ABS = 12, OM_SDF_SQRT = 13,
    OM_SDF_LEN2 = 14, /* b = pop, a = pop, push sqrt(a*a + b*b + 0*0)       = Vec3(a,b,0).length()       */
    OM_SDF_LEN3 = 15, /* c, b, a = pop x3, push sqrt(a*a + b*b + c*c)       = Vec3(a,b,c).length()       */
    OM_SDF_STORE = 16, /* reg[arg] = pop                                                                 */
    OM_SDF_LOAD = 17   /* push reg[arg]                                                                  */
};
#define OM_SDF_STACK 8
#define OM_SDF_REGS 4
#define OM_SDF_MAX_WORDS 256
#define OM_SDF_INSN(op, arg) ((uint32_t)(op) | ((uint32_t)(arg) << 8))
typedef struct om_dyn_marched {
    float m_world_to_local[16]; /* to_local(p)  = this . p           (Mat4x4::dot, mat4x4.rs:45-48)       */
    float m_local_to_world[16]; /* to_world(n)  = this . n           (applied to the local normal, w = 0) */
    float world_f;              /* to_world_f(f) = f * world_f                                           */
    /* Bounded::build_world_bounding_box (bounding_box.rs:69-73): a world box that contains every point
     * with |sdf| < 0.005, or -INF/+INF = the trait's default "draw always".  Rays that miss the union of
     * the marched objects' boxes skip the march (they could neither start nor end on a marched surface).  */
    float bounds_lo[3], bounds_hi[3];
    om_material material;
} om_dyn_marched;

/* `Arc<dyn Traced + Send + Sync>` as data (traced_objects, hits.rs:39,91-95; trait traced.rs:33-35).  Compiled host code cannot
 * run on the device, so a user type that wants to be traced there describes itself as a QUADRIC SURFACE in its own frame,
 *     q(p) = A x^2 + B y^2 + C z^2 + D xy + E xz + F yz + G x + H y + I z + J = 0,   q[10] = {A .. J},
 * optionally cut by an axis-aligned box in that frame (clip_lo / clip_hi, +-INF = uncut): ellipsoids, cylinders, cones,
 * paraboloids, hyperboloids, and slabs or caps of them.  hit(r, t_min, t_max) is defined like the reference's own Sphere::hit
 * (traced.rs:38-63), which it reproduces bit for bit for q = {1,1,1,0,0,0,0,0,0,-1}: the ray goes to the local frame without
 * being renormalised (ray.rs:17-19), a = q2(d), half_b = o.Q.d + (G,H,I).d / 2, c = q(o), the roots (-half_b -+ sqrt(half_b^2 -
 * a c)) / a are tried nearer first and the first one inside [t_min, t_max] AND inside the clip box is the hit (a == 0: the one
 * root -c / (2 half_b)); point = M . p_local; normal = unit(M . grad q(p_local) / 2) - the local gradient carried by the
 * local->world matrix as Sphere and Cube do (traced.rs:59,296), not its inverse transpose.  Objects are scanned after the
 * parallelograms, as hits.rs:280-285 scans traced_objects after the typed lists (later object wins on equal t).
 * bounds_lo / bounds_hi: the world box `Bounded::build_world_bounding_box` returns (bounding_box.rs:69-73); +-INF = the trait's
 * default "draw always" (the object is then tested against every ray).  Object id = OM_OBJ_ID(OM_TYPE_DYN_TRACED, index).     */
typedef struct om_dyn_traced {
    float m_local_to_world[16];
    float m_world_to_local[16];
    float q[10];
    float clip_lo[3], clip_hi[3];
    float bounds_lo[3], bounds_hi[3];
    om_material material;
} om_dyn_traced;

/* camera.rs:9-29 Camera (all fields, same order)                                             */
typedef struct om_camera {
    float origin[3];
    float horizontal[3];
    float vertical[3];
    float lower_left_corner[3];
    float u_of_plane[3];
    float v_of_plane[3];
    float w_of_plane[3];
    float lens_radius;
    float aspect_ratio;
    float focus_dist;
    float viewport_width;
    float viewport_height;
} om_camera;

/* render_thread.rs:9-17 Stats, as read by the viewer (main.rs:228-232,249,287,377).
 * `c` of Pixel (render_thread.rs:44) is never written by the reference and is not mirrored.  */
typedef struct om_pixel {
    float sum[3];
    uint32_t n;
    float avg_depth;
    uint32_t bad_avgs;
    uint64_t bloom; /* BloomFilter.state, utils.rs:76-78 */
    uint8_t color[3];
    uint8_t _pad[5];
} om_pixel; /* 40 bytes */

/* Object identity.  The reference uses the object's address (utils.rs:110); across the ABI an
 * object is (type tag << 32) | (index in its list + 1); 0 stays "sky" (render_thread.rs:122). */
enum {
    OM_TYPE_SPHERE = 1, OM_TYPE_CUBE = 2, OM_TYPE_TRIANGLE = 3, OM_TYPE_INFINITE_PLANE = 4,
    OM_TYPE_PARALLELOGRAM = 5, OM_TYPE_MARCHED_SPHERE = 6, OM_TYPE_MARCHED_BOX = 7,
    OM_TYPE_MARCHED_TORUS = 8, OM_TYPE_DYN_MARCHED = 9, OM_TYPE_DYN_TRACED = 10
};
#define OM_OBJ_ID(type, index) ((((uint64_t)(type)) << 32) | ((uint64_t)(index) + 1u))

/* ------------------------------------------------------------------------------------------ */
/* render flags                                                                                */
enum {
    /* retire a pixel when its 8-bit colour was unchanged for 5 samples in a row — the ONLY
     * behaviour of the reference (render_thread.rs:31-38,194-198).  Off = fixed spp.          */
    OM_RENDER_ADAPTIVE = 1u << 0,
    /* Opt OUT of the parity arithmetic.  By default every result-bearing expression is evaluated
     * in the reference's order without FMA contraction and with IEEE div/sqrt, which makes the
     * output bit-identical to the CPU oracle sample by sample (FMAs are used only for culling).
     * With this flag the same kernels run compiled with FMA contraction everywhere: a few percent
     * faster, statistically equivalent, not bit-identical.                                     */
    OM_RENDER_FAST_MATH = 1u << 1,
    /* closest hit by linear scan in the reference's list order (hits.rs:270-286), no BVH.     */
    OM_RENDER_BRUTE_FORCE = 1u << 2,
    /* wavefront pipeline instead of the persistent path-regeneration kernel: per bounce an extend
     * kernel (BVH traversal, rays refilled from a queue), a shade kernel (sky / scatter by material
     * class, Stats::add, next sample or pixel; also advances the marched rays a bounded number of
     * steps) and a ballot+popc live-ray compaction; the last few thousand paths are finished by the
     * persistent kernel.  Same results bit for bit.  om_render_launch blocks until the tail starts. */
    OM_RENDER_WAVEFRONT = 1u << 3,
    /* keep primitives and BVH in global memory even if they fit in shared memory.             */
    OM_RENDER_NO_SMEM_SCENE = 1u << 4,
    /* instrumented kernels: count segments, BVH node visits, primitive tests, SDF evaluations
     * (the executed-work figures of the FP32 roofline accounting); slower.                    */
    OM_RENDER_COUNTERS = 1u << 5,
    /* force the persistent path-regeneration kernel.  With neither this nor OM_RENDER_WAVEFRONT the
     * library picks: wavefront when the scene image does not fit in shared memory (the BVH is then
     * walked through L2, which wants many more rays in flight) or when it has more than 8 marched
     * objects and no distance BVH over them (OM_FREEZE_NO_MARCHED_BVH; full-scan SDF marches stall
     * the persistent kernel's warps); persistent kernel otherwise.                                */
    OM_RENDER_PERSISTENT = 1u << 6,
    /* persistent kernel only: the 8 warps of each SM sub-partition meet at a named barrier at the top of
     * the path loop and between the closest-hit query and shading, so that the warps sharing an
     * instruction cache run the same part of the kernel together (fewer instruction-cache misses, at the
     * price of waiting for the group's slowest warp).  Same results bit for bit.  With neither flag the
     * library picks: on for scenes that march 1-8 objects without a distance BVH (measured on B200:
     * C1 +29 %, C2 +6 %, C4 +5 %), off otherwise (spheres only -13 %, 65 marched objects -33 %).         */
    OM_RENDER_PHASE_ALIGN = 1u << 7,
    OM_RENDER_NO_PHASE_ALIGN = 1u << 8,
    /* the CTA-resident wavefront ("path pool", csrc/om_pool.cuh): every CTA keeps a pool of path slots in shared memory
     * beside the scene image and alternates an extend phase (BVH traversal of compacted rays) with a shade phase that
     * takes the resolved rays one KIND at a time (marching, hit on a Lambertian / metal / dielectric sphere, other
     * primitive, miss), so scatter, Stats::add and the next camera ray run with full warps.  The iterative-queue form
     * of ray_color's recursion (render_thread.rs:128-143) without leaving the SM.  Needs the scene image, the traversal
     * stacks and the pool to fit shared memory (else OM_ERR_INVALID) and max_depth < 65536.  Bit-identical output.   */
    OM_RENDER_POOL = 1u << 9
};

typedef struct om_render_params {
    uint32_t width;       /* image_width   (render_thread.rs:146)                              */
    uint32_t height;      /* image_height                                                      */
    uint32_t spp_max;     /* samples_per_pixel: a MAXIMUM in adaptive mode                      */
    uint32_t max_depth;   /* max_depth     (main.rs:146)                                       */
    float tmin;           /* main.rs:205                                                       */
    float tmax;           /* main.rs:206                                                       */
    uint64_t seed;        /* Philox key; replaces the OS-seeded thread_rng (utils.rs:25)        */
    uint32_t flags;       /* OM_RENDER_*                                                       */
    /* spp-slice partition: this handle renders samples [sample_base, sample_base+sample_count)
     * of every pixel it owns.  sample_count 0 means spp_max - sample_base.                    */
    uint32_t sample_base;
    uint32_t sample_count;
    /* interleaved-chunk partition, the analogue of `tid`/`assigned_thread` (main.rs:172-189):
     * pixel p belongs to this handle iff (p / OM_CHUNK_PIXELS) % part_world == part_rank.      */
    uint32_t part_rank;
    uint32_t part_world;  /* 0 or 1 = own every pixel                                          */
    int32_t device;       /* CUDA device ordinal, -1 = current device                          */
} om_render_params;

#define OM_CHUNK_PIXELS 2730u /* 32 KiB / sizeof(Color), main.rs:174-175 */

typedef struct om_scene om_scene;   /* HittableList, then FrozenHittableList after freeze       */
typedef struct om_render om_render; /* the N render threads + their shared Vec<Pixel>           */

/* ------------------------------------------------------------------------------------------ */
/* scene: HittableList (hits.rs:37-43) — one list per concrete type, `world += &obj` ≙ add_*   */
/* Lifetime: an om_render handle refers to its scene's device images for as long as it lives (the Rust side holds the
 * FrozenHittableList in an Arc, main.rs:197-204).  While any handle created from a scene exists, om_scene_freeze,
 * om_scene_clear and om_scene_destroy on that scene fail with OM_ERR_INVALID and change nothing: destroy the handles first. */
OM_API om_status om_scene_create(om_scene** out);                 /* HittableList::new hits.rs:72 */
OM_API om_status om_scene_clear(om_scene* s);                     /* HittableList::clear hits.rs:81 */
OM_API om_status om_scene_destroy(om_scene* s);                   /* NULL is fine                */
OM_API om_status om_scene_add_spheres(om_scene* s, const om_sphere* a, size_t n);          /* hits.rs:101-105 */
OM_API om_status om_scene_add_cubes(om_scene* s, const om_cube* a, size_t n);
OM_API om_status om_scene_add_triangles(om_scene* s, const om_barycentric* a, size_t n);
OM_API om_status om_scene_add_infinite_planes(om_scene* s, const om_infinite_plane* a, size_t n);
OM_API om_status om_scene_add_parallelograms(om_scene* s, const om_barycentric* a, size_t n);
OM_API om_status om_scene_add_marched_spheres(om_scene* s, const om_marched_sphere* a, size_t n);
OM_API om_status om_scene_add_marched_boxes(om_scene* s, const om_marched_box* a, size_t n);
OM_API om_status om_scene_add_marched_tori(om_scene* s, const om_marched_torus* a, size_t n);
/* `world += Arc<dyn Marched>` (hits.rs:96-100) for a user type described as data: one object and the
 * n_words-word program of its local_sdf.  OM_ERR_INVALID for a malformed program (unknown opcode, stack
 * under/overflow, register out of range or read before written, more than OM_SDF_MAX_WORDS words, result
 * stack not exactly one value).  Dyn marched objects are scanned after the three typed marched lists, as
 * in hits.rs:312-320; object id = OM_OBJ_ID(9, index).                                                  */
OM_API om_status om_scene_add_dyn_marched(om_scene* s, const om_dyn_marched* obj, const uint32_t* code, size_t n_words);
/* number of dyn marched objects, and a copy of object `index` with its program (code may be NULL to ask
 * for *n_words only; on entry *n_words is the capacity of code).                                         */
OM_API om_status om_scene_dyn_marched_count(const om_scene* s, uint64_t* count);
OM_API om_status om_scene_get_dyn_marched(const om_scene* s, uint64_t index, om_dyn_marched* obj, uint32_t* code, size_t* n_words);
/* `world += Arc<dyn Traced>` (hits.rs:91-95) for user types described as data (om_dyn_traced): copies n objects.  OM_ERR_INVALID
 * for a non-finite matrix or coefficient, an all-zero quadric, or clip_lo > clip_hi.                                            */
OM_API om_status om_scene_add_dyn_traced(om_scene* s, const om_dyn_traced* a, size_t n);
OM_API om_status om_scene_dyn_traced_count(const om_scene* s, uint64_t* count);
/* A NATIVE trait object (compiled host code behind `Arc<dyn Traced>` / `Arc<dyn Marched>` that does not describe itself as
 * om_dyn_traced / om_dyn_marched) cannot cross to a device.  Always returns OM_ERR_UNSUPPORTED; present so the binding is total. */
OM_API om_status om_scene_add_dyn_object(om_scene* s, const void* trait_object);
/* counts per list, order: spheres,cubes,triangles,infinite_planes,parallelograms,
 * marched_spheres,marched_boxes,marched_tori */
OM_API om_status om_scene_counts(const om_scene* s, uint64_t counts[8]);
/* copy list `which` (0 spheres .. 7 marched_tori, the order of om_scene_counts; 8 = the om_dyn_traced objects) back to the caller;
 * nbytes must be count * sizeof(element).  Lets a host hand a generated scene to another consumer. */
OM_API om_status om_scene_get_list(const om_scene* s, uint32_t which, void* out, size_t nbytes);
/* HittableList::freeze (hits.rs:87-89).  Builds the BVH that subsumes CameraHash
 * (camera_hash.rs, hits.rs:116-185) from the objects' world bounding boxes
 * (bounding_box.rs:69-73).  `cam` may be NULL (the BVH does not depend on the camera).
 * flags (never change a result, only what gets culled):
 *   OM_FREEZE_NO_ALWAYS_LIST  scene-sized objects go into the BVH too
 *   OM_FREEZE_RAY_PADDING / OM_FREEZE_STATIC_PADDING  force / forbid per-ray box padding: the
 *     margin that makes the boxes cover the reference's f32 round-off grows with the distance
 *     from the ray origin; scenes of tiny objects spread over a large extent traverse several
 *     times fewer nodes when it is applied per ray instead of baked in for the worst case.
 *     Default: chosen from the scene.                                                            */
#define OM_FREEZE_NO_ALWAYS_LIST (1u << 0)
#define OM_FREEZE_RAY_PADDING    (1u << 1)
#define OM_FREEZE_STATIC_PADDING (1u << 2)
/* no distance BVH over the marched objects (min_j |sdf_j| by the full scan outside the candidate grid) */
#define OM_FREEZE_NO_MARCHED_BVH (1u << 3)
/* build the BVH over the traced primitives on the current CUDA device (Morton order + Karras' parallel
 * hierarchy) instead of the host's binned-SAH builder: a 10^6-sphere freeze drops from seconds to tens of
 * milliseconds, the tree traverses somewhat slower.  Needs a device (OM_ERR_NO_DEVICE / OM_ERR_CUDA).     */
#define OM_FREEZE_GPU_BVH (1u << 4)
OM_API om_status om_scene_freeze(om_scene* s, const om_camera* cam, uint32_t flags);
/* statistics of the frozen scene: [0] bvh nodes, [1] bvh depth, [2] primitives in the BVH,
 * [3] primitives kept in the always-test list, [4] bytes of the device scene image,
 * [5] bit 0: the hot part of that image fits in shared memory, bit 1: per-ray box padding is on
 * (OM_FREEZE_RAY_PADDING), [6] cells of the marched-object candidate grid
 * (0 = no grid), [7] entries in its lists (rounded up to 4) */
OM_API om_status om_scene_stats(const om_scene* s, uint64_t stats[8]);

/* ------------------------------------------------------------------------------------------ */
/* render: render_thread::render (render_thread.rs:145-202)                                    */
OM_API om_status om_render_create(om_render** out, const om_scene* frozen, const om_camera* cam,
                                  const om_render_params* params);
OM_API om_status om_render_destroy(om_render* r); /* NULL is fine */
/* Run this handle's work on a caller-owned CUDA stream (a cudaStream_t, e.g. torch's current
 * stream) instead of the handle's own; NULL switches back.  The caller keeps ownership.        */
OM_API om_status om_render_set_stream(om_render* r, void* cuda_stream);
/* Trace up to `spp_this_pass` more samples of every live pixel (sample-major passes, the
 * reference's outer loop render_thread.rs:176).  Asynchronous.  0 = all remaining samples.
 * Tile scheduling: the first fixed-spp launch of 32 samples or more on the persistent kernel is split
 * into a pilot launch of 1/64 of the samples, which measures what every 8x4 tile costs, and the
 * rest, whose 16x16-pixel blocks are then handed out most expensive first (a warp owns a tile for all
 * of its samples, so the launch ends when the last expensive tile does).  That call blocks for the
 * duration of the pilot; the order is kept for later launches of the handle.  Same pixels either way. */
OM_API om_status om_render_launch(om_render* r, uint32_t spp_this_pass);
OM_API om_status om_render_sync(om_render* r);
/* the AtomicU64 progress counter (main.rs:148, render_thread.rs:196-198): +1 per sample, plus
 * all left-over samples of a pixel at the moment it retires.  Synchronises.                   */
OM_API om_status om_render_samples_done(om_render* r, uint64_t* credited);
/* samples actually traced (what Msamples/s counts).  Synchronises.                            */
OM_API om_status om_render_samples_traced(om_render* r, uint64_t* traced);
/* all 16 device counters: [0] credited [1] traced [2] (internal) [3] segments [4] BVH node visits
 * [5] primitive tests [6] SDF evaluations [7] march steps [8] normal evaluations
 * [9] segments that entered the march loop.  [3..9] need OM_RENDER_COUNTERS.
 * Host side: [10] kernels launched by om_render_launch since create, [11] the flags in force
 * (with OM_RENDER_WAVEFRONT or OM_RENDER_PERSISTENT as chosen when the caller set neither),
 * [12] 1 if that pipeline keeps the scene image in shared memory.  Device again (OM_RENDER_COUNTERS,
 * persistent kernel): [13] path-loop iterations summed over warps, [14] the same summed over the lanes
 * that carried a path ([14] / (32 * [13]) = lane occupancy of the loop).  Host: [15] CUDA graph launches among [10]'s
 * kernels (the wavefront pipeline replays 48 iterations = 144 kernels per graph launch).  Synchronises.  */
OM_API om_status om_render_perf_counters(om_render* r, uint64_t out[16]);
/* device time of the launches since the last call to this function, in ms (CUDA events on the
 * render stream).  Synchronises.                                                              */
OM_API om_status om_render_elapsed_ms(om_render* r, float* ms, uint32_t* launches);
/* Vec<Pixel> read-out (what draw_to_sdl reads racily, main.rs:371).  `n` must be width*height. */
OM_API om_status om_render_read_pixels(om_render* r, om_pixel* host, size_t n);
/* MODE_NORMAL view (main.rs:374-381): tightly packed RGB24, width*height*3 bytes.             */
OM_API om_status om_render_read_rgb(om_render* r, uint8_t* host_rgb, size_t nbytes);
/* One refresh of the viewer's pixel buffer (draw_to_sdl, main.rs:361-437) computed on the device: tightly
 * packed RGB24.  mode 0 MODE_NORMAL, 1 MODE_SHOW_SAMPLES, 2 MODE_SAMPLE_WEIGHTED_BLUR, 3 MODE_SHOW_DEPTH,
 * 4 MODE_DEPTH_WEIGHTED_BLUR, 5 MODE_SHOW_IDS, 6 MODE_ID_WEIGHTED_BLUR (the 3x3 cross-weighted filters of
 * main.rs:219-343; they need width, height >= 3).  The SDL window itself is not part of this library. */
OM_API om_status om_render_read_view(om_render* r, uint32_t mode, uint8_t* host_rgb, size_t nbytes);
/* om_render_read_view without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments, same bytes out.  The device-side RGB image and the reduction words of modes 1 / 3
 * live in one file-scope device array of OM_VIEW_NOALLOC_MAX_PIXELS * 3 + 256 bytes (6 MiB + 256 B) that every GPU
 * which loads the library holds until the process exits; the call performs no cudaMalloc / cudaFree and calls no
 * allocating library routine.  Images of more than OM_VIEW_NOALLOC_MAX_PIXELS pixels are rejected with OM_ERR_INVALID
 * (use om_render_read_view).  All calls on one device share the array, and it keeps the previous call's contents (the
 * call re-initialises what it reads): calls that overlap on the GPU - two handles on different streams of the same
 * device - race; serialise them.  Synchronises the handle's stream.                                                    */
#define OM_VIEW_NOALLOC_MAX_PIXELS 2097152u /* 2 Mi pixels: covers 1920 x 1080 */
OM_API om_status om_render_read_view_noalloc(om_render* r, uint32_t mode, uint8_t* host_rgb, size_t nbytes);
/* F12 of the viewer (surface.save_bmp, main.rs:473-476): 24-bit BMP of an RGB24 buffer.             */
OM_API om_status om_write_bmp(const char* path, const uint8_t* rgb, uint32_t width, uint32_t height);
/* reset every pixel to Pixel::new() (render_thread.rs:47-50) and the counters to 0.           */
OM_API om_status om_render_reset(om_render* r);

/* Parity hook: primary-ray closest hit for every pixel — (obj_id, t), t = +INF and id 0 on a
 * miss — i.e. what ray_color returns as (depthf, obj_id) (render_thread.rs:132-135).
 * mode 0: pixel-centre ray (i_rand = j_rand = 0.5) without lens offset;
 * mode 1: the ray sample `sample_index` of the pixel would use (Philox jitter + thin lens).   */
OM_API om_status om_render_primary(om_render* r, uint32_t mode, uint32_t sample_index,
                                   uint64_t* ids, float* t, size_t n);

/* Multi-GPU plumbing.  Export the accumulators into / import from caller-owned DEVICE memory
 * (e.g. a torch tensor) so that any collective library can combine them.
 *   acc   : width*height*8 floats  {sum.r,sum.g,sum.b,n, n*avg_depth, bad_avgs, 0, 0}
 *   bloom : width*height uint64
 * om_render_import_accum replaces the handle's pixels by the combined values and re-derives
 * colour (utils.rs:12-17).                                                                   */
OM_API om_status om_render_export_accum(om_render* r, float* dev_acc, uint64_t* dev_bloom);
OM_API om_status om_render_import_accum(om_render* r, const float* dev_acc, const uint64_t* dev_bloom);
/* Adaptive sampling across GPUs (render_thread.rs:68-102,194-200 at multi-GPU scale): pixels retire at
 * different times, so the live ones are re-dealt between passes.  Every handle is created over the WHOLE
 * frame (no static partition); before each pass the caller gives all of them the same chunk -> owner
 * table (one byte per OM_CHUNK_PIXELS pixels) and its own id, a pass then renders the live pixels of
 * the chunks it owns; afterwards om_render_export_state(only_owned = 1) writes 8 words per pixel
 * {sum.rgb, n, avg_depth, bad_avgs, bloom lo, hi} bit for bit with zeros for other owners' pixels, an
 * integer SUM over the ranks (NCCL all-reduce) reassembles the frame, om_render_import_state loads it
 * into every handle, and om_render_live_per_chunk says how many pixels of each chunk still want samples
 * (the input of the next deal).  Results are those of one handle rendering alone, bit for bit.
 * With a table that does not change, a handle needs only its own chunks' state between passes: the exchange
 * can then wait until the frame is wanted (after the last pass, or whenever a viewer refreshes).
 * Adaptive passes over few live pixels (at most 250 000 on the handle) are rendered by a second kernel in
 * which 2..32 lanes share a pixel and fold their samples into its Stats in sample order - the same bits,
 * without a pass lasting as long as the hardest pixel's chain of samples.
 * owners == NULL returns the handle to the static rule of om_render_params.                          */
OM_API om_status om_render_set_chunk_owners(om_render* r, const uint8_t* owners, size_t n_chunks, uint32_t my_id);
OM_API om_status om_render_export_state(om_render* r, uint32_t* dev_state, uint32_t only_owned);
OM_API om_status om_render_import_state(om_render* r, const uint32_t* dev_state);
OM_API om_status om_render_live_per_chunk(om_render* r, uint32_t* host_counts, size_t n_chunks);
/* dst[i] |= src[k][i] for k < nsrc: the bitwise-OR NCCL does not have. Device pointers.       */
OM_API om_status om_bloom_or(uint64_t* dev_dst, const uint64_t* dev_src, size_t nsrc, size_t n,
                             int32_t device);
/* The same OR over width*height masks on the HANDLE'S stream, without synchronising: ordered like every other call on the
 * handle (om_bloom_or runs on the legacy default stream and synchronises the device, which does not order it against a
 * non-blocking stream that a collective was issued on).                                                                */
OM_API om_status om_render_bloom_or(om_render* r, uint64_t* dev_dst, const uint64_t* dev_src, size_t nsrc);

/* Parity hook: closest hit of arbitrary rays against a frozen scene, on the device, with the
 * semantics of FrozenHittableList::hit (hits.rs:270-334).  dirs are normalised first like
 * Ray::new (ray.rs:11-13).  Host arrays: origins/dirs n*3 floats in; ids n, t n, points n*3,
 * normals n*3 out (any out pointer may be NULL).  flags: OM_RENDER_FAST_MATH / _BRUTE_FORCE /
 * _NO_SMEM_SCENE.                                                                             */
OM_API om_status om_trace_rays(const om_scene* frozen, const float* origins, const float* dirs,
                               size_t n, float tmin, float tmax, uint32_t flags, int32_t device,
                               uint64_t* ids, float* t, float* points, float* normals);

/* ------------------------------------------------------------------------------------------ */
/* Host-side constructors: the reference keeps these in Rust; they are exported for hosts that
 * are not the Rust crate (the C++ mirror in ottomarcher.hpp, the Python tests and bench).     */
OM_API void om_camera_new(om_camera* out, const float lookfrom[3], const float lookat[3],
                          const float vup[3], float vfov_in_degrees, float aspect_ratio,
                          float aperture, float focus_dist);              /* camera.rs:38-59   */
OM_API void om_material_lambertian(om_material* out, const float albedo[3]);   /* materials.rs:26 */
OM_API void om_material_metal(om_material* out, const float albedo[3], float fuzz); /* :29-34  */
OM_API void om_material_dielectric(om_material* out, float ior);               /* materials.rs:35 */
OM_API void om_sphere_new(om_sphere* out, const float m_local_to_world[16], const om_material* m); /* traced.rs:22 */
OM_API void om_sphere_new_with_radius(om_sphere* out, const float o[3], float r, const om_material* m); /* traced.rs:26 */
OM_API void om_cube_new(om_cube* out, const float m_local_to_world[16], const om_material* m);   /* traced.rs:238 */
OM_API void om_cube_new_with_length(om_cube* out, const float o[3], float len, const om_material* m); /* traced.rs:242 */
OM_API void om_infinite_plane_new(om_infinite_plane* out, const float center[3], const float normal[3],
                                  const om_material* m);                       /* traced.rs:86  */
OM_API void om_barycentric_new3points(om_barycentric* out, const float origin[3], const float upoint[3],
                                      const float vpoint[3], const om_material* m); /* traced.rs:148 */
OM_API void om_marched_torus_new(om_marched_torus* out, const float m_local_to_world[16],
                                 const float local_sizes[3], const om_material* m); /* marched.rs:115 */
/* m4x4!(..) helpers (mat4x4.rs:88-123,167-209).  out = a ^ b is om_mat4_mul.                  */
OM_API void om_mat4_identity(float out[16]);
OM_API void om_mat4_translate(float out[16], float x, float y, float z);
OM_API void om_mat4_scale(float out[16], float x, float y, float z);
OM_API void om_mat4_rotate_x(float out[16], float f);
OM_API void om_mat4_rotate_y(float out[16], float f);
OM_API void om_mat4_rotate_z(float out[16], float f);
OM_API void om_mat4_mul(float out[16], const float a[16], const float b[16]);
OM_API void om_mat4_fast_homogenous_inverse(float out[16], const float m[16]);

/* BloomFilter::get_hash (utils.rs:93-107) and the shuffled 2x2 jitter table of
 * render_thread.rs:164-174 (code bit1 = jitter i, bit0 = jitter j), as the library computes them. */
OM_API uint64_t om_bloom_get_hash(uint64_t id);
OM_API void om_jitter_table(uint32_t spp, uint64_t seed, uint8_t* out);

/* Seeded scene generators (synthetic inputs for tests and benchmarks).  They
 * append to `s`.  Kinds:
 *   0 random_scene of main.rs:37-100 (Philox in place of thread_rng)
 *   1 basic_scene  of main.rs:102-110
 *   2 rtiow_classic: random_scene without torus / parallelogram / triangle / cube
 *   3 marched scene (marched ground sphere + 8x8 grid of marched sphere/box/torus)
 *   4 random_scene with >= 60 % dielectric small spheres (glass-heavy)
 *   5 n_param spheres in a 200 x 4 x 200 slab over the ground sphere (BVH stress)             */
OM_API om_status om_scene_generate(om_scene* s, uint32_t kind, uint64_t seed, uint64_t n_param);

/* ------------------------------------------------------------------------------------------ */
OM_API const char* om_last_error(void); /* thread-local, never NULL */
OM_API const char* om_version(void);
OM_API int32_t om_device_count(void);
/* The library keeps the large device blocks of destroyed render handles (framebuffer, read-out staging, wavefront
 * pools; at most 4 GiB per process) for the next om_render_create of the same size: cudaMalloc/cudaFree of a frame
 * cost up to a fifth of a render.  This releases them to the driver.                                          */
OM_API om_status om_trim_device_cache(void);
/* FP32 FMA microbenchmark on `device`: achieved TFLOP/s (2 flop per FMA lane) and SM clock.    */
OM_API om_status om_bench_fp32_peak(int32_t device, float* tflops, float* ms);

#ifdef __cplusplus
}
#endif
#endif /* OTTOMARCHER_H */
