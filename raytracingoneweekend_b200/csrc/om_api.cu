// C ABI of include/ottomarcher.h: handles, device memory, streams, launches.
// There is no CPU fallback anywhere in this file: without a CUDA device every compute entry
// point returns OM_ERR_NO_DEVICE.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <map>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/ottomarcher.h"
#include "om_host.h"
#include "om_kernels.h"
#include "om_scene.h"

namespace {

thread_local std::string g_err;
om_status fail(om_status s, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return s;
}
#define CU(call)                                                                                      \
    do {                                                                                              \
        cudaError_t e_ = (call);                                                                      \
        if (e_ != cudaSuccess) {                                                                      \
            om_status st_ = (e_ == cudaErrorMemoryAllocation) ? OM_ERR_OOM                            \
                          : (e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver) ? OM_ERR_NO_DEVICE : OM_ERR_CUDA; \
            return fail(st_, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
        }                                                                                             \
    } while (0)

struct DevScene {
    float4* hot = nullptr;
    float4* cold = nullptr;
};

struct DeviceProps {
    int sms = 0;
    size_t smem_optin = 0;
};
om_status device_props(int dev, DeviceProps& p) {
    int v = 0;
    CU(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev));
    p.sms = v;
    CU(cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    p.smem_optin = size_t(v);
    return OM_OK;
}
// Device blocks of 1 MiB and more (framebuffers, read-out staging, wavefront pools) are kept after om_render_destroy and
// handed to the next om_render_create that asks for the same size on the same device: cudaMalloc / cudaFree of a 66 MB
// frame were measured at 10-280 ms per call on B200 hosts, a fifth of a 1024-spp C2 render.  At most DEV_CACHE_MAX bytes per
// process are kept; om_trim_device_cache() releases them.  A block is only reused after its owner synchronised its stream
// (om_render_destroy does).
constexpr size_t DEV_CACHE_MIN_BLOCK = size_t(1) << 20, DEV_CACHE_MAX = size_t(4) << 30;
struct DevBlockCache {
    std::mutex mu;
    std::multimap<std::pair<int, size_t>, void*> free_blocks;
    std::map<void*, std::pair<int, size_t>> live;
    size_t cached = 0;
} g_dev_cache;
cudaError_t dev_alloc_raw(void** p, size_t bytes) {
    *p = nullptr;
    if (bytes >= DEV_CACHE_MIN_BLOCK) {
        int dev = 0;
        cudaGetDevice(&dev);
        std::lock_guard<std::mutex> lk(g_dev_cache.mu);
        auto it = g_dev_cache.free_blocks.find({dev, bytes});
        if (it != g_dev_cache.free_blocks.end()) {
            *p = it->second;
            g_dev_cache.free_blocks.erase(it);
            g_dev_cache.cached -= bytes;
            g_dev_cache.live[*p] = {dev, bytes};
            return cudaSuccess;
        }
    }
    cudaError_t e = cudaMalloc(p, bytes);
    if (e != cudaSuccess && bytes >= DEV_CACHE_MIN_BLOCK) {  // out of memory with blocks parked in the cache: release them and retry
        cudaGetLastError();
        {
            std::lock_guard<std::mutex> lk(g_dev_cache.mu);
            int prev = 0;
            cudaGetDevice(&prev);
            for (auto& kv : g_dev_cache.free_blocks) { cudaSetDevice(kv.first.first); cudaFree(kv.second); }
            cudaSetDevice(prev);
            g_dev_cache.free_blocks.clear();
            g_dev_cache.cached = 0;
        }
        e = cudaMalloc(p, bytes);
    }
    if (e == cudaSuccess && bytes >= DEV_CACHE_MIN_BLOCK) {
        int dev = 0;
        cudaGetDevice(&dev);
        std::lock_guard<std::mutex> lk(g_dev_cache.mu);
        g_dev_cache.live[*p] = {dev, bytes};
    }
    return e;
}
template <class T>
cudaError_t dev_alloc(T** p, size_t bytes) { return dev_alloc_raw(reinterpret_cast<void**>(p), bytes); }
void dev_free(void* p) {
    if (!p) return;
    {
        std::lock_guard<std::mutex> lk(g_dev_cache.mu);
        auto it = g_dev_cache.live.find(p);
        if (it != g_dev_cache.live.end()) {
            const std::pair<int, size_t> key = it->second;
            g_dev_cache.live.erase(it);
            if (g_dev_cache.cached + key.second <= DEV_CACHE_MAX) {
                g_dev_cache.free_blocks.insert({key, p});
                g_dev_cache.cached += key.second;
                return;
            }
        }
    }
    cudaFree(p);
}

om_status pick_device(int32_t want, int& dev) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail(OM_ERR_NO_DEVICE, "no CUDA device (%s); this library has no CPU path", e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    }
    if (want < 0) { CU(cudaGetDevice(&dev)); return OM_OK; }
    if (want >= n) return fail(OM_ERR_INVALID, "device %d out of range (have %d)", want, n);
    dev = want;
    return OM_OK;
}

}  // namespace

struct om_scene {
    omh::SceneLists lists;
    bool frozen = false;
    oms::FrozenScene fz;
    std::mutex mu;
    std::map<int, DevScene> dev;  // uploaded images per device
    // live om_render handles: each keeps raw pointers into `dev` and re-reads fz.hdr at every launch, so the scene may not be
    // re-frozen, cleared or destroyed under them (om_scene_freeze / _clear / _destroy return OM_ERR_INVALID while this is > 0)
    std::atomic<int> handles{0};
    ~om_scene() { release_device(); }
    void release_device() {
        for (auto& kv : dev) {
            int prev = 0;
            cudaGetDevice(&prev);
            cudaSetDevice(kv.first);
            cudaFree(kv.second.hot);
            cudaFree(kv.second.cold);
            cudaSetDevice(prev);
        }
        dev.clear();
    }
};

namespace {
om_status scene_on_device(om_scene* s, int dev, DevScene& out) {
    std::lock_guard<std::mutex> lk(s->mu);
    auto it = s->dev.find(dev);
    if (it != s->dev.end()) { out = it->second; return OM_OK; }
    DevScene d;
    CU(cudaSetDevice(dev));
    CU(cudaMalloc(&d.hot, s->fz.hot.size() * sizeof(float)));
    CU(cudaMalloc(&d.cold, s->fz.cold.size() * sizeof(float)));
    CU(cudaMemcpy(d.hot, s->fz.hot.data(), s->fz.hot.size() * sizeof(float), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d.cold, s->fz.cold.data(), s->fz.cold.size() * sizeof(float), cudaMemcpyHostToDevice));
    // a pageable-memory cudaMemcpy may return once the data sits in the driver's staging buffer; the render streams are
    // non-blocking (no implicit order with the legacy stream), so wait for the DMA itself before anyone launches on the image
    CU(cudaDeviceSynchronize());
    s->dev[dev] = d;
    out = d;
    return OM_OK;
}
}  // namespace

extern "C" {

const char* om_last_error(void) { return g_err.c_str(); }
const char* om_version(void) { return "ottomarcher-b200 0.1 (sm_100a)"; }
int32_t om_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

// ---- scene ---------------------------------------------------------------------------------------
om_status om_trim_device_cache(void) {
    std::lock_guard<std::mutex> lk(g_dev_cache.mu);
    int prev = 0;
    cudaGetDevice(&prev);
    for (auto& kv : g_dev_cache.free_blocks) { cudaSetDevice(kv.first.first); cudaFree(kv.second); }
    cudaSetDevice(prev);
    g_dev_cache.free_blocks.clear();
    g_dev_cache.cached = 0;
    return OM_OK;
}
om_status om_scene_create(om_scene** out) {
    if (!out) return fail(OM_ERR_INVALID, "om_scene_create: out is NULL");
    *out = new (std::nothrow) om_scene();
    return *out ? OM_OK : fail(OM_ERR_OOM, "om_scene_create: out of host memory");
}
om_status om_scene_destroy(om_scene* s) {
    if (s && s->handles.load() > 0) return fail(OM_ERR_INVALID, "om_scene_destroy: %d render handle(s) still use this scene (om_render_destroy them first)", s->handles.load());
    delete s;
    return OM_OK;
}
om_status om_scene_clear(om_scene* s) {
    if (!s) return fail(OM_ERR_INVALID, "om_scene_clear: NULL scene");
    if (s->handles.load() > 0) return fail(OM_ERR_INVALID, "om_scene_clear: %d render handle(s) still use this scene (om_render_destroy them first)", s->handles.load());
    s->lists = omh::SceneLists();
    s->frozen = false;
    s->release_device();
    return OM_OK;
}
#define OM_ADD(NAME, TYPE, FIELD)                                                                   \
    om_status NAME(om_scene* s, const TYPE* a, size_t n) {                                          \
        if (!s || (!a && n)) return fail(OM_ERR_INVALID, #NAME ": NULL argument");                  \
        if (s->frozen) return fail(OM_ERR_INVALID, #NAME ": scene is frozen (om_scene_clear first)"); \
        if (s->lists.FIELD.size() + n > uint64_t(omd::REF_MAX_OBJECTS)) return fail(OM_ERR_INVALID, #NAME ": too many objects"); \
        s->lists.FIELD.insert(s->lists.FIELD.end(), a, a + n);                                      \
        return OM_OK;                                                                               \
    }
OM_ADD(om_scene_add_spheres, om_sphere, spheres)
OM_ADD(om_scene_add_cubes, om_cube, cubes)
OM_ADD(om_scene_add_triangles, om_barycentric, triangles)
OM_ADD(om_scene_add_infinite_planes, om_infinite_plane, infinite_planes)
OM_ADD(om_scene_add_parallelograms, om_barycentric, parallelograms)
OM_ADD(om_scene_add_marched_spheres, om_marched_sphere, marched_spheres)
OM_ADD(om_scene_add_marched_boxes, om_marched_box, marched_boxes)
OM_ADD(om_scene_add_marched_tori, om_marched_torus, marched_tori)
om_status om_scene_add_dyn_marched(om_scene* s, const om_dyn_marched* obj, const uint32_t* code, size_t n_words) {
    if (!s || !obj || !code) return fail(OM_ERR_INVALID, "om_scene_add_dyn_marched: NULL argument");
    if (s->frozen) return fail(OM_ERR_INVALID, "om_scene_add_dyn_marched: scene is frozen (om_scene_clear first)");
    const char* why = omh::check_sdf_program(code, n_words);
    if (why[0]) return fail(OM_ERR_INVALID, "om_scene_add_dyn_marched: %s", why);
    omh::SceneLists& L = s->lists;
    if (L.dyn_marched.size() + 1 > uint64_t(omd::REF_MAX_OBJECTS) || L.dyn_code.size() + n_words > 0x3FFFFFFFull) return fail(OM_ERR_INVALID, "om_scene_add_dyn_marched: too many objects");
    if (L.dyn_code_first.empty()) L.dyn_code_first.push_back(0u);
    L.dyn_marched.push_back(*obj);
    L.dyn_code.insert(L.dyn_code.end(), code, code + n_words);
    L.dyn_code_first.push_back(uint32_t(L.dyn_code.size()));
    return OM_OK;
}
om_status om_scene_add_dyn_traced(om_scene* s, const om_dyn_traced* a, size_t n) {
    if (!s || (!a && n)) return fail(OM_ERR_INVALID, "om_scene_add_dyn_traced: NULL argument");
    if (s->frozen) return fail(OM_ERR_INVALID, "om_scene_add_dyn_traced: scene is frozen (om_scene_clear first)");
    if (s->lists.dyn_traced.size() + n > uint64_t(omd::REF_MAX_OBJECTS)) return fail(OM_ERR_INVALID, "om_scene_add_dyn_traced: too many objects");
    for (size_t i = 0; i < n; i++) {
        const char* why = omh::check_dyn_traced(a[i]);
        if (why[0]) return fail(OM_ERR_INVALID, "om_scene_add_dyn_traced: object %zu: %s", i, why);
    }
    s->lists.dyn_traced.insert(s->lists.dyn_traced.end(), a, a + n);
    return OM_OK;
}
om_status om_scene_dyn_traced_count(const om_scene* s, uint64_t* count) {
    if (!s || !count) return fail(OM_ERR_INVALID, "om_scene_dyn_traced_count: NULL argument");
    *count = s->lists.dyn_traced.size();
    return OM_OK;
}
om_status om_scene_dyn_marched_count(const om_scene* s, uint64_t* count) {
    if (!s || !count) return fail(OM_ERR_INVALID, "om_scene_dyn_marched_count: NULL argument");
    *count = s->lists.dyn_marched.size();
    return OM_OK;
}
om_status om_scene_get_dyn_marched(const om_scene* s, uint64_t index, om_dyn_marched* obj, uint32_t* code, size_t* n_words) {
    if (!s || !n_words) return fail(OM_ERR_INVALID, "om_scene_get_dyn_marched: NULL argument");
    const omh::SceneLists& L = s->lists;
    if (index >= L.dyn_marched.size()) return fail(OM_ERR_INVALID, "om_scene_get_dyn_marched: object %llu does not exist", (unsigned long long)index);
    const size_t first = L.dyn_code_first[index], n = L.dyn_code_first[index + 1] - first;
    if (obj) *obj = L.dyn_marched[index];
    if (code) {
        if (*n_words < n) return fail(OM_ERR_INVALID, "om_scene_get_dyn_marched: room for %zu words, program has %zu", *n_words, n);
        memcpy(code, L.dyn_code.data() + first, n * sizeof(uint32_t));
    }
    *n_words = n;
    return OM_OK;
}
om_status om_scene_add_dyn_object(om_scene*, const void*) {
    return fail(OM_ERR_UNSUPPORTED, "native Arc<dyn Traced>/Arc<dyn Marched> objects (hits.rs:91-100) cannot cross to the device; describe a Marched type as data with om_scene_add_dyn_marched");
}
om_status om_scene_counts(const om_scene* s, uint64_t c[8]) {
    if (!s || !c) return fail(OM_ERR_INVALID, "om_scene_counts: NULL argument");
    c[0] = s->lists.spheres.size(); c[1] = s->lists.cubes.size(); c[2] = s->lists.triangles.size(); c[3] = s->lists.infinite_planes.size();
    c[4] = s->lists.parallelograms.size(); c[5] = s->lists.marched_spheres.size(); c[6] = s->lists.marched_boxes.size(); c[7] = s->lists.marched_tori.size();
    return OM_OK;
}
om_status om_scene_get_list(const om_scene* s, uint32_t which, void* out, size_t nbytes) {
    if (!s || (!out && nbytes)) return fail(OM_ERR_INVALID, "om_scene_get_list: NULL argument");
    const void* src = nullptr;
    size_t have = 0;
    const omh::SceneLists& L = s->lists;
    switch (which) {
        case 0: src = L.spheres.data(); have = L.spheres.size() * sizeof(om_sphere); break;
        case 1: src = L.cubes.data(); have = L.cubes.size() * sizeof(om_cube); break;
        case 2: src = L.triangles.data(); have = L.triangles.size() * sizeof(om_barycentric); break;
        case 3: src = L.infinite_planes.data(); have = L.infinite_planes.size() * sizeof(om_infinite_plane); break;
        case 4: src = L.parallelograms.data(); have = L.parallelograms.size() * sizeof(om_barycentric); break;
        case 5: src = L.marched_spheres.data(); have = L.marched_spheres.size() * sizeof(om_marched_sphere); break;
        case 6: src = L.marched_boxes.data(); have = L.marched_boxes.size() * sizeof(om_marched_box); break;
        case 7: src = L.marched_tori.data(); have = L.marched_tori.size() * sizeof(om_marched_torus); break;
        case 8: src = L.dyn_traced.data(); have = L.dyn_traced.size() * sizeof(om_dyn_traced); break;
        default: return fail(OM_ERR_INVALID, "om_scene_get_list: list %u does not exist", which);
    }
    if (nbytes != have) return fail(OM_ERR_INVALID, "om_scene_get_list: nbytes %zu != %zu", nbytes, have);
    if (have) memcpy(out, src, have);
    return OM_OK;
}
om_status om_scene_generate(om_scene* s, uint32_t kind, uint64_t seed, uint64_t n_param) {
    if (!s) return fail(OM_ERR_INVALID, "om_scene_generate: NULL scene");
    if (s->frozen) return fail(OM_ERR_INVALID, "om_scene_generate: scene is frozen");
    if (kind == 5 && n_param > uint64_t(omd::REF_MAX_OBJECTS) - 16u) return fail(OM_ERR_INVALID, "om_scene_generate: too many spheres");
    if (!omh::generate(s->lists, kind, seed, n_param)) return fail(OM_ERR_INVALID, "om_scene_generate: unknown kind %u", kind);
    return OM_OK;
}
om_status om_scene_freeze(om_scene* s, const om_camera* cam, uint32_t flags) {
    if (!s) return fail(OM_ERR_INVALID, "om_scene_freeze: NULL scene");
    if (s->handles.load() > 0) return fail(OM_ERR_INVALID, "om_scene_freeze: %d render handle(s) still use this scene (om_render_destroy them first)", s->handles.load());
    s->release_device();
    s->frozen = false;
    const std::string err = oms::freeze(s->lists, cam, flags, s->fz);
    if (!err.empty()) return fail(om_device_count() > 0 ? OM_ERR_CUDA : OM_ERR_NO_DEVICE, "om_scene_freeze: %s", err.c_str());
    s->frozen = true;
    return OM_OK;
}
om_status om_scene_stats(const om_scene* s, uint64_t st[8]) {
    if (!s || !st || !s->frozen) return fail(OM_ERR_INVALID, "om_scene_stats: scene not frozen");
    st[0] = s->fz.hdr.n_nodes; st[1] = s->fz.hdr.bvh_depth; st[2] = s->fz.hdr.n_bvh_prims; st[3] = s->fz.hdr.n_large;
    st[4] = uint64_t(s->fz.hdr.hot_units) * 16u + uint64_t(s->fz.hdr.cold_units) * 16u;
    // bit 0: the scene image and k_render's traversal stacks fit the 227 KB of shared memory of one CTA
    st[5] = ((uint64_t(s->fz.hdr.hot_units) * 16u + om_exact::render_stack_bytes(s->fz.hdr.stack_need) + 2048u <= 227u * 1024u && s->fz.hdr.n_nodes <= 16384u) ? 1 : 0) | (s->fz.hdr.dyn_pad != 0.f ? 2 : 0);
    st[6] = uint64_t(s->fz.hdr.mgrid_dim[0]) * s->fz.hdr.mgrid_dim[1] * s->fz.hdr.mgrid_dim[2];
    st[7] = st[6] ? uint64_t(s->fz.hdr.cold_units - s->fz.hdr.off_mgrid_lists) * 4u : 0u;
    return OM_OK;
}

// ---- host constructors -----------------------------------------------------------------------------
void om_camera_new(om_camera* out, const float lookfrom[3], const float lookat[3], const float vup[3], float vfov, float aspect, float aperture, float focus_dist) {
    omh::camera_new(out, lookfrom, lookat, vup, vfov, aspect, aperture, focus_dist);
}
void om_material_lambertian(om_material* out, const float a[3]) { *out = omh::lambertian(a[0], a[1], a[2]); }
void om_material_metal(om_material* out, const float a[3], float fuzz) { *out = omh::metal(a[0], a[1], a[2], fuzz); }
void om_material_dielectric(om_material* out, float ior) { *out = omh::dielectric(ior); }
void om_sphere_new(om_sphere* out, const float m[16], const om_material* mat) { omh::sphere_new(out, m, *mat); }
void om_sphere_new_with_radius(om_sphere* out, const float o[3], float r, const om_material* mat) { omh::sphere_new_with_radius(out, o, r, *mat); }
void om_cube_new(om_cube* out, const float m[16], const om_material* mat) { omh::cube_new(out, m, *mat); }
void om_cube_new_with_length(om_cube* out, const float o[3], float len, const om_material* mat) { omh::cube_new_with_length(out, o, len, *mat); }
void om_infinite_plane_new(om_infinite_plane* out, const float c[3], const float n[3], const om_material* mat) { omh::infinite_plane_new(out, c, n, *mat); }
void om_barycentric_new3points(om_barycentric* out, const float o[3], const float u[3], const float v[3], const om_material* mat) { omh::barycentric_new3points(out, o, u, v, *mat); }
void om_marched_torus_new(om_marched_torus* out, const float m[16], const float sizes[3], const om_material* mat) { omh::marched_torus_new(out, m, sizes, *mat); }
void om_mat4_identity(float o[16]) { omh::mat4_identity(o); }
void om_mat4_translate(float o[16], float x, float y, float z) { omh::mat4_translate(o, x, y, z); }
void om_mat4_scale(float o[16], float x, float y, float z) { omh::mat4_scale(o, x, y, z); }
void om_mat4_rotate_x(float o[16], float f) { omh::mat4_rotate_x(o, f); }
void om_mat4_rotate_y(float o[16], float f) { omh::mat4_rotate_y(o, f); }
void om_mat4_rotate_z(float o[16], float f) { omh::mat4_rotate_z(o, f); }
void om_mat4_mul(float o[16], const float a[16], const float b[16]) { omh::mat4_mul(o, a, b); }
void om_mat4_fast_homogenous_inverse(float o[16], const float m[16]) { omh::mat4_fast_homogenous_inverse(o, m); }

}  // extern "C"

#include "om_api_render.inl"
