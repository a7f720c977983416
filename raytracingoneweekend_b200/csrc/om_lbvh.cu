// BVH build on the device: Morton codes of the box centres, radix sort, Karras' parallel
// hierarchy, bottom-up box fit, nodes emitted in the renderer's 64-byte layout.  Used by om_scene_freeze
// with OM_FREEZE_GPU_BVH; the host binned-SAH builder (om_bvh.cpp) stays the default because its trees
// traverse faster — the BVH only culls, so either tree gives bit-identical images.
#include <cuda_runtime.h>
#include <stdio.h>

#include <cub/device/device_radix_sort.cuh>

#include "om_bvh.h"

namespace omb {
namespace {

constexpr int T = 256;
inline int blocks_for(size_t n) { return int((n + T - 1) / T); }

struct Box6 { float lo[3], hi[3]; };
// boxes written by other CTAs a moment ago: read them past L1
__device__ __forceinline__ Box6 load_cg(const Box6* p) {
    const float* f = reinterpret_cast<const float*>(p);
    Box6 b;
    b.lo[0] = __ldcg(f); b.lo[1] = __ldcg(f + 1); b.lo[2] = __ldcg(f + 2); b.hi[0] = __ldcg(f + 3); b.hi[1] = __ldcg(f + 4); b.hi[2] = __ldcg(f + 5);
    return b;
}

__device__ __forceinline__ unsigned long long spread21(unsigned long long v) {  // 21 bits -> every third bit
    v &= 0x1FFFFFull;
    v = (v | (v << 32)) & 0x1F00000000FFFFull;
    v = (v | (v << 16)) & 0x1F0000FF0000FFull;
    v = (v | (v << 8)) & 0x100F00F00F00F00Full;
    v = (v | (v << 4)) & 0x10C30C30C30C30C3ull;
    v = (v | (v << 2)) & 0x1249249249249249ull;
    return v;
}
__global__ void k_morton(const Box6* boxes, uint32_t n, float3 lo, float3 inv, unsigned long long* keys, uint32_t* vals) {
    const uint32_t i = blockIdx.x * T + threadIdx.x;
    if (i >= n) return;
    const Box6 b = boxes[i];
    const float s = 2097152.f;  // 2^21
    const float ux = (0.5f * (b.lo[0] + b.hi[0]) - lo.x) * inv.x, uy = (0.5f * (b.lo[1] + b.hi[1]) - lo.y) * inv.y, uz = (0.5f * (b.lo[2] + b.hi[2]) - lo.z) * inv.z;
    const unsigned long long qx = (unsigned long long)fminf(fmaxf(ux * s, 0.f), s - 1.f), qy = (unsigned long long)fminf(fmaxf(uy * s, 0.f), s - 1.f),
                             qz = (unsigned long long)fminf(fmaxf(uz * s, 0.f), s - 1.f);
    keys[i] = (spread21(qx) << 2) | (spread21(qy) << 1) | spread21(qz);
    vals[i] = i;
}

// length of the common prefix of the keys at sorted positions i and j; equal keys are told apart by position
__device__ __forceinline__ int delta(const unsigned long long* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const unsigned long long a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz((unsigned)i ^ (unsigned)j);
    return __clzll((long long)(a ^ b));
}
// Karras 2012: inner node i covers a range of sorted leaves found from the prefix lengths around position i.
// child encoding here: >= 0 inner node, < 0 leaf ~k (k = sorted position)
__global__ void k_hierarchy(const unsigned long long* keys, int n, int2* child, int* parent_inner, int* parent_leaf) {
    const int i = blockIdx.x * T + threadIdx.x;
    if (i >= n - 1) return;
    const int d = delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) / 2;; t = (t + 1) / 2) {
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t <= 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const int left = lo == gamma ? ~gamma : gamma, right = hi == gamma + 1 ? ~(gamma + 1) : gamma + 1;
    child[i] = make_int2(left, right);
    if (left < 0) parent_leaf[gamma] = i; else parent_inner[gamma] = i;
    if (right < 0) parent_leaf[gamma + 1] = i; else parent_inner[gamma + 1] = i;
    if (i == 0) parent_inner[0] = -1;
}
// boxes of the inner nodes, bottom-up: the second thread to reach a node owns it
__global__ void k_fit(const Box6* boxes, const uint32_t* sorted, int n, const int2* child, const int* parent_inner, const int* parent_leaf, Box6* nbox, int* flags,
                      int* max_depth) {
    const int k = blockIdx.x * T + threadIdx.x;
    if (k >= n) return;
    int p = parent_leaf[k];
    int depth = 1;
    {   // this leaf's distance from the root
        int q = p;
        while (q >= 0) { depth++; q = parent_inner[q]; }
        atomicMax(max_depth, depth);
    }
    while (p >= 0) {
        __threadfence();
        if (atomicAdd(&flags[p], 1) == 0) return;
        const int2 c = child[p];
        const Box6 a = c.x < 0 ? boxes[sorted[~c.x]] : load_cg(nbox + c.x), b = c.y < 0 ? boxes[sorted[~c.y]] : load_cg(nbox + c.y);
        Box6 u;
        for (int x = 0; x < 3; x++) { u.lo[x] = fminf(a.lo[x], b.lo[x]); u.hi[x] = fmaxf(a.hi[x], b.hi[x]); }
        nbox[p] = u;
        p = parent_inner[p];
    }
}
__global__ void k_emit(const Box6* boxes, const uint32_t* sorted, const uint32_t* refs, int n, const int2* child, const Box6* nbox, Node* nodes) {
    const int i = blockIdx.x * T + threadIdx.x;
    if (i >= n - 1) return;
    const int2 c = child[i];
    const Box6 a = c.x < 0 ? boxes[sorted[~c.x]] : nbox[c.x], b = c.y < 0 ? boxes[sorted[~c.y]] : nbox[c.y];
    Node nd;
    for (int x = 0; x < 3; x++) { nd.lmin[x] = a.lo[x]; nd.lmax[x] = a.hi[x]; nd.rmin[x] = b.lo[x]; nd.rmax[x] = b.hi[x]; }
    nd.left = c.x < 0 ? ~int32_t(refs[sorted[~c.x]]) : c.x;
    nd.right = c.y < 0 ? ~int32_t(refs[sorted[~c.y]]) : c.y;
    nd.pad0 = nd.pad1 = 0;
    nodes[i] = nd;
}

}  // namespace

bool build_gpu(const std::vector<uint32_t>& refs, const std::vector<Aabb>& boxes, BuildResult& out, uint32_t max_depth, char* err, size_t errlen) {
    static_assert(sizeof(Aabb) == sizeof(Box6), "box layout");
    out.nodes.clear();
    out.depth = 0;
    const size_t n = refs.size();
    auto fail = [&](const char* what, cudaError_t e) { if (err) snprintf(err, errlen, "%s%s%s", what, e != cudaSuccess ? ": " : "", e != cudaSuccess ? cudaGetErrorString(e) : ""); return false; };
    if (n < 2 || n > 0x3FFFFFFFu) return fail("GPU BVH build needs 2 .. 2^30 primitives", cudaSuccess);
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (size_t i = 0; i < n; i++)
        for (int a = 0; a < 3; a++) {
            const float c = 0.5f * (boxes[i].lo[a] + boxes[i].hi[a]);
            if (!(c - c == 0.f)) return fail("GPU BVH build: non-finite box", cudaSuccess);
            lo[a] = fminf(lo[a], c); hi[a] = fmaxf(hi[a], c);
        }
    char* buf = nullptr;
    void* tmp = nullptr;
    auto al = [](size_t b) { return (b + 255) & ~size_t(255); };
    const size_t b_box = al(n * sizeof(Box6)), b_key = al(n * 8), b_val = al(n * 4), b_child = al(n * 8), b_par = al(n * 4), b_nodes = al(n * sizeof(Node));
    const size_t total = 2 * b_box + 2 * b_key + 3 * b_val + b_child + 2 * b_par + b_par + 256 + b_nodes;
    cudaError_t e = cudaMalloc(&buf, total);
    if (e != cudaSuccess) return fail("GPU BVH build: cudaMalloc", e);
    char* p = buf;
    auto take = [&](size_t b) { char* r = p; p += b; return r; };
    Box6* d_box = (Box6*)take(b_box); Box6* d_nbox = (Box6*)take(b_box);
    unsigned long long* d_key = (unsigned long long*)take(b_key); unsigned long long* d_key2 = (unsigned long long*)take(b_key);
    uint32_t* d_val = (uint32_t*)take(b_val); uint32_t* d_val2 = (uint32_t*)take(b_val); uint32_t* d_refs = (uint32_t*)take(b_val);
    int2* d_child = (int2*)take(b_child);
    int* d_pin = (int*)take(b_par); int* d_pleaf = (int*)take(b_par); int* d_flags = (int*)take(b_par);
    int* d_depth = (int*)take(256);
    Node* d_nodes = (Node*)take(b_nodes);
    auto done = [&](bool ok) { cudaFree(tmp); cudaFree(buf); return ok; };
#define LB(x) do { e = (x); if (e != cudaSuccess) return done(fail("GPU BVH build: " #x, e)); } while (0)
    LB(cudaMemcpy(d_box, boxes.data(), n * sizeof(Box6), cudaMemcpyHostToDevice));
    LB(cudaMemcpy(d_refs, refs.data(), n * 4, cudaMemcpyHostToDevice));
    LB(cudaMemset(d_flags, 0, n * 4));
    LB(cudaMemset(d_depth, 0, 4));
    const float3 flo = make_float3(lo[0], lo[1], lo[2]);
    // one scale for the three axes: cubic Morton cells.  Per-axis scaling would cut a flat scene (C5's 200 x 4 x 200 slab)
    // into pancakes, which traverse twice as slowly.
    const float ext = fmaxf(hi[0] - lo[0], fmaxf(hi[1] - lo[1], hi[2] - lo[2]));
    const float iv = ext > 0.f ? 1.f / ext : 0.f;
    const float3 inv = make_float3(iv, iv, iv);
    k_morton<<<blocks_for(n), T>>>(d_box, (uint32_t)n, flo, inv, d_key, d_val);
    size_t tmp_bytes = 0;
    LB(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, d_key, d_key2, d_val, d_val2, (int)n, 0, 63));
    LB(cudaMalloc(&tmp, tmp_bytes));
    LB(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, d_key, d_key2, d_val, d_val2, (int)n, 0, 63));
    k_hierarchy<<<blocks_for(n - 1), T>>>(d_key2, (int)n, d_child, d_pin, d_pleaf);
    k_fit<<<blocks_for(n), T>>>(d_box, d_val2, (int)n, d_child, d_pin, d_pleaf, d_nbox, d_flags, d_depth);
    k_emit<<<blocks_for(n - 1), T>>>(d_box, d_val2, d_refs, (int)n, d_child, d_nbox, d_nodes);
    LB(cudaGetLastError());
    int depth = 0;
    LB(cudaMemcpy(&depth, d_depth, 4, cudaMemcpyDeviceToHost));
    if ((uint32_t)depth > max_depth) return done(fail("GPU BVH build: tree deeper than the traversal stack", cudaSuccess));
    out.nodes.resize(n - 1);
    LB(cudaMemcpy(out.nodes.data(), d_nodes, (n - 1) * sizeof(Node), cudaMemcpyDeviceToHost));
#undef LB
    out.depth = (uint32_t)depth;
    return done(true);
}

}  // namespace omb
