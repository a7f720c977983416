#include <type_traits>
// Host-side launch wrappers, compiled once per build of om_device.cuh (namespace OM_NS).
namespace OM_NS {

template <class K>
static cudaError_t prep(K kern, size_t smem_bytes) {
    return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
}

// k_render<SMEM, COUNT, MA>: MA = 1 when the scene carries a candidate grid / distance BVH over its marched objects,
// MA = 3 when it has dyn marched objects (om_dyn_marched), else 0
template <class F>
static void with_render_kernel(bool smem, bool count, int ma, F f) {
    auto pick = [&](auto SM, auto CN) {
        constexpr bool sm = decltype(SM)::value, cn = decltype(CN)::value;
        if (ma == 1) f(k_render<sm, cn, 1>);
        else if (ma == 3) f(k_render<sm, cn, 3>);
        else f(k_render<sm, cn, 0>);
    };
    if (smem) { if (count) pick(std::true_type{}, std::true_type{}); else pick(std::true_type{}, std::false_type{}); }
    else { if (count) pick(std::false_type{}, std::true_type{}); else pick(std::false_type{}, std::false_type{}); }
}
static int scene_march_variant(const omk::RenderArgs& a) {
    if (a.scene.n[omd::PT_MDYN] != 0u) return 3;
    return (a.scene.mgrid_dim[0] | a.scene.n_mbvh_nodes) != 0u ? 1 : 0;
}

size_t render_stack_bytes(uint32_t stack_need) { return size_t(stack_need) * 4u * RENDER_THREADS; }
size_t hook_stack_bytes(uint32_t stack_need) { return size_t(stack_need) * 4u * CTA_THREADS; }

// resident CTAs per SM of the render kernel for this configuration (the same for both MA instances: one launch bound)
int render_occupancy(bool smem, bool count, size_t smem_bytes) {
    int n = 0;
    cudaError_t e = cudaSuccess;
    with_render_kernel(smem, count, 0, [&](auto kern) {
        if (smem) prep(kern, smem_bytes);
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, RENDER_THREADS, smem ? smem_bytes : 0);
    });
    return e == cudaSuccess ? n : 0;
}

cudaError_t launch_render(const omk::RenderArgs& a, bool smem, bool count, int grid, size_t smem_bytes, cudaStream_t st) {
    with_render_kernel(smem, count, scene_march_variant(a), [&](auto kern) {
        if (smem) prep(kern, smem_bytes);
        kern<<<grid, RENDER_THREADS, smem ? smem_bytes : 0, st>>>(a);
    });
    return cudaGetLastError();
}

// k_render_lanes<SMEM, MA> (adaptive passes: a work slot per lane; same block size and shared-memory layout as k_render)
cudaError_t launch_render_lanes(const omk::RenderArgs& a, bool smem, int grid, size_t smem_bytes, cudaStream_t st) {
    const int ma = scene_march_variant(a);
    auto go = [&](auto kern) {
        if (smem) prep(kern, smem_bytes);
        kern<<<grid, RENDER_THREADS, smem ? smem_bytes : 0, st>>>(a);
    };
    if (smem) { if (ma == 1) go(k_render_lanes<true, 1>); else if (ma == 3) go(k_render_lanes<true, 3>); else go(k_render_lanes<true, 0>); }
    else { if (ma == 1) go(k_render_lanes<false, 1>); else if (ma == 3) go(k_render_lanes<false, 3>); else go(k_render_lanes<false, 0>); }
    return cudaGetLastError();
}

// k_render_spread<SMEM, MA> (adaptive passes with few live pixels; same block size and shared-memory layout as k_render)
cudaError_t launch_render_spread(const omk::RenderArgs& a, bool smem, int grid, size_t smem_bytes, cudaStream_t st) {
    const int ma = scene_march_variant(a);
    auto go = [&](auto kern) {
        if (smem) prep(kern, smem_bytes);
        kern<<<grid, RENDER_THREADS, smem ? smem_bytes : 0, st>>>(a);
    };
    if (smem) { if (ma == 1) go(k_render_spread<true, 1>); else if (ma == 3) go(k_render_spread<true, 3>); else go(k_render_spread<true, 0>); }
    else { if (ma == 1) go(k_render_spread<false, 1>); else if (ma == 3) go(k_render_spread<false, 3>); else go(k_render_spread<false, 0>); }
    return cudaGetLastError();
}

cudaError_t launch_primary(const omk::RenderArgs& a, bool smem, int grid, size_t smem_bytes, cudaStream_t st, uint32_t mode, uint32_t sample_index,
                           unsigned long long* ids, float* t) {
    if (smem) { prep(k_primary<true>, smem_bytes); k_primary<true><<<grid, CTA_THREADS, smem_bytes, st>>>(a, mode, sample_index, ids, t); }
    else k_primary<false><<<grid, CTA_THREADS, 0, st>>>(a, mode, sample_index, ids, t);
    return cudaGetLastError();
}

cudaError_t launch_trace(const omk::TraceArgs& a, bool smem, int grid, size_t smem_bytes, cudaStream_t st) {
    if (smem) { prep(k_trace_rays<true>, smem_bytes); k_trace_rays<true><<<grid, CTA_THREADS, smem_bytes, st>>>(a); }
    else k_trace_rays<false><<<grid, CTA_THREADS, 0, st>>>(a);
    return cudaGetLastError();
}

// ---- CTA-resident wavefront (om_pool.cuh) ----
size_t pool_extra_smem(uint32_t stack_need) { return pool_extra_bytes(stack_need); }
template <class F>
static void with_pool_kernel(bool count, int ma, F f) {
    if (count) { if (ma == 1) f(k_render_pool<true, 1>); else if (ma == 3) f(k_render_pool<true, 3>); else f(k_render_pool<true, 0>); }
    else { if (ma == 1) f(k_render_pool<false, 1>); else if (ma == 3) f(k_render_pool<false, 3>); else f(k_render_pool<false, 0>); }
}
int pool_occupancy(bool count, size_t smem_bytes) {
    int n = 0;
    cudaError_t e = cudaSuccess;
    with_pool_kernel(count, 0, [&](auto kern) {
        e = prep(kern, smem_bytes);
        if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, POOL_T, smem_bytes);
    });
    if (e != cudaSuccess) cudaGetLastError();
    return e == cudaSuccess ? n : 0;
}
cudaError_t launch_render_pool(const omk::RenderArgs& a, bool count, int grid, size_t smem_bytes, cudaStream_t st) {
    with_pool_kernel(count, scene_march_variant(a), [&](auto kern) {
        prep(kern, smem_bytes);
        kern<<<grid, POOL_T, smem_bytes, st>>>(a);
    });
    return cudaGetLastError();
}

// ---- wavefront pipeline (om_wavefront.cuh) ----
size_t wf_stage_bytes() { return WF_STAGE_BYTES; }
int wf_extend_occupancy(bool smem, bool count, size_t scene_bytes) {
    int n = 0;
    cudaError_t e;
    const size_t total = (smem ? scene_bytes : 0) + WF_STAGE_BYTES;
    if (smem) {
        if (count) { prep(k_wf_extend<true, true>, total); e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_wf_extend<true, true>, WF_EXT_THREADS, total); }
        else { prep(k_wf_extend<true, false>, total); e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_wf_extend<true, false>, WF_EXT_THREADS, total); }
    } else {
        // the scene is walked through L1 / L2: ask for the smallest shared-memory carve-out that holds the staging lines of the resident
        // CTAs, the rest of the 256 KB goes to L1 (the kernel waits on node fetches: profiles/r2_experiments.md section 8)
        const int carve = getenv("OM_WF_NO_CARVEOUT") ? -1 : int((size_t(OM_WF_EXT_MIN_CTAS) * (total + 1024u) * 100u + 228u * 1024u - 1u) / (228u * 1024u));
        if (carve >= 0) {
            cudaFuncSetAttribute(k_wf_extend<false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            cudaFuncSetAttribute(k_wf_extend<false, false>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
        }
        if (count) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_wf_extend<false, true>, WF_EXT_THREADS, total);
        else e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_wf_extend<false, false>, WF_EXT_THREADS, total);
    }
    return e == cudaSuccess ? n : 0;
}
cudaError_t launch_wf_extend(const omk::RenderArgs& a, const omk::WfArgs& w, bool smem, bool count, int grid, size_t scene_bytes, cudaStream_t st) {
    const size_t total = (smem ? scene_bytes : 0) + WF_STAGE_BYTES;
    if (smem) {
        if (count) k_wf_extend<true, true><<<grid, WF_EXT_THREADS, total, st>>>(a, w);
        else k_wf_extend<true, false><<<grid, WF_EXT_THREADS, total, st>>>(a, w);
    } else {
        if (count) k_wf_extend<false, true><<<grid, WF_EXT_THREADS, total, st>>>(a, w);
        else k_wf_extend<false, false><<<grid, WF_EXT_THREADS, total, st>>>(a, w);
    }
    return cudaGetLastError();
}
cudaError_t launch_wf_shade(const omk::RenderArgs& a, const omk::WfArgs& w, bool count, int grid, cudaStream_t st) {
    if (count) k_wf_shade<true><<<grid, WF_THREADS, 0, st>>>(a, w);
    else k_wf_shade<false><<<grid, WF_THREADS, 0, st>>>(a, w);
    return cudaGetLastError();
}
cudaError_t launch_wf_compact(const omk::WfArgs& w, int grid, cudaStream_t st) {
    k_wf_compact<<<grid, WF_THREADS, 0, st>>>(w);
    return cudaGetLastError();
}

}  // namespace OM_NS
