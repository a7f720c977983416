// Host-side launch wrappers, compiled once per build of om_device.cuh (namespace OM_NS).
namespace OM_NS {

template <class K>
static cudaError_t prep(K kern, size_t smem_bytes) {
    return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
}

// resident CTAs per SM of the render kernel for this configuration
int render_occupancy(bool smem, bool count, size_t smem_bytes) {
    int n = 0;
    cudaError_t e;
    if (smem) {
        if (count) { prep(k_render<true, true>, smem_bytes); e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_render<true, true>, RENDER_THREADS, smem_bytes); }
        else { prep(k_render<true, false>, smem_bytes); e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_render<true, false>, RENDER_THREADS, smem_bytes); }
    } else {
        if (count) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_render<false, true>, RENDER_THREADS, 0);
        else e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_render<false, false>, RENDER_THREADS, 0);
    }
    return e == cudaSuccess ? n : 0;
}

cudaError_t launch_render(const omk::RenderArgs& a, bool smem, bool count, int grid, size_t smem_bytes, cudaStream_t st) {
    if (smem) {
        if (count) { prep(k_render<true, true>, smem_bytes); k_render<true, true><<<grid, RENDER_THREADS, smem_bytes, st>>>(a); }
        else { prep(k_render<true, false>, smem_bytes); k_render<true, false><<<grid, RENDER_THREADS, smem_bytes, st>>>(a); }
    } else {
        if (count) k_render<false, true><<<grid, RENDER_THREADS, 0, st>>>(a);
        else k_render<false, false><<<grid, RENDER_THREADS, 0, st>>>(a);
    }
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
