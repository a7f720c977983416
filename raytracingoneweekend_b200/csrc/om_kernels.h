// Declarations of the launch wrappers of both kernel builds (see om_launch.inl).
#pragma once
#include "om_args.h"

#define OM_DECLARE_KERNELS(NS)                                                                                                     \
    namespace NS {                                                                                                                 \
    /* shared-memory traversal stacks behind the staged scene image (TravStack<true, ..>): bytes per CTA of k_render / of the parity hooks */ \
    size_t render_stack_bytes(uint32_t stack_need);                                                                                \
    size_t hook_stack_bytes(uint32_t stack_need);                                                                                  \
    int render_occupancy(bool smem, bool count, size_t smem_bytes);                                                                \
    cudaError_t launch_render(const omk::RenderArgs& a, bool smem, bool count, int grid, size_t smem_bytes, cudaStream_t st);      \
    cudaError_t launch_render_spread(const omk::RenderArgs& a, bool smem, int grid, size_t smem_bytes, cudaStream_t st);           \
    cudaError_t launch_render_lanes(const omk::RenderArgs& a, bool smem, int grid, size_t smem_bytes, cudaStream_t st);            \
    /* CTA-resident wavefront (om_pool.cuh): shared-memory bytes behind the scene image, and the launch */                        \
    size_t pool_extra_smem(uint32_t stack_need);                                                                                   \
    int pool_occupancy(bool count, size_t smem_bytes);                                                                             \
    cudaError_t launch_render_pool(const omk::RenderArgs& a, bool count, int grid, size_t smem_bytes, cudaStream_t st);            \
    int wf_extend_occupancy(bool smem, bool count, size_t smem_bytes);                                                             \
    cudaError_t launch_wf_extend(const omk::RenderArgs& a, const omk::WfArgs& w, bool smem, bool count, int grid, size_t smem_bytes, cudaStream_t st); \
    size_t wf_stage_bytes();                                                                                                       \
    cudaError_t launch_wf_compact(const omk::WfArgs& w, int grid, cudaStream_t st);                                              \
    cudaError_t launch_wf_shade(const omk::RenderArgs& a, const omk::WfArgs& w, bool count, int grid, cudaStream_t st);          \
    cudaError_t launch_primary(const omk::RenderArgs& a, bool smem, int grid, size_t smem_bytes, cudaStream_t st, uint32_t mode,   \
                               uint32_t sample_index, unsigned long long* ids, float* t);                                         \
    cudaError_t launch_trace(const omk::TraceArgs& a, bool smem, int grid, size_t smem_bytes, cudaStream_t st);                    \
    }
OM_DECLARE_KERNELS(om_exact)
OM_DECLARE_KERNELS(om_fast)
