// Declarations of the launch wrappers of both kernel builds (see om_launch.inl).
#pragma once
#include "om_args.h"

#define OM_DECLARE_KERNELS(NS)                                                                                                     \
    namespace NS {                                                                                                                 \
    int render_occupancy(bool smem, bool count, size_t smem_bytes);                                                                \
    cudaError_t launch_render(const omk::RenderArgs& a, bool smem, bool count, int grid, size_t smem_bytes, cudaStream_t st);      \
    cudaError_t launch_primary(const omk::RenderArgs& a, bool smem, int grid, size_t smem_bytes, cudaStream_t st, uint32_t mode,   \
                               uint32_t sample_index, unsigned long long* ids, float* t);                                         \
    cudaError_t launch_trace(const omk::TraceArgs& a, bool smem, int grid, size_t smem_bytes, cudaStream_t st);                    \
    }
OM_DECLARE_KERNELS(om_exact)
OM_DECLARE_KERNELS(om_fast)
