// Kernel argument blocks shared by the two builds of om_device.cuh and by the host API.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ottomarcher.h"
#include "om_dev_types.h"

namespace omk {
using omd::SceneHeader;

namespace ctr {
// slots of RenderArgs::counters; the CTR_SEGMENTS.. ones are only written by the COUNT kernels
enum { CTR_CREDITED = 0, CTR_TRACED = 1, CTR_TICKET = 2, CTR_SEGMENTS = 3, CTR_NODE_VISITS = 4, CTR_PRIM_TESTS = 5,
       CTR_SDF_EVALS = 6, CTR_MARCH_STEPS = 7, CTR_NORMAL_EVALS = 8, CTR_MARCH_ENTERED = 9, CTR_COUNT = 16 };

}  // namespace ctr

struct RenderArgs {
    SceneHeader scene;
    const float4* hot_g;
    const float4* cold_g;
    om_camera cam;
    uint32_t width, height, npx;
    uint32_t spp_budget;   // samples this handle owns per pixel
    uint32_t spp_pass;     // at most this many more samples per pixel in this launch
    uint32_t sample_base;  // global index of the handle's first sample (spp-slice partition)
    uint32_t max_depth;
    float tmin, tmax;
    uint32_t seed_lo, seed_hi;
    uint32_t flags;        // OM_RENDER_*
    uint32_t part_rank, part_world;
    uint32_t n_work;       // work slots (pixels padded to whole 8x4 tiles / chunks)
    uint32_t tiles_x;
    const uint8_t* jitter; // spp_max entries, bit1 = jitter i, bit0 = jitter j
    float4* acc;           // per pixel {sum.r, sum.g, sum.b, n (uint bits)}
    uint4* aux;            // per pixel {avg_depth (float bits), bad_avgs, bloom lo, bloom hi}
    unsigned long long* counters;
    const uint32_t* live_list;  // adaptive mode: compacted list of live pixels, or nullptr
    const uint32_t* live_count;
};

struct TraceArgs {
    SceneHeader scene;
    const float4* hot_g;
    const float4* cold_g;
    const float* origins;
    const float* dirs;
    uint32_t n;
    float tmin, tmax;
    uint32_t flags;
    unsigned long long* ids;
    float* t;
    float* points;
    float* normals;
};

}  // namespace omk
