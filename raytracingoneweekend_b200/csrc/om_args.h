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
       CTR_SDF_EVALS = 6, CTR_MARCH_STEPS = 7, CTR_NORMAL_EVALS = 8, CTR_MARCH_ENTERED = 9,
       CTR_WARP_ITERS = 13 /* path-loop iterations of k_render, per warp */, CTR_LANE_ITERS = 14 /* ... summed over the lanes that carried a path */, CTR_COUNT = 16 };

}  // namespace ctr

struct RenderArgs {
    SceneHeader scene;
    const float4* hot_g;
    const float4* cold_g;
    om_camera cam;
    uint32_t width, height, npx;
    uint32_t spp_budget;   // samples this handle owns per pixel
    uint32_t spp_pass;     // at most this many more samples per pixel in this launch
    uint32_t n_target;     // if non-zero: bring every live pixel to min(spp_budget, n_target) samples instead (wavefront tail)
    uint32_t sample_base;  // global index of the handle's first sample (spp-slice partition)
    uint32_t max_depth;
    float tmin, tmax;
    uint32_t seed_lo, seed_hi;
    uint32_t rk[20];       // Philox round keys: rk[2r] = seed_lo + r * 0x9E3779B9, rk[2r+1] = seed_hi + r * 0xBB67AE85 (r = 0..9)
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
    // Tile scheduling (om_api_render.inl, refresh_block_order): the order in which the 16 x 16-pixel blocks are handed out
    // (most expensive first) or nullptr = row order; and where k_render leaves each tile's cost (cycles / 16) for the next order
    const uint32_t* block_order;
    uint32_t* tile_cost;
    // Adaptive passes over a compacted live list (k_render_spread, om_device.cuh): when at most spread_max pixels are live, `spread`
    // lanes share a pixel, each traces one of its next samples, and the results are folded into the pixel's Stats in sample order.
    // 0 = never.  Both kernels are launched; each looks at *live_count and the one whose turn it is not returns at once.
    uint32_t spread_max, spread;
};

// Wavefront pipeline (om_wavefront.cuh).  A SLOT carries one path; it is bound to one pixel until
// that pixel has all the samples of this launch, then takes the next unstarted pixel.
namespace wf {
// per-iteration counter set (three sets rotate: iteration i uses set i % 3, pushes the next
// iteration's rays into set (i+1) % 3 and clears set (i+2) % 3)
// every counter sits on its own 128-byte line: same-line atomics serialise in one L2 slice
enum { C_STRIDE = 32, C_N_RAY = 0, C_N_MARCH = 1 * C_STRIDE, C_CUR_RAY = 2 * C_STRIDE, C_N_SHADE = 3 * C_STRIDE /* + C_STRIDE * class */, C_SET_WORDS = 8 * C_STRIDE };
// shade classes: what a finished segment needs next (a warp of k_wf_shade only ever sees one class)
enum { CLS_SKY = 0, CLS_LAMBERTIAN = 1, CLS_METAL = 2, CLS_DIELECTRIC = 3, CLS_OTHER = 4 /* hit on anything but a Sphere */, N_CLS = 5 };
// global words after the three sets
enum { G_NEXT_WORK = 3 * C_SET_WORDS, G_WORDS = 3 * C_SET_WORDS + C_STRIDE };
}  // namespace wf
struct WfArgs {
    uint32_t n_slots;
    uint32_t it;          // iteration index
    uint32_t n_target;    // every live pixel is brought to min(spp_budget, n_target) samples
    uint32_t init;        // shade kernel: 1 = first launch (every slot takes a pixel; nothing to shade)
    float4* s_ray;        // 2 per slot: {origin.xyz, sample index (uint bits)}, {direction.xyz (unit), pixel (uint bits, 0xFFFFFFFF = free)}
    float4* s_st;         // 4 per slot: {throughput.rgb, primary-hit t}, {segment, Philox draw, bloom lo, hi},
                          //             {hit t, hit ref, march t, march iterations left (0xFFFFFFFF: unstuck first)}, unused
    uint8_t* s_live;      // 1 = the slot has a ray for the next iteration (set by k_wf_shade, gathered and cleared by k_wf_compact)
    uint32_t* q_ray[2];   // slots with a ray to extend (ping-pong)
    uint32_t* q_march[2]; // slots whose ray enters the marched objects' bounds before its traced hit (ping-pong: marches span iterations)
    uint32_t* q_shade[2]; // each: wf::N_CLS queues of n_slots entries: slots whose segment is resolved, by shade class (ping-pong:
                          // a march that ends during iteration i queues its slot for the shading of iteration i + 1)
    uint32_t* ctr;        // wf::G_WORDS words
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
