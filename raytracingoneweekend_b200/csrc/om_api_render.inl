// Second half of om_api.cu: framebuffer kernels and the om_render handle.

namespace {

constexpr int UT = 256;  // threads of the small framebuffer kernels
inline int ublocks(size_t n) { return int((n + UT - 1) / UT); }

// utils.rs:12-17 + vec3.rs:73-76
__device__ __forceinline__ uint32_t d_as_u8(float f) {
    if (!(f == f) || f <= 0.f) return 0u;
    if (f >= 255.f) return 255u;
    return (uint32_t)f;
}
__device__ __forceinline__ uint32_t d_tone(float sum, float inv_n) { return d_as_u8(fminf(fmaxf(sqrtf(sum * inv_n), 0.f), 0.999f) * 256.0f); }
__device__ __forceinline__ void d_color(float4 a, uint32_t c[3]) {
    const uint32_t n = __float_as_uint(a.w);
    if (n == 0u) { c[0] = c[1] = c[2] = 0u; return; }  // Stats::new
    const float inv_n = 1.f / (float)n;                 // Div<f32> = multiply by the reciprocal
    c[0] = d_tone(a.x, inv_n); c[1] = d_tone(a.y, inv_n); c[2] = d_tone(a.z, inv_n);
}

// Vec<Pixel> view of the accumulators (om_pixel, 40 bytes = 10 words)
__global__ void k_pack_pixels(const float4* __restrict__ acc, const uint4* __restrict__ aux, uint32_t* __restrict__ out, uint32_t npx) {
    const uint32_t i = blockIdx.x * UT + threadIdx.x;
    if (i >= npx) return;
    const float4 a = acc[i];
    const uint4 x = aux[i];
    uint32_t c[3];
    d_color(a, c);
    uint32_t* o = out + 10u * i;
    o[0] = __float_as_uint(a.x); o[1] = __float_as_uint(a.y); o[2] = __float_as_uint(a.z); o[3] = __float_as_uint(a.w);
    o[4] = x.x; o[5] = x.y; o[6] = x.z; o[7] = x.w;
    o[8] = c[0] | (c[1] << 8) | (c[2] << 16);
    o[9] = 0u;
}
// MODE_NORMAL view, main.rs:374-381
__global__ void k_rgb(const float4* __restrict__ acc, uint8_t* __restrict__ rgb, uint32_t npx) {
    const uint32_t i = blockIdx.x * UT + threadIdx.x;
    if (i >= npx) return;
    uint32_t c[3];
    d_color(acc[i], c);
    rgb[3u * i] = (uint8_t)c[0]; rgb[3u * i + 1] = (uint8_t)c[1]; rgb[3u * i + 2] = (uint8_t)c[2];
}
__global__ void k_export_accum(const float4* __restrict__ acc, const uint4* __restrict__ aux, float4* __restrict__ out, unsigned long long* __restrict__ bloom, uint32_t npx) {
    const uint32_t i = blockIdx.x * UT + threadIdx.x;
    if (i >= npx) return;
    const float4 a = acc[i];
    const uint4 x = aux[i];
    const float n = (float)__float_as_uint(a.w);
    const float d = __uint_as_float(x.x);
    out[2u * i] = make_float4(a.x, a.y, a.z, n);
    out[2u * i + 1] = make_float4(n > 0.f ? n * d : 0.f, (float)x.y, 0.f, 0.f);
    if (bloom) bloom[i] = ((unsigned long long)x.w << 32) | x.z;
}
__global__ void k_import_accum(float4* __restrict__ acc, uint4* __restrict__ aux, const float4* __restrict__ in, const unsigned long long* __restrict__ bloom, uint32_t npx) {
    const uint32_t i = blockIdx.x * UT + threadIdx.x;
    if (i >= npx) return;
    const float4 a = in[2u * i], b = in[2u * i + 1];
    const uint32_t n = (uint32_t)(a.w + 0.5f);
    acc[i] = make_float4(a.x, a.y, a.z, __uint_as_float(n));
    uint4 x = aux[i];
    x.x = __float_as_uint(n ? b.x / a.w : 0.f);
    x.y = (uint32_t)(b.y + 0.5f);
    if (bloom) { x.z = (uint32_t)bloom[i]; x.w = (uint32_t)(bloom[i] >> 32); }
    aux[i] = x;
}
__global__ void k_bloom_or(unsigned long long* __restrict__ dst, const unsigned long long* __restrict__ src, uint32_t nsrc, size_t n) {
    const size_t i = (size_t)blockIdx.x * UT + threadIdx.x;
    if (i >= n) return;
    unsigned long long v = dst[i];
    for (uint32_t k = 0; k < nsrc; k++) v |= src[(size_t)k * n + i];
    dst[i] = v;
}
// ThreadPixels::add_run / swap_buffers (render_thread.rs:92-101) for the whole image: the list
// of pixels that still want samples, by warp ballot + popc and one atomic per warp.
// Who owns pixel i: the chunk owner table when one is set (om_render_set_chunk_owners), else the interleaved-chunk rule.
__device__ __forceinline__ bool pixel_is_mine(uint32_t i, uint32_t part_rank, uint32_t part_world, const uint8_t* __restrict__ owners, uint32_t my_id) {
    if (owners) return owners[i / OM_CHUNK_PIXELS] == my_id;
    return part_world <= 1u || ((i / OM_CHUNK_PIXELS) % part_world) == part_rank;
}
// The list is written in the render kernel's tile order (a warp here = an 8x4 tile, a CTA = a 16x16 block), so that the 32 live pixels
// a render warp takes from it are neighbours in both directions: strips of 32 pixels of one image row mixed more kinds of pixels per
// warp (lane occupancy of the adaptive passes 0.59 against 0.72 for tiles; the warp waits for its longest pixel).
__global__ void k_compact_live(const float4* __restrict__ acc, const uint4* __restrict__ aux, uint32_t width, uint32_t height, uint32_t budget, uint32_t adaptive,
                               uint32_t part_rank, uint32_t part_world, const uint8_t* __restrict__ owners, uint32_t my_id, uint32_t* __restrict__ list,
                               uint32_t* __restrict__ count) {
    const uint32_t t = blockIdx.x * UT + threadIdx.x, tile = t >> 5, in = t & 31u;
    const uint32_t tiles_x = (width + 7u) / 8u, bx_n = (tiles_x + 1u) >> 1, blk = tile >> 3, k = tile & 7u;
    const uint32_t x = ((blk % bx_n) * 2u + (k & 1u)) * 8u + (in & 7u), y = ((blk / bx_n) * 4u + (k >> 1)) * 4u + (in >> 3);
    const uint32_t i = y * width + x;
    bool live = false;
    if (x < width && y < height) {
        const bool mine = pixel_is_mine(i, part_rank, part_world, owners, my_id);
        const uint32_t n = __float_as_uint(acc[i].w);
        live = mine && n < budget && !(adaptive && aux[i].y >= 5u);
    }
    const uint32_t m = __ballot_sync(0xffffffffu, live);
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t base = 0;
    if (lane == 0u && m) base = atomicAdd(count, (uint32_t)__popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (live) list[base + __popc(m & ((1u << lane) - 1u))] = i;
}
// ---- multi-GPU adaptive scheduling: raw pixel state in and out, live pixels per chunk -------------------------------
// 8 words per pixel = {sum.r, sum.g, sum.b, n, avg_depth, bad_avgs, bloom lo, bloom hi}, bit for bit; pixels of other
// owners are written as zeros so that an integer SUM over the ranks reassembles the frame exactly.
__global__ void k_export_state(const float4* __restrict__ acc, const uint4* __restrict__ aux, uint4* __restrict__ out, uint32_t npx, uint32_t only_owned,
                               uint32_t part_rank, uint32_t part_world, const uint8_t* __restrict__ owners, uint32_t my_id) {
    const uint32_t i = blockIdx.x * UT + threadIdx.x;
    if (i >= npx) return;
    uint4 a = make_uint4(0u, 0u, 0u, 0u), b = a;
    if (!only_owned || pixel_is_mine(i, part_rank, part_world, owners, my_id)) {
        const float4 f = acc[i];
        a = make_uint4(__float_as_uint(f.x), __float_as_uint(f.y), __float_as_uint(f.z), __float_as_uint(f.w));
        b = aux[i];
    }
    out[2u * i] = a; out[2u * i + 1u] = b;
}
__global__ void k_import_state(float4* __restrict__ acc, uint4* __restrict__ aux, const uint4* __restrict__ in, uint32_t npx) {
    const uint32_t i = blockIdx.x * UT + threadIdx.x;
    if (i >= npx) return;
    const uint4 a = in[2u * i];
    acc[i] = make_float4(__uint_as_float(a.x), __uint_as_float(a.y), __uint_as_float(a.z), __uint_as_float(a.w));
    aux[i] = in[2u * i + 1u];
}
__global__ void k_live_per_chunk(const float4* __restrict__ acc, const uint4* __restrict__ aux, uint32_t npx, uint32_t budget, uint32_t adaptive,
                                 uint32_t* __restrict__ counts) {
    const uint32_t i = blockIdx.x * UT + threadIdx.x;
    bool live = false;
    if (i < npx) {
        const uint32_t n = __float_as_uint(acc[i].w);
        live = n < budget && !(adaptive && aux[i].y >= 5u);
    }
    // one atomic per warp and chunk (a warp's 32 consecutive pixels lie in one chunk or two): an atomic per live pixel on 760
    // counters took 0.58 ms for a full-HD frame, an eighth of a pass of the 8-GPU adaptive render
    const uint32_t c = live ? i / OM_CHUNK_PIXELS : 0xFFFFFFFFu;
    const unsigned peers = __match_any_sync(0xffffffffu, c);
    if (live && (threadIdx.x & 31u) == (uint32_t)(__ffs(peers) - 1)) atomicAdd(&counts[c], (uint32_t)__popc(peers));
}
// ---- draw_to_sdl's display modes (main.rs:361-437) and its three weighted 3x3 filters (main.rs:219-343) ----
// reductions of MODE_SHOW_SAMPLES / MODE_SHOW_DEPTH: red[0] = max n (starts at 1), red[1] = bits of the largest finite
// avg_depth >= 0 (bits order like values there), red[2] = 1 if one was seen (else the reference keeps -1)
__global__ void k_view_reduce(const float4* __restrict__ acc, const uint4* __restrict__ aux, uint32_t npx, uint32_t* __restrict__ red) {
    const uint32_t i = blockIdx.x * UT + threadIdx.x;
    uint32_t n = 1u, dbits = 0u, seen = 0u;
    if (i < npx) {
        n = max(1u, __float_as_uint(acc[i].w));
        const float d = __uint_as_float(aux[i].x);
        if (d > -1.f && !isinf(d)) { seen = 1u; dbits = d > 0.f ? __float_as_uint(d) : 0u; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        n = max(n, __shfl_xor_sync(0xffffffffu, n, o));
        dbits = max(dbits, __shfl_xor_sync(0xffffffffu, dbits, o));
        seen |= __shfl_xor_sync(0xffffffffu, seen, o);
    }
    if ((threadIdx.x & 31u) == 0u) { atomicMax(&red[0], n); if (seen) { atomicMax(&red[1], dbits); atomicOr(&red[2], 1u); } }
}
struct ViewPx {
    float sx, sy, sz, n, depth;
    unsigned long long bloom;
};
__device__ __forceinline__ ViewPx view_px(const float4* acc, const uint4* aux, uint32_t i) {
    const float4 a = acc[i];
    const uint4 x = aux[i];
    ViewPx p;
    p.sx = a.x; p.sy = a.y; p.sz = a.z; p.n = (float)__float_as_uint(a.w); p.depth = __uint_as_float(x.x);
    p.bloom = ((unsigned long long)x.w << 32) | x.z;
    return p;
}
__device__ __forceinline__ uint32_t view_tone(float r, float g, float b) {  // normalize_color(..).to_u8x3()
    return d_as_u8(fminf(fmaxf(sqrtf(r), 0.f), 0.999f) * 256.0f) | (d_as_u8(fminf(fmaxf(sqrtf(g), 0.f), 0.999f) * 256.0f) << 8) |
           (d_as_u8(fminf(fmaxf(sqrtf(b), 0.f), 0.999f) * 256.0f) << 16);
}
__global__ void k_view(const float4* __restrict__ acc, const uint4* __restrict__ aux, uint8_t* __restrict__ rgb, uint32_t W, uint32_t H, uint32_t mode,
                       const uint32_t* __restrict__ red) {
    const uint32_t idx = blockIdx.x * UT + threadIdx.x;
    if (idx >= W * H) return;
    const int i = (int)(idx % W), j = (int)(idx / W);
    uint32_t c = 0;
    if (mode == 0u) {  // MODE_NORMAL
        uint32_t cc[3];
        d_color(acc[idx], cc);
        c = cc[0] | (cc[1] << 8) | (cc[2] << 16);
    } else if (mode == 1u) {  // MODE_SHOW_SAMPLES main.rs:383-398
        const float smpls = (float)__float_as_uint(acc[idx].w) / (float)red[0];
        c = view_tone(smpls, smpls, smpls);
    } else if (mode == 3u) {  // MODE_SHOW_DEPTH main.rs:399-418
        const float max_depth = red[2] ? __uint_as_float(red[1]) : -1.f;
        const float depth_0to1 = __uint_as_float(aux[idx].x) / max_depth;
        const bool is_inf = isinf(depth_0to1);
        c = view_tone(is_inf ? 0.f : depth_0to1, is_inf ? 1.f : depth_0to1, is_inf ? 0.f : depth_0to1);
    } else if (mode == 5u) {  // MODE_SHOW_IDS main.rs:428-436, utils.rs:45-69
        const uint4 x = aux[idx];
        const unsigned long long st = ((unsigned long long)x.w << 32) | x.z;
        unsigned long long id1 = st & 0xFFFFFFFFull;
        id1 ^= id1 << 13; id1 ^= id1 >> 7; id1 ^= id1 << 17;
        unsigned long long id2 = st >> 32;
        id2 ^= id2 << 13; id2 ^= id2 >> 17; id2 ^= id2 << 5;
        const unsigned long long id = (id2 << 32) ^ id1 ^ (id1 * id2);
        uint32_t b[8];
        for (int k = 0; k < 8; k++) b[k] = (uint32_t)((id >> (8 * k)) & 0xFFull);
        c = (b[0] ^ b[7] ^ b[3]) | ((b[1] ^ b[4] ^ b[5]) << 8) | ((b[2] ^ b[6]) << 16);
    } else {  // the weighted 3x3 filters: 2 samples, 4 depth, 6 ids (apply_box_filter_ij_*, main.rs:222-318)
        const int min_x = i == 0 ? 0 : -1, max_x = i == (int)W - 1 ? 0 : 1, min_y = j == 0 ? 0 : -1, max_y = j == (int)H - 1 ? 0 : 1;  // main.rs:324-342
        const ViewPx ctr = view_px(acc, aux, idx);
        const float SQRT2_INV = 0.7071067811865475244f;
        if (mode == 4u && isinf(ctr.depth)) {
            uint32_t cc[3];
            d_color(acc[idx], cc);
            c = cc[0] | (cc[1] << 8) | (cc[2] << 16);
        } else {
            float total_weight = 0.f, cr = 0.f, cg = 0.f, cb = 0.f;
            for (int y = min_y; y <= max_y; y++) {
                for (int x = min_x; x <= max_x; x++) {
                    const ViewPx q = view_px(acc, aux, (uint32_t)((i + x) + (j + y) * (int)W));
                    const float is_diagonal = (x != 0 && y != 0) ? 1.f : 0.f;
                    if (mode == 2u) {
                        const float diag_w = 1.f - (1.f - SQRT2_INV) * is_diagonal;
                        total_weight += q.n * diag_w;
                        cr = cr + q.sx * diag_w; cg = cg + q.sy * diag_w; cb = cb + q.sz * diag_w;
                    } else {
                        float w;
                        if (mode == 4u) w = 1.f / (1.f + fabsf(q.depth - ctr.depth));
                        else w = ((q.bloom == ctr.bloom) ? 1.f : 0.f) + (((q.bloom & ctr.bloom) == ctr.bloom) ? 1.f : 0.f);
                        const float diag_w = w * (1.f - (1.f - SQRT2_INV) * is_diagonal);
                        total_weight += diag_w;
                        const float inv_n = 1.f / q.n;  // Div<f32> = multiply by the reciprocal
                        cr = cr + diag_w * (q.sx * inv_n); cg = cg + diag_w * (q.sy * inv_n); cb = cb + diag_w * (q.sz * inv_n);
                    }
                }
            }
            const float inv_w = 1.f / total_weight;
            c = view_tone(cr * inv_w, cg * inv_w, cb * inv_w);
        }
    }
    rgb[3u * idx] = (uint8_t)(c & 0xFFu); rgb[3u * idx + 1] = (uint8_t)((c >> 8) & 0xFFu); rgb[3u * idx + 2] = (uint8_t)((c >> 16) & 0xFFu);
}
// dependent-free FMA chains: the FP32 roofline denominator
__global__ void k_fma_peak(float* out, int iters) {
    float a[8];
#pragma unroll
    for (int k = 0; k < 8; k++) a[k] = 1.0f + 0.001f * (float)(threadIdx.x + k);
    const float m = 0.9999f, c = 0.0001f;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = fmaf(a[k], m, c);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; k++) s += a[k];
    if (s == 12345.678f) out[0] = s;
}

struct EventPair {
    cudaEvent_t a = nullptr, b = nullptr;
};

}  // namespace

struct om_render {
    om_scene* scene = nullptr;
    om_camera cam;
    om_render_params p;
    int dev = 0;
    DeviceProps props;
    DevScene ds;
    uint32_t npx = 0, budget = 0;
    cudaStream_t stream = nullptr;      // the stream work is issued on
    cudaStream_t own_stream = nullptr;  // created by the handle
    float4* acc = nullptr;
    uint4* aux = nullptr;
    uint8_t* jitter = nullptr;
    unsigned long long* counters = nullptr;
    uint8_t* owners = nullptr;   // chunk -> owner id table (om_render_set_chunk_owners), device
    uint32_t owner_id = 0;
    uint32_t* chunk_counts = nullptr;
    uint32_t* live_list = nullptr;
    uint32_t* live_count = nullptr;
    bool have_live = false;
    void* staging = nullptr;  // device staging for read-outs
    size_t staging_bytes = 0;
    std::vector<EventPair> pending, pool;
    float acc_ms = 0.f;
    uint32_t launches = 0;
    uint64_t kernels = 0;         // kernels launched by om_render_launch since create (om_render_perf_counters [10])
    uint64_t graph_launches = 0;  // of which through CUDA graphs: graph launches (om_render_perf_counters [15])
    bool use_smem = false;       // k_render: scene image in shared memory (one 1024-thread CTA per SM)
    bool use_smem_hook = false;  // parity-hook kernels (two 256-thread CTAs per SM)
    size_t smem_bytes = 0;       // k_render: scene image + traversal stacks
    size_t smem_hook_bytes = 0;  // parity hooks: the same for their 256-thread CTAs
    int grid = 0;
    // wavefront pipeline (OM_RENDER_WAVEFRONT)
    omk::WfArgs wf;               // device arrays; n_slots == 0: not allocated
    int wf_grid_extend = 0, wf_grid_small = 0;
    bool wf_smem = false;
    bool use_pool = false;       // OM_RENDER_POOL: k_render_pool instead of k_render
    size_t pool_smem_bytes = 0;  // scene image + traversal stacks + slot pool + queues
    int pool_grid = 0;
    uint32_t* wf_host_ctr = nullptr;  // pinned: wf::G_WORDS words per in-flight batch (2 batches)
    cudaEvent_t wf_ev[2] = {nullptr, nullptr};
    uint32_t n_uniform = 0;       // samples every live pixel has after the launches so far
    // tile scheduling: see refresh_block_order
    uint32_t* block_order = nullptr;
    uint32_t* tile_cost = nullptr;
    bool have_order = false;
};

namespace {
om_status ensure_staging(om_render* r, size_t bytes) {
    if (r->staging_bytes >= bytes) return OM_OK;
    if (r->staging) dev_free(r->staging);
    r->staging = nullptr;
    r->staging_bytes = 0;
    CU(dev_alloc(&r->staging, bytes));
    r->staging_bytes = bytes;
    return OM_OK;
}
omk::RenderArgs make_args(const om_render* r) {
    omk::RenderArgs a;
    memset(&a, 0, sizeof a);
    a.scene = r->scene->fz.hdr;
    a.hot_g = r->ds.hot;
    a.cold_g = r->ds.cold;
    a.cam = r->cam;
    a.width = r->p.width; a.height = r->p.height; a.npx = r->npx;
    a.spp_budget = r->budget;
    a.sample_base = r->p.sample_base;
    a.max_depth = r->p.max_depth;
    a.tmin = r->p.tmin; a.tmax = r->p.tmax;
    a.seed_lo = uint32_t(r->p.seed); a.seed_hi = uint32_t(r->p.seed >> 32);
    for (uint32_t k = 0; k < 10; k++) { a.rk[2 * k] = a.seed_lo + k * 0x9E3779B9u; a.rk[2 * k + 1] = a.seed_hi + k * 0xBB67AE85u; }
    a.flags = r->p.flags;
    a.block_order = r->have_order ? r->block_order : nullptr;
    a.tile_cost = nullptr;
    a.part_rank = r->p.part_rank; a.part_world = r->p.part_world ? r->p.part_world : 1u;
    if (a.part_world > 1u) {
        const uint32_t chunks = (r->npx + OM_CHUNK_PIXELS - 1) / OM_CHUNK_PIXELS;
        const uint32_t mine = (chunks + a.part_world - 1 - a.part_rank) / a.part_world;
        a.n_work = mine * OM_CHUNK_PIXELS;
        a.tiles_x = 0;
    } else {
        a.tiles_x = (r->p.width + 7u) / 8u;
        // tiles are handed out in blocks of 2 x 4 (16 x 16 pixels, see work_to_pixel): whole blocks, the surplus maps to no pixel
        a.n_work = ((a.tiles_x + 1u) / 2u) * ((((r->p.height + 3u) / 4u) + 3u) / 4u) * 8u * 32u;
    }
    a.jitter = r->jitter;
    a.acc = r->acc; a.aux = r->aux;
    a.counters = r->counters;
    return a;
}
om_status drain_events(om_render* r) {
    for (EventPair& e : r->pending) {
        CU(cudaEventSynchronize(e.b));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, e.a, e.b));
        r->acc_ms += ms;
        r->pool.push_back(e);
    }
    r->pending.clear();
    return OM_OK;
}
}  // namespace

extern "C" {

uint64_t om_bloom_get_hash(uint64_t id) { return omh::bloom_get_hash(id); }
void om_jitter_table(uint32_t spp, uint64_t seed, uint8_t* out) {
    std::vector<uint8_t> t;
    omh::jitter_table(spp, seed, t);
    if (spp) memcpy(out, t.data(), spp);
}

om_status om_render_destroy(om_render* r) {
    if (!r) return OM_OK;
    int prev = 0;
    cudaGetDevice(&prev);
    cudaSetDevice(r->dev);
    if (r->stream) cudaStreamSynchronize(r->stream);
    for (EventPair& e : r->pending) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
    for (EventPair& e : r->pool) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
    cudaFree(r->block_order); cudaFree(r->tile_cost); dev_free(r->acc); dev_free(r->aux); cudaFree(r->jitter); cudaFree(r->counters); dev_free(r->live_list); cudaFree(r->live_count); cudaFree(r->owners); cudaFree(r->chunk_counts); dev_free(r->staging);
    if (r->wf.n_slots) {
        dev_free(r->wf.s_ray); dev_free(r->wf.s_st); dev_free(r->wf.s_live);
        dev_free(r->wf.q_ray[0]); dev_free(r->wf.q_ray[1]); dev_free(r->wf.q_march[0]); dev_free(r->wf.q_march[1]); dev_free(r->wf.q_shade[0]); dev_free(r->wf.q_shade[1]); cudaFree(r->wf.ctr);
    }
    if (r->wf_host_ctr) cudaFreeHost(r->wf_host_ctr);
    for (int i = 0; i < 2; i++) if (r->wf_ev[i]) cudaEventDestroy(r->wf_ev[i]);
    if (r->own_stream) cudaStreamDestroy(r->own_stream);
    cudaSetDevice(prev);
    if (r->scene) r->scene->handles.fetch_sub(1);
    delete r;
    return OM_OK;
}

om_status om_render_create(om_render** out, const om_scene* frozen, const om_camera* cam, const om_render_params* params) {
    if (!out || !frozen || !cam || !params) return fail(OM_ERR_INVALID, "om_render_create: NULL argument");
    *out = nullptr;
    if (!frozen->frozen) return fail(OM_ERR_INVALID, "om_render_create: scene is not frozen (om_scene_freeze)");
    const om_render_params& p = *params;
    if (p.width < 2 || p.height < 2) return fail(OM_ERR_INVALID, "om_render_create: image must be at least 2x2 (u,v divide by W-1,H-1; render_thread.rs:190-191)");
    if (uint64_t(p.width) * p.height > 0x7FFFFFFFull) return fail(OM_ERR_INVALID, "om_render_create: image too large");
    if (p.spp_max == 0) return fail(OM_ERR_INVALID, "om_render_create: spp_max is 0");
    if (p.sample_base >= p.spp_max || uint64_t(p.sample_base) + p.sample_count > p.spp_max) return fail(OM_ERR_INVALID, "om_render_create: sample slice outside [0, spp_max)");
    if (p.part_world > 1 && p.part_rank >= p.part_world) return fail(OM_ERR_INVALID, "om_render_create: part_rank >= part_world");
    if ((p.flags & OM_RENDER_ADAPTIVE) && (p.sample_base != 0 || (p.sample_count != 0 && p.sample_count != p.spp_max)))
        return fail(OM_ERR_INVALID, "om_render_create: adaptive retirement needs all of a pixel's samples in order; use the chunk partition, not an spp slice");
    int dev = 0;
    om_status st = pick_device(p.device, dev);
    if (st != OM_OK) return st;
    om_render* r = new (std::nothrow) om_render();
    if (!r) return fail(OM_ERR_OOM, "om_render_create: out of host memory");
    r->cam = *cam;
    r->p = p;
    if ((p.flags & OM_RENDER_WAVEFRONT) && (p.flags & OM_RENDER_PERSISTENT)) { delete r; return fail(OM_ERR_INVALID, "om_render_create: OM_RENDER_WAVEFRONT and OM_RENDER_PERSISTENT exclude each other"); }
    r->scene = const_cast<om_scene*>(frozen);
    r->scene->handles.fetch_add(1);  // released by om_render_destroy (also on the failure paths below)
    r->dev = dev;
    memset(&r->wf, 0, sizeof r->wf);
    r->npx = p.width * p.height;
    r->budget = p.sample_count ? p.sample_count : (p.spp_max - p.sample_base);
#define OM_TRY(call)                                 \
    do {                                             \
        om_status s_ = (call);                       \
        if (s_ != OM_OK) { std::string keep = g_err; om_render_destroy(r); g_err = keep; return s_; } \
    } while (0)
    auto body = [&]() -> om_status {
        CU(cudaSetDevice(dev));
        om_status s2 = device_props(dev, r->props);
        if (s2 != OM_OK) return s2;
        s2 = scene_on_device(r->scene, dev, r->ds);
        if (s2 != OM_OK) return s2;
        CU(cudaStreamCreateWithFlags(&r->own_stream, cudaStreamNonBlocking));
        r->stream = r->own_stream;
        CU(dev_alloc(&r->acc, size_t(r->npx) * sizeof(float4)));
        CU(dev_alloc(&r->aux, size_t(r->npx) * sizeof(uint4)));
        CU(cudaMalloc(&r->jitter, p.spp_max));
        CU(cudaMalloc(&r->counters, omk::ctr::CTR_COUNT * sizeof(unsigned long long)));
        CU(cudaMemsetAsync(r->acc, 0, size_t(r->npx) * sizeof(float4), r->stream));
        CU(cudaMemsetAsync(r->aux, 0, size_t(r->npx) * sizeof(uint4), r->stream));
        CU(cudaMemsetAsync(r->counters, 0, omk::ctr::CTR_COUNT * sizeof(unsigned long long), r->stream));
        std::vector<uint8_t> jt;
        omh::jitter_table(p.spp_max, p.seed, jt);
        CU(cudaMemcpyAsync(r->jitter, jt.data(), p.spp_max, cudaMemcpyHostToDevice, r->stream));
        CU(cudaStreamSynchronize(r->stream));
        if (p.flags & OM_RENDER_ADAPTIVE) {
            CU(dev_alloc(&r->live_list, size_t(r->npx) * sizeof(uint32_t)));
            CU(cudaMalloc(&r->live_count, sizeof(uint32_t)));
        }
        // launch shape: persistent grid of one CTA per SM, hot image in shared memory when it fits
        const bool exact = (p.flags & OM_RENDER_FAST_MATH) == 0, count = (p.flags & OM_RENDER_COUNTERS) != 0;
        const size_t hot_bytes = size_t(r->scene->fz.hdr.hot_units) * 16u;
        // (per-ray box padding lives in the global-memory variants only, see node_slab)
        const bool may_stage = !(p.flags & OM_RENDER_NO_SMEM_SCENE) && r->scene->fz.hdr.dyn_pad == 0.f;
        // a staged scene brings the traversal stacks into shared memory with it (TravStack<true, ..>: 4-byte entries, 14-bit node index)
        const uint32_t stack_need = r->scene->fz.hdr.stack_need;
        const bool small_tree = r->scene->fz.hdr.n_nodes <= 16384u;
        r->smem_bytes = hot_bytes + om_exact::render_stack_bytes(stack_need);  // each launch wrapper uses these only when its kernel keeps the scene in shared memory
        r->smem_hook_bytes = hot_bytes + om_exact::hook_stack_bytes(stack_need);
        r->use_smem = may_stage && small_tree && r->smem_bytes + 2048u <= r->props.smem_optin;
        r->use_smem_hook = may_stage && small_tree && r->smem_hook_bytes + 1024u <= r->props.smem_optin / 2;
        int occ = exact ? om_exact::render_occupancy(r->use_smem, count, r->smem_bytes) : om_fast::render_occupancy(r->use_smem, count, r->smem_bytes);
        if (occ < 1) {
            cudaError_t e = cudaGetLastError();
            return fail(OM_ERR_CUDA, "render kernel cannot be resident on device %d (%s)", dev, cudaGetErrorString(e));
        }
        r->grid = occ * r->props.sms;
        if (p.flags & OM_RENDER_POOL) {
            if (p.flags & OM_RENDER_WAVEFRONT) return fail(OM_ERR_INVALID, "om_render_create: OM_RENDER_POOL and OM_RENDER_WAVEFRONT exclude each other");
            r->pool_smem_bytes = hot_bytes + (exact ? om_exact::pool_extra_smem(stack_need) : om_fast::pool_extra_smem(stack_need));
            const bool fits = may_stage && small_tree && p.max_depth < 65536u && r->pool_smem_bytes + 1024u <= r->props.smem_optin;
            const int o3 = fits ? (exact ? om_exact::pool_occupancy(count, r->pool_smem_bytes) : om_fast::pool_occupancy(count, r->pool_smem_bytes)) : 0;
            if (o3 < 1) return fail(OM_ERR_INVALID, "om_render_create: OM_RENDER_POOL needs the scene image, the stacks and the path pool in shared memory (%zu bytes wanted, %zu available)",
                                    r->pool_smem_bytes, r->props.smem_optin);
            r->use_pool = true;
            r->pool_grid = o3 * r->props.sms;
            r->p.flags |= OM_RENDER_PERSISTENT;  // (not the global wavefront; the automatic choices below do not apply)
        }
        if (!(p.flags & (OM_RENDER_WAVEFRONT | OM_RENDER_PERSISTENT))) {  // automatic choice (see OM_RENDER_PERSISTENT)
            const bool fits = hot_bytes + om_exact::wf_stage_bytes() + 1024u <= r->props.smem_optin / 2;
            // many marched objects without a distance BVH: long SDF loops stall the persistent kernel's warps
            const bool slow_march = r->scene->fz.hdr.n_marched > 8u && r->scene->fz.hdr.n_mbvh_nodes == 0u;
            if (!fits || slow_march) r->p.flags |= OM_RENDER_WAVEFRONT;
        }
        if (!(p.flags & (OM_RENDER_PHASE_ALIGN | OM_RENDER_NO_PHASE_ALIGN))) {  // automatic choice (see OM_RENDER_PHASE_ALIGN)
            const omd::SceneHeader& sh = r->scene->fz.hdr;
            const bool accel = (sh.mgrid_dim[0] | sh.n_mbvh_nodes) != 0u;
            if (sh.n_marched >= 1u && sh.n_marched <= 8u && !accel && sh.n[omd::PT_MDYN] == 0u) r->p.flags |= OM_RENDER_PHASE_ALIGN;
        }
        if (r->p.flags & OM_RENDER_NO_PHASE_ALIGN) r->p.flags &= ~(uint32_t)OM_RENDER_PHASE_ALIGN;
        if (r->p.flags & OM_RENDER_WAVEFRONT) {
            // slot pool: resident in L2 (72 B of path state per slot), never more slots than pixels
            uint32_t want = 1u << 20;
            if (const char* e = getenv("OM_WF_SLOTS")) { const long v = atol(e); if (v >= 1024 && v <= (1l << 24)) want = (uint32_t)v; }  // tuning knob
            const uint32_t n_slots = std::min<uint32_t>(want, (r->npx + 31u) & ~31u);
            omk::WfArgs& w = r->wf;
            CU(dev_alloc(&w.s_ray, size_t(n_slots) * 32)); CU(dev_alloc(&w.s_st, size_t(n_slots) * 64)); CU(dev_alloc(&w.s_live, (size_t(n_slots) + 511u) & ~size_t(511)));
            CU(dev_alloc(&w.q_ray[0], size_t(n_slots) * 4)); CU(dev_alloc(&w.q_ray[1], size_t(n_slots) * 4)); CU(dev_alloc(&w.q_march[0], size_t(n_slots) * 4)); CU(dev_alloc(&w.q_march[1], size_t(n_slots) * 4));
            for (int i = 0; i < 2; i++) CU(dev_alloc(&w.q_shade[i], size_t(n_slots) * 4 * omk::wf::N_CLS));
            CU(cudaMalloc(&w.ctr, omk::wf::G_WORDS * 4));
            w.n_slots = n_slots;
            CU(cudaMallocHost(&r->wf_host_ctr, 2 * omk::wf::G_WORDS * 4));
            for (int i = 0; i < 2; i++) CU(cudaEventCreateWithFlags(&r->wf_ev[i], cudaEventDisableTiming));
            r->wf_smem = may_stage && hot_bytes + om_exact::wf_stage_bytes() + 1024u <= r->props.smem_optin / 2;
            int o2 = exact ? om_exact::wf_extend_occupancy(r->wf_smem, count, r->wf_smem ? hot_bytes : 0) : om_fast::wf_extend_occupancy(r->wf_smem, count, r->wf_smem ? hot_bytes : 0);
            if (o2 < 1) return fail(OM_ERR_CUDA, "wavefront extend kernel cannot be resident on device %d", dev);
            r->wf_grid_extend = o2 * r->props.sms;
            r->wf_grid_small = 8 * r->props.sms;
        }
        return OM_OK;
    };
    OM_TRY(body());
#undef OM_TRY
    *out = r;
    return OM_OK;
}

om_status om_render_set_stream(om_render* r, void* cuda_stream) {
    if (!r) return fail(OM_ERR_INVALID, "om_render_set_stream: NULL handle");
    CU(cudaSetDevice(r->dev));
    CU(cudaStreamSynchronize(r->stream));
    om_status s = drain_events(r);
    if (s != OM_OK) return s;
    r->stream = cuda_stream ? (cudaStream_t)cuda_stream : r->own_stream;
    return OM_OK;
}

om_status om_render_reset(om_render* r) {
    if (!r) return fail(OM_ERR_INVALID, "om_render_reset: NULL handle");
    CU(cudaSetDevice(r->dev));
    CU(cudaMemsetAsync(r->acc, 0, size_t(r->npx) * sizeof(float4), r->stream));
    CU(cudaMemsetAsync(r->aux, 0, size_t(r->npx) * sizeof(uint4), r->stream));
    CU(cudaMemsetAsync(r->counters, 0, omk::ctr::CTR_COUNT * sizeof(unsigned long long), r->stream));
    r->have_live = false;
    r->n_uniform = 0;
    return OM_OK;
}

namespace {
// The wavefront loop: iterations are launched in batches; after each batch the counters are copied to
// pinned memory and the host decides, one batch behind the GPU, whether live rays remain.  When few
// are left the persistent kernel finishes the pixels (it restarts their unfinished sample from the
// Philox counters, so nothing is lost or counted twice).
om_status run_wavefront(om_render* r, omk::RenderArgs& a, uint32_t n_target) {
    const bool exact = (r->p.flags & OM_RENDER_FAST_MATH) == 0, count = (r->p.flags & OM_RENDER_COUNTERS) != 0;
    omk::WfArgs w = r->wf;
    w.n_target = n_target;
    const size_t smem = r->wf_smem ? size_t(r->scene->fz.hdr.hot_units) * 16u : 0;  // the extend kernel stages the scene image only (its stacks are in local memory)
    CU(cudaMemsetAsync(w.ctr, 0, omk::wf::G_WORDS * 4, r->stream));
    CU(cudaMemsetAsync(w.s_live, 0, (size_t(w.n_slots) + 511u) & ~size_t(511), r->stream));
    w.it = 0; w.init = 1;
    cudaError_t e = exact ? om_exact::launch_wf_shade(a, w, count, r->wf_grid_small, r->stream) : om_fast::launch_wf_shade(a, w, count, r->wf_grid_small, r->stream);
    if (e == cudaSuccess) e = om_exact::launch_wf_compact(w, r->wf_grid_small, r->stream);
    if (e != cudaSuccess) return fail(OM_ERR_CUDA, "k_wf_shade(init) launch failed: %s", cudaGetErrorString(e));
    r->kernels += 2;
    w.init = 0;
    const uint32_t BATCH = 16, TAIL = 24576;
    // OM_WF_TIMING=1 (development): device time of each kernel of iterations 8..39, printed to stderr
    const bool timing = getenv("OM_WF_TIMING") != nullptr;
    const uint32_t TIMED = 32;
    std::vector<cudaEvent_t> tev;
    if (timing) { tev.resize(4 * TIMED); for (auto& ev : tev) CU(cudaEventCreate(&ev)); }
    // The kernels use the iteration index only as it % 3 (counter set) and it & 1 (queue parity), so a run of iterations whose length
    // is a multiple of 6 has the same arguments wherever it starts: GRAPH_ITERS iterations are captured ONCE per call into a CUDA graph
    // and replayed - 144 kernels per launch instead of one (C5: 6339 launches per frame -> 45).  Iterations past the end of the frame
    // find empty queues and return at once.  OM_WF_NO_GRAPH=1 (or the timing mode) keeps the kernel-by-kernel loop.
    const uint32_t GRAPH_ITERS = 48;
    struct GraphOwner {  // destroyed on every way out of this function, after the work it has in flight
        cudaGraphExec_t exec = nullptr;
        cudaStream_t stream = nullptr;
        ~GraphOwner() { if (exec) { cudaStreamSynchronize(stream); cudaGraphExecDestroy(exec); } }
    } graph_owner;
    graph_owner.stream = r->stream;
    cudaGraphExec_t& graph_exec = graph_owner.exec;
    if (!timing && !getenv("OM_WF_NO_GRAPH")) {
        cudaGraph_t graph = nullptr;
        e = cudaStreamBeginCapture(r->stream, cudaStreamCaptureModeThreadLocal);
        if (e == cudaSuccess) {
            for (uint32_t k = 0; k < GRAPH_ITERS && e == cudaSuccess; k++) {
                w.it = k;
                e = exact ? om_exact::launch_wf_extend(a, w, r->wf_smem, count, r->wf_grid_extend, smem, r->stream)
                          : om_fast::launch_wf_extend(a, w, r->wf_smem, count, r->wf_grid_extend, smem, r->stream);
                if (e == cudaSuccess) e = exact ? om_exact::launch_wf_shade(a, w, count, r->wf_grid_small, r->stream) : om_fast::launch_wf_shade(a, w, count, r->wf_grid_small, r->stream);
                if (e == cudaSuccess) e = om_exact::launch_wf_compact(w, r->wf_grid_small, r->stream);
            }
            const cudaError_t e2 = cudaStreamEndCapture(r->stream, &graph);
            if (e == cudaSuccess) e = e2;
        }
        if (e == cudaSuccess) e = cudaGraphInstantiate(&graph_exec, graph, 0);
        if (graph) cudaGraphDestroy(graph);
        if (e != cudaSuccess) {  // (an unsupported capture is not an error of the render: fall back to plain launches)
            cudaGetLastError();
            graph_exec = nullptr;
        }
    }
    uint32_t it = 0;
    bool pending[2] = {false, false};
    uint32_t it_at[2] = {0, 0};
    for (uint32_t batch = 0;; batch++) {
        const int slot = batch & 1;
        if (pending[slot]) {  // result of batch - 2
            CU(cudaEventSynchronize(r->wf_ev[slot]));
            pending[slot] = false;
            const uint32_t* h = r->wf_host_ctr + slot * omk::wf::G_WORDS;
            const uint32_t* hs = h + (it_at[slot] % 3u) * omk::wf::C_SET_WORDS;
            uint32_t live = hs[omk::wf::C_N_RAY] + hs[omk::wf::C_N_MARCH];
            for (int c = 0; c < omk::wf::N_CLS; c++) live += hs[omk::wf::C_N_SHADE + omk::wf::C_STRIDE * c];
            if (live < TAIL) break;
        }
        if (graph_exec) {  // one launch = GRAPH_ITERS iterations (their arguments only depend on it % 6, see above)
            e = cudaGraphLaunch(graph_exec, r->stream);
            if (e != cudaSuccess) return fail(OM_ERR_CUDA, "wavefront graph launch failed: %s", cudaGetErrorString(e));
            it += GRAPH_ITERS;
            r->kernels += 3 * GRAPH_ITERS;
            r->graph_launches++;
        } else
        for (uint32_t k = 0; k < BATCH; k++, it++) {
            w.it = it;
            if (timing && it >= 8 && it < 8 + TIMED) CU(cudaEventRecord(tev[4 * (it - 8)], r->stream));
            e = exact ? om_exact::launch_wf_extend(a, w, r->wf_smem, count, r->wf_grid_extend, smem, r->stream)
                      : om_fast::launch_wf_extend(a, w, r->wf_smem, count, r->wf_grid_extend, smem, r->stream);
            if (timing && it >= 8 && it < 8 + TIMED) CU(cudaEventRecord(tev[4 * (it - 8) + 1], r->stream));
            if (e == cudaSuccess) e = exact ? om_exact::launch_wf_shade(a, w, count, r->wf_grid_small, r->stream) : om_fast::launch_wf_shade(a, w, count, r->wf_grid_small, r->stream);
            if (timing && it >= 8 && it < 8 + TIMED) CU(cudaEventRecord(tev[4 * (it - 8) + 2], r->stream));
            if (e == cudaSuccess) e = om_exact::launch_wf_compact(w, r->wf_grid_small, r->stream);
            if (timing && it >= 8 && it < 8 + TIMED) CU(cudaEventRecord(tev[4 * (it - 8) + 3], r->stream));
            if (e != cudaSuccess) return fail(OM_ERR_CUDA, "wavefront launch failed: %s", cudaGetErrorString(e));
            r->kernels += 3;
        }
        // snapshot: the rays of iteration `it` are counted in set it % 3
        uint32_t* h = r->wf_host_ctr + slot * omk::wf::G_WORDS;
        CU(cudaMemcpyAsync(h, w.ctr, omk::wf::G_WORDS * 4, cudaMemcpyDeviceToHost, r->stream));
        it_at[slot] = it;
        CU(cudaEventRecord(r->wf_ev[slot], r->stream));
        pending[slot] = true;
    }
    for (int i = 0; i < 2; i++)
        if (pending[i]) { CU(cudaEventSynchronize(r->wf_ev[i])); }
    if (graph_exec) { CU(cudaStreamSynchronize(r->stream)); cudaGraphExecDestroy(graph_exec); graph_exec = nullptr; }
    if (timing && it >= 8 + TIMED) {
        CU(cudaStreamSynchronize(r->stream));
        float t_ext = 0, t_sh = 0, t_co = 0, t_gap = 0;
        for (uint32_t i = 0; i < TIMED; i++) {
            float x;
            cudaEventElapsedTime(&x, tev[4 * i], tev[4 * i + 1]); t_ext += x;
            cudaEventElapsedTime(&x, tev[4 * i + 1], tev[4 * i + 2]); t_sh += x;
            cudaEventElapsedTime(&x, tev[4 * i + 2], tev[4 * i + 3]); t_co += x;
            if (i + 1 < TIMED) { cudaEventElapsedTime(&x, tev[4 * i + 3], tev[4 * i + 4]); t_gap += x; }
        }
        fprintf(stderr, "[om wavefront] per iteration (us): extend %.1f  shade %.1f  compact %.1f  gap %.1f  (slots %u)\n", 1e3f * t_ext / TIMED, 1e3f * t_sh / TIMED,
                1e3f * t_co / TIMED, 1e3f * t_gap / (TIMED - 1), w.n_slots);
    }
    for (auto& ev : tev) cudaEventDestroy(ev);
    // tail: the persistent kernel brings every pixel to the target
    a.n_target = n_target;
    e = exact ? om_exact::launch_render(a, r->use_smem, count, r->grid, r->smem_bytes, r->stream) : om_fast::launch_render(a, r->use_smem, count, r->grid, r->smem_bytes, r->stream);
    if (e != cudaSuccess) return fail(OM_ERR_CUDA, "k_render (tail) launch failed: %s", cudaGetErrorString(e));
    r->kernels++;
    return OM_OK;
}
}  // namespace

namespace {
// Tile scheduling.  A warp owns a tile for all of its samples and tile costs are heavy-tailed (C2: the tiles on the marched torus
// take several times the average), so the launch ends when the last expensive tile does: measured with time stamps, the average
// warp was out of tickets at 144 ms of a 154 ms launch in row order and at 147 of 184 ms when the expensive band came last.  A
// fixed-spp launch of 32 samples or more on the persistent kernel therefore starts with a PILOT launch of 1/64 of the samples in
// which k_render files every tile's cost; the blocks are then handed out most expensive first for the rest of this launch and for
// later launches of the handle.  Same pixels in any order (the passes of a handle are bit-identical to one launch).
bool can_order_blocks(const om_render* r) {
    if (r->p.flags & (OM_RENDER_WAVEFRONT | OM_RENDER_ADAPTIVE)) return false;
    if (r->use_pool) return false;  // the pool hands out pixels one by one: nothing to order
    if (r->p.part_world > 1u || r->owners || getenv("OM_NO_BLOCK_ORDER")) return false;
    const uint32_t tiles_x = (r->p.width + 7u) / 8u, tiles_y = (r->p.height + 3u) / 4u;
    return ((tiles_x + 1u) / 2u) * ((tiles_y + 3u) / 4u) >= 8u * (uint32_t)r->props.sms;
}
om_status refresh_block_order(om_render* r) {
    const uint32_t tiles_x = (r->p.width + 7u) / 8u, tiles_y = (r->p.height + 3u) / 4u, nb = ((tiles_x + 1u) / 2u) * ((tiles_y + 3u) / 4u);
    std::vector<uint32_t> cost(size_t(nb) * 8u);
    CU(cudaMemcpyAsync(cost.data(), r->tile_cost, cost.size() * 4u, cudaMemcpyDeviceToHost, r->stream));
    CU(cudaStreamSynchronize(r->stream));
    std::vector<uint64_t> key(nb);
    for (uint32_t b = 0; b < nb; b++) {
        uint64_t c = 0;
        for (uint32_t i = 0; i < 8u; i++) c = std::max<uint64_t>(c, cost[size_t(b) * 8u + i]);  // a block is a run of 8 tiles for 8 warps: its longest tile
        key[b] = c;
    }
    std::vector<uint32_t> order(nb);
    for (uint32_t b = 0; b < nb; b++) order[b] = b;
    std::stable_sort(order.begin(), order.end(), [&](uint32_t x, uint32_t y) { return key[x] > key[y]; });
    if (!r->block_order) CU(cudaMalloc(&r->block_order, size_t(nb) * 4u));
    CU(cudaMemcpyAsync(r->block_order, order.data(), size_t(nb) * 4u, cudaMemcpyHostToDevice, r->stream));
    CU(cudaStreamSynchronize(r->stream));
    r->have_order = true;
    return OM_OK;
}
om_status launch_pass(om_render* r, uint32_t spp_this_pass, bool pilot);
}  // namespace

om_status om_render_launch(om_render* r, uint32_t spp_this_pass) {
    if (!r) return fail(OM_ERR_INVALID, "om_render_launch: NULL handle");
    CU(cudaSetDevice(r->dev));
    const uint32_t want = spp_this_pass ? spp_this_pass : r->budget;
    const uint32_t todo = (uint32_t)std::min<uint64_t>(r->budget, uint64_t(r->n_uniform) + want) - r->n_uniform;
    if (!r->have_order && todo >= 32u && can_order_blocks(r)) {
        const uint32_t pilot = std::max(1u, todo / 64u);
        const uint32_t tiles_x = (r->p.width + 7u) / 8u, tiles_y = (r->p.height + 3u) / 4u, nb = ((tiles_x + 1u) / 2u) * ((tiles_y + 3u) / 4u);
        if (!r->tile_cost) CU(cudaMalloc(&r->tile_cost, size_t(nb) * 8u * 4u));
        CU(cudaMemsetAsync(r->tile_cost, 0, size_t(nb) * 8u * 4u, r->stream));
        om_status s = launch_pass(r, pilot, true);
        if (s != OM_OK) return s;
        s = refresh_block_order(r);
        if (s != OM_OK) return s;
        return launch_pass(r, todo - pilot, false);
    }
    return launch_pass(r, spp_this_pass, false);
}

namespace {
// one CTA of k_compact_live per 16x16 block of pixels
int compact_blocks(const om_render* r) { const uint32_t tx = (r->p.width + 7u) / 8u, ty = (r->p.height + 3u) / 4u; return int(((tx + 1u) / 2u) * ((ty + 3u) / 4u)); }
#ifndef OM_SPREAD_MAX_LIVE_DEFAULT
#define OM_SPREAD_MAX_LIVE_DEFAULT 250000u  // live pixels at or below which an adaptive pass runs on k_render_spread (measured, profiles/r2_experiments.md section 9)
#endif
om_status launch_pass(om_render* r, uint32_t spp_this_pass, bool pilot) {
    omk::RenderArgs a = make_args(r);
    if (pilot) a.tile_cost = r->tile_cost;
    a.spp_pass = spp_this_pass ? spp_this_pass : r->budget;
    const bool adaptive = (r->p.flags & OM_RENDER_ADAPTIVE) != 0;
    if (r->owners) {  // scheduled ownership: the pixels to walk are decided now, from the state and the table of this moment
        if (!r->live_list) { CU(dev_alloc(&r->live_list, size_t(r->npx) * sizeof(uint32_t))); CU(cudaMalloc(&r->live_count, sizeof(uint32_t))); }
        CU(cudaMemsetAsync(r->live_count, 0, sizeof(uint32_t), r->stream));
        k_compact_live<<<compact_blocks(r), UT, 0, r->stream>>>(r->acc, r->aux, r->p.width, r->p.height, r->budget, adaptive ? 1u : 0u, 0u, 1u, r->owners, r->owner_id, r->live_list, r->live_count);
        CU(cudaGetLastError());
        r->kernels++;
        r->have_live = true;
    }
    if ((adaptive || r->owners) && r->have_live) { a.live_list = r->live_list; a.live_count = r->live_count; }
    EventPair ev;
    if (!r->pool.empty()) { ev = r->pool.back(); r->pool.pop_back(); }
    else { CU(cudaEventCreate(&ev.a)); CU(cudaEventCreate(&ev.b)); }
    CU(cudaMemsetAsync(r->counters + omk::ctr::CTR_TICKET, 0, sizeof(unsigned long long), r->stream));
    CU(cudaEventRecord(ev.a, r->stream));
    const bool exact = (r->p.flags & OM_RENDER_FAST_MATH) == 0, count = (r->p.flags & OM_RENDER_COUNTERS) != 0;
    const uint32_t n_target = (uint32_t)std::min<uint64_t>(r->budget, uint64_t(r->n_uniform) + a.spp_pass);
    if (r->p.flags & OM_RENDER_WAVEFRONT) {
        om_status s = run_wavefront(r, a, n_target);
        if (s != OM_OK) { r->pool.push_back(ev); return s; }
    } else {
        cudaError_t e;
        if (r->use_pool) e = exact ? om_exact::launch_render_pool(a, count, r->pool_grid, r->pool_smem_bytes, r->stream) : om_fast::launch_render_pool(a, count, r->pool_grid, r->pool_smem_bytes, r->stream);
        else {
            // Adaptive passes over a live list: when few pixels are left a pass is as long as its hardest pixel's chain of samples; then
            // k_render_spread lets several lanes share a pixel.  Which of the two kernels renders the pass is decided ON THE DEVICE from
            // the live count (no host synchronisation): both are launched, the other one returns at once.
            const bool spread = adaptive && a.live_list != nullptr && !count && !pilot && !getenv("OM_NO_SPREAD");
            if (spread) {
                uint32_t sp = 1u;
                while (sp * 2u <= 32u && sp * 2u <= a.spp_pass) sp *= 2u;
                const char* mx = getenv("OM_SPREAD_MAX_LIVE");
                a.spread = sp;
                a.spread_max = sp >= 2u ? (mx ? (uint32_t)strtoul(mx, nullptr, 10) : OM_SPREAD_MAX_LIVE_DEFAULT) : 0u;
            }
            // adaptive passes: a slot per lane (k_render_lanes) - a warp that owns a tile idles behind the pixel that retires last
            // (passes of up to 64 samples: +6 % at 16, +3 % at 32, +1 % at 64; one 256-sample launch is bound by its hardest pixel's chain either way)
            const bool lanes = adaptive && !count && !pilot && a.spp_pass <= 64u && !getenv("OM_NO_LANES");
            if (lanes) e = exact ? om_exact::launch_render_lanes(a, r->use_smem, r->grid, r->smem_bytes, r->stream) : om_fast::launch_render_lanes(a, r->use_smem, r->grid, r->smem_bytes, r->stream);
            else e = exact ? om_exact::launch_render(a, r->use_smem, count, r->grid, r->smem_bytes, r->stream)
                           : om_fast::launch_render(a, r->use_smem, count, r->grid, r->smem_bytes, r->stream);
            if (e == cudaSuccess && a.spread_max != 0u) {
                e = exact ? om_exact::launch_render_spread(a, r->use_smem, r->grid, r->smem_bytes, r->stream) : om_fast::launch_render_spread(a, r->use_smem, r->grid, r->smem_bytes, r->stream);
                r->kernels++;
            }
        }
        if (e != cudaSuccess) { r->pool.push_back(ev); return fail(OM_ERR_CUDA, "k_render launch failed: %s", cudaGetErrorString(e)); }
        r->kernels++;
    }
    r->n_uniform = n_target;
    if (adaptive && !r->owners) {  // next pass walks only the pixels that still want samples
        CU(cudaMemsetAsync(r->live_count, 0, sizeof(uint32_t), r->stream));
        k_compact_live<<<compact_blocks(r), UT, 0, r->stream>>>(r->acc, r->aux, r->p.width, r->p.height, r->budget, 1u, r->p.part_rank, r->p.part_world ? r->p.part_world : 1u, nullptr, 0u, r->live_list, r->live_count);
        CU(cudaGetLastError());
        r->kernels++;
        r->have_live = true;
    }
    CU(cudaEventRecord(ev.b, r->stream));
    r->pending.push_back(ev);
    r->launches++;
    return OM_OK;
}
}  // namespace

om_status om_render_sync(om_render* r) {
    if (!r) return fail(OM_ERR_INVALID, "om_render_sync: NULL handle");
    CU(cudaSetDevice(r->dev));
    CU(cudaStreamSynchronize(r->stream));
    return OM_OK;
}
static om_status read_counter(om_render* r, int slot, uint64_t* out) {
    if (!r || !out) return fail(OM_ERR_INVALID, "counter read: NULL argument");
    CU(cudaSetDevice(r->dev));
    unsigned long long v = 0;
    CU(cudaMemcpyAsync(&v, r->counters + slot, sizeof v, cudaMemcpyDeviceToHost, r->stream));
    CU(cudaStreamSynchronize(r->stream));
    *out = v;
    return OM_OK;
}
om_status om_render_samples_done(om_render* r, uint64_t* credited) { return read_counter(r, omk::ctr::CTR_CREDITED, credited); }
om_status om_render_samples_traced(om_render* r, uint64_t* traced) { return read_counter(r, omk::ctr::CTR_TRACED, traced); }
om_status om_render_perf_counters(om_render* r, uint64_t out[16]) {
    if (!r || !out) return fail(OM_ERR_INVALID, "om_render_perf_counters: NULL argument");
    CU(cudaSetDevice(r->dev));
    CU(cudaMemcpyAsync(out, r->counters, 16 * sizeof(uint64_t), cudaMemcpyDeviceToHost, r->stream));
    CU(cudaStreamSynchronize(r->stream));
    out[10] = r->kernels;
    out[11] = r->p.flags;  // with the pipeline that was chosen when the caller left it open
    out[12] = (r->p.flags & OM_RENDER_WAVEFRONT) ? (r->wf_smem ? 1 : 0) : (r->use_smem ? 1 : 0);
    out[15] = r->graph_launches;
    return OM_OK;
}
om_status om_render_elapsed_ms(om_render* r, float* ms, uint32_t* launches) {
    if (!r) return fail(OM_ERR_INVALID, "om_render_elapsed_ms: NULL handle");
    CU(cudaSetDevice(r->dev));
    om_status s = drain_events(r);
    if (s != OM_OK) return s;
    if (ms) *ms = r->acc_ms;
    if (launches) *launches = r->launches;
    r->acc_ms = 0.f;
    r->launches = 0;
    return OM_OK;
}

om_status om_render_read_pixels(om_render* r, om_pixel* host, size_t n) {
    if (!r || !host) return fail(OM_ERR_INVALID, "om_render_read_pixels: NULL argument");
    if (n != r->npx) return fail(OM_ERR_INVALID, "om_render_read_pixels: n must be width*height");
    static_assert(sizeof(om_pixel) == 40, "om_pixel layout");
    CU(cudaSetDevice(r->dev));
    om_status s = ensure_staging(r, n * sizeof(om_pixel));
    if (s != OM_OK) return s;
    k_pack_pixels<<<ublocks(n), UT, 0, r->stream>>>(r->acc, r->aux, (uint32_t*)r->staging, r->npx);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(host, r->staging, n * sizeof(om_pixel), cudaMemcpyDeviceToHost, r->stream));
    CU(cudaStreamSynchronize(r->stream));
    return OM_OK;
}
om_status om_render_read_rgb(om_render* r, uint8_t* host_rgb, size_t nbytes) {
    if (!r || !host_rgb) return fail(OM_ERR_INVALID, "om_render_read_rgb: NULL argument");
    if (nbytes != size_t(r->npx) * 3u) return fail(OM_ERR_INVALID, "om_render_read_rgb: nbytes must be width*height*3");
    CU(cudaSetDevice(r->dev));
    om_status s = ensure_staging(r, nbytes);
    if (s != OM_OK) return s;
    k_rgb<<<ublocks(r->npx), UT, 0, r->stream>>>(r->acc, (uint8_t*)r->staging, r->npx);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(host_rgb, r->staging, nbytes, cudaMemcpyDeviceToHost, r->stream));
    CU(cudaStreamSynchronize(r->stream));
    return OM_OK;
}

om_status om_render_read_view(om_render* r, uint32_t mode, uint8_t* host_rgb, size_t nbytes) {
    if (!r || !host_rgb) return fail(OM_ERR_INVALID, "om_render_read_view: NULL argument");
    if (mode > 6u) return fail(OM_ERR_INVALID, "om_render_read_view: mode must be 0..6 (main.rs:361-368)");
    if (nbytes != size_t(r->npx) * 3u) return fail(OM_ERR_INVALID, "om_render_read_view: nbytes must be width*height*3");
    if ((mode == 2u || mode == 4u || mode == 6u) && (r->p.width < 3u || r->p.height < 3u))
        return fail(OM_ERR_INVALID, "om_render_read_view: the 3x3 filters need an image of at least 3x3 (the reference's border loops, main.rs:324-342)");
    CU(cudaSetDevice(r->dev));
    om_status s = ensure_staging(r, nbytes + 16u);
    if (s != OM_OK) return s;
    uint32_t* red = (uint32_t*)((uint8_t*)r->staging + ((nbytes + 3u) & ~size_t(3)));
    const uint32_t init[3] = {1u, 0u, 0u};
    CU(cudaMemcpyAsync(red, init, sizeof init, cudaMemcpyHostToDevice, r->stream));
    if (mode == 1u || mode == 3u) {
        k_view_reduce<<<ublocks(r->npx), UT, 0, r->stream>>>(r->acc, r->aux, r->npx, red);
        CU(cudaGetLastError());
    }
    k_view<<<ublocks(r->npx), UT, 0, r->stream>>>(r->acc, r->aux, (uint8_t*)r->staging, r->p.width, r->p.height, mode, red);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(host_rgb, r->staging, nbytes, cudaMemcpyDeviceToHost, r->stream));
    CU(cudaStreamSynchronize(r->stream));
    return OM_OK;
}

// ---- om_render_read_view_noalloc ---------------------------------------------------------------------------------
// The same refresh as om_render_read_view for callers that may not allocate (CUDA graph capture, tight serving loops):
// the device-side RGB image and the three reduction words live in one file-scope device array instead of the handle's
// staging block.  No counterpart in the reference.
// Layout: bytes [0, 3 * npx) = the RGB24 image, the reduction words at OM_VIEW_NOALLOC_RED_OFFSET.  Every GPU that loads
// the library holds the array (6 MiB + 256 B) until the process exits.  All calls on a device share it and it keeps the
// last call's contents: the reduction words are re-initialised by k_view_red_init at the start of every call, the image
// bytes are overwritten in full before they are read.  Calls that overlap on the GPU (two handles on different streams
// of one device) race on it.
constexpr size_t OM_VIEW_NOALLOC_RED_OFFSET = size_t(OM_VIEW_NOALLOC_MAX_PIXELS) * 3u;
static __device__ __align__(256) unsigned char om_view_scratch[OM_VIEW_NOALLOC_RED_OFFSET + 256u];
// device code may name the array directly
__global__ void k_view_red_init() {
    uint32_t* red = reinterpret_cast<uint32_t*>(om_view_scratch + OM_VIEW_NOALLOC_RED_OFFSET);
    if (threadIdx.x == 0u) { red[0] = 1u; red[1] = 0u; red[2] = 0u; }  // max n starts at 1, no finite depth seen yet
}
namespace {
// Host code needs the array's DEVICE address: in host code `om_view_scratch` is only a stand-in for the symbol.
// Looked up once per device, with that device current.
constexpr int OM_MAX_DEVICES = 64;
std::mutex g_view_scratch_mu;
unsigned char* g_view_scratch_addr[OM_MAX_DEVICES] = {nullptr};
om_status view_scratch_on_device(int dev, unsigned char** out) {
    if (dev < 0 || dev >= OM_MAX_DEVICES) return fail(OM_ERR_INVALID, "om_render_read_view_noalloc: device %d out of range", dev);
    std::lock_guard<std::mutex> lk(g_view_scratch_mu);
    if (!g_view_scratch_addr[dev]) {
        CU(cudaSetDevice(dev));
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, om_view_scratch);
        if (e != cudaSuccess || !p) return fail(OM_ERR_CUDA, "om_render_read_view_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
        g_view_scratch_addr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_view_scratch_addr[dev];
    return OM_OK;
}
}  // namespace

om_status om_render_read_view_noalloc(om_render* r, uint32_t mode, uint8_t* host_rgb, size_t nbytes) {
    if (!r || !host_rgb) return fail(OM_ERR_INVALID, "om_render_read_view_noalloc: NULL argument");
    if (mode > 6u) return fail(OM_ERR_INVALID, "om_render_read_view_noalloc: mode must be 0..6 (main.rs:361-368)");
    if (nbytes != size_t(r->npx) * 3u) return fail(OM_ERR_INVALID, "om_render_read_view_noalloc: nbytes must be width*height*3");
    if (r->npx > OM_VIEW_NOALLOC_MAX_PIXELS)
        return fail(OM_ERR_INVALID, "om_render_read_view_noalloc: %u pixels exceed the fixed scratch of OM_VIEW_NOALLOC_MAX_PIXELS = %u (use om_render_read_view)",
                    r->npx, (unsigned)OM_VIEW_NOALLOC_MAX_PIXELS);
    if ((mode == 2u || mode == 4u || mode == 6u) && (r->p.width < 3u || r->p.height < 3u))
        return fail(OM_ERR_INVALID, "om_render_read_view_noalloc: the 3x3 filters need an image of at least 3x3 (the reference's border loops, main.rs:324-342)");
    CU(cudaSetDevice(r->dev));
    unsigned char* scratch = nullptr;
    const om_status s = view_scratch_on_device(r->dev, &scratch);
    if (s != OM_OK) return s;
    uint32_t* red = reinterpret_cast<uint32_t*>(scratch + OM_VIEW_NOALLOC_RED_OFFSET);
    k_view_red_init<<<1, 32, 0, r->stream>>>();
    CU(cudaGetLastError());
    if (mode == 1u || mode == 3u) {
        k_view_reduce<<<ublocks(r->npx), UT, 0, r->stream>>>(r->acc, r->aux, r->npx, red);
        CU(cudaGetLastError());
    }
    k_view<<<ublocks(r->npx), UT, 0, r->stream>>>(r->acc, r->aux, scratch, r->p.width, r->p.height, mode, red);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(host_rgb, scratch, nbytes, cudaMemcpyDeviceToHost, r->stream));
    CU(cudaStreamSynchronize(r->stream));
    return OM_OK;
}

om_status om_write_bmp(const char* path, const uint8_t* rgb, uint32_t width, uint32_t height) {
    if (!path || !rgb || !width || !height) return fail(OM_ERR_INVALID, "om_write_bmp: bad argument");
    FILE* f = fopen(path, "wb");
    if (!f) return fail(OM_ERR_INVALID, "om_write_bmp: cannot open %s", path);
    const uint32_t row = (width * 3u + 3u) & ~3u, data = row * height, off = 54u, size = off + data;
    uint8_t h[54] = {0};
    auto put32 = [&](int o, uint32_t v) { h[o] = uint8_t(v); h[o + 1] = uint8_t(v >> 8); h[o + 2] = uint8_t(v >> 16); h[o + 3] = uint8_t(v >> 24); };
    h[0] = 'B'; h[1] = 'M'; put32(2, size); put32(10, off); put32(14, 40u); put32(18, width); put32(22, height);
    h[26] = 1; h[28] = 24; put32(34, data); put32(38, 2835u); put32(42, 2835u);
    bool ok = fwrite(h, 1, 54, f) == 54;
    std::vector<uint8_t> line(row, 0);
    for (uint32_t y = 0; ok && y < height; y++) {  // BMP rows run bottom-up, pixels are BGR
        const uint8_t* src = rgb + size_t(height - 1 - y) * width * 3u;
        for (uint32_t x = 0; x < width; x++) { line[3 * x] = src[3 * x + 2]; line[3 * x + 1] = src[3 * x + 1]; line[3 * x + 2] = src[3 * x]; }
        ok = fwrite(line.data(), 1, row, f) == row;
    }
    ok = (fclose(f) == 0) && ok;
    return ok ? OM_OK : fail(OM_ERR_INVALID, "om_write_bmp: write to %s failed", path);
}

om_status om_render_primary(om_render* r, uint32_t mode, uint32_t sample_index, uint64_t* ids, float* t, size_t n) {
    if (!r || !ids || !t) return fail(OM_ERR_INVALID, "om_render_primary: NULL argument");
    if (n != r->npx) return fail(OM_ERR_INVALID, "om_render_primary: n must be width*height");
    if (mode > 1 || (mode == 1 && sample_index >= r->p.spp_max)) return fail(OM_ERR_INVALID, "om_render_primary: bad mode / sample index");
    CU(cudaSetDevice(r->dev));
    om_status s = ensure_staging(r, n * 12u);
    if (s != OM_OK) return s;
    unsigned long long* d_ids = (unsigned long long*)r->staging;
    float* d_t = (float*)(d_ids + n);
    omk::RenderArgs a = make_args(r);
    const int grid = std::min(ublocks(n), r->props.sms * 8);
    cudaError_t e = !(r->p.flags & OM_RENDER_FAST_MATH) ? om_exact::launch_primary(a, r->use_smem_hook, grid, r->smem_hook_bytes, r->stream, mode, sample_index, d_ids, d_t)
                                                   : om_fast::launch_primary(a, r->use_smem_hook, grid, r->smem_hook_bytes, r->stream, mode, sample_index, d_ids, d_t);
    if (e != cudaSuccess) return fail(OM_ERR_CUDA, "k_primary launch failed: %s", cudaGetErrorString(e));
    CU(cudaMemcpyAsync(ids, d_ids, n * 8u, cudaMemcpyDeviceToHost, r->stream));
    CU(cudaMemcpyAsync(t, d_t, n * 4u, cudaMemcpyDeviceToHost, r->stream));
    CU(cudaStreamSynchronize(r->stream));
    return OM_OK;
}

om_status om_render_export_accum(om_render* r, float* dev_acc, uint64_t* dev_bloom) {
    if (!r || !dev_acc) return fail(OM_ERR_INVALID, "om_render_export_accum: NULL argument");
    CU(cudaSetDevice(r->dev));
    k_export_accum<<<ublocks(r->npx), UT, 0, r->stream>>>(r->acc, r->aux, (float4*)dev_acc, (unsigned long long*)dev_bloom, r->npx);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(r->stream));
    return OM_OK;
}
om_status om_render_import_accum(om_render* r, const float* dev_acc, const uint64_t* dev_bloom) {
    if (!r || !dev_acc) return fail(OM_ERR_INVALID, "om_render_import_accum: NULL argument");
    CU(cudaSetDevice(r->dev));
    k_import_accum<<<ublocks(r->npx), UT, 0, r->stream>>>(r->acc, r->aux, (const float4*)dev_acc, (const unsigned long long*)dev_bloom, r->npx);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(r->stream));
    return OM_OK;
}
om_status om_render_set_chunk_owners(om_render* r, const uint8_t* owners, size_t n_chunks, uint32_t my_id) {
    if (!r) return fail(OM_ERR_INVALID, "om_render_set_chunk_owners: NULL handle");
    CU(cudaSetDevice(r->dev));
    const size_t chunks = (size_t(r->npx) + OM_CHUNK_PIXELS - 1) / OM_CHUNK_PIXELS;
    if (!owners) {  // back to the static rule of om_render_params
        CU(cudaStreamSynchronize(r->stream));
        cudaFree(r->owners); r->owners = nullptr; r->have_live = false;
        return OM_OK;
    }
    if (n_chunks != chunks) return fail(OM_ERR_INVALID, "om_render_set_chunk_owners: the image has %zu chunks of %u pixels, got %zu", chunks, OM_CHUNK_PIXELS, n_chunks);
    if (r->p.part_world > 1u || r->p.sample_base || r->p.sample_count) return fail(OM_ERR_INVALID, "om_render_set_chunk_owners: the handle must own the whole frame (no static partition)");
    if (!r->owners) CU(cudaMalloc(&r->owners, chunks));
    CU(cudaMemcpyAsync(r->owners, owners, chunks, cudaMemcpyHostToDevice, r->stream));
    CU(cudaStreamSynchronize(r->stream));
    r->owner_id = my_id;
    return OM_OK;
}
om_status om_render_export_state(om_render* r, uint32_t* dev_state, uint32_t only_owned) {
    if (!r || !dev_state) return fail(OM_ERR_INVALID, "om_render_export_state: NULL argument");
    CU(cudaSetDevice(r->dev));
    k_export_state<<<ublocks(r->npx), UT, 0, r->stream>>>(r->acc, r->aux, (uint4*)dev_state, r->npx, only_owned, r->p.part_rank, r->p.part_world ? r->p.part_world : 1u, r->owners, r->owner_id);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(r->stream));
    return OM_OK;
}
om_status om_render_import_state(om_render* r, const uint32_t* dev_state) {
    if (!r || !dev_state) return fail(OM_ERR_INVALID, "om_render_import_state: NULL argument");
    CU(cudaSetDevice(r->dev));
    k_import_state<<<ublocks(r->npx), UT, 0, r->stream>>>(r->acc, r->aux, (const uint4*)dev_state, r->npx);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(r->stream));
    r->have_live = false;
    return OM_OK;
}
om_status om_render_live_per_chunk(om_render* r, uint32_t* host_counts, size_t n_chunks) {
    if (!r || !host_counts) return fail(OM_ERR_INVALID, "om_render_live_per_chunk: NULL argument");
    CU(cudaSetDevice(r->dev));
    const size_t chunks = (size_t(r->npx) + OM_CHUNK_PIXELS - 1) / OM_CHUNK_PIXELS;
    if (n_chunks != chunks) return fail(OM_ERR_INVALID, "om_render_live_per_chunk: the image has %zu chunks, got %zu", chunks, n_chunks);
    if (!r->chunk_counts) CU(cudaMalloc(&r->chunk_counts, chunks * sizeof(uint32_t)));
    CU(cudaMemsetAsync(r->chunk_counts, 0, chunks * sizeof(uint32_t), r->stream));
    k_live_per_chunk<<<ublocks(r->npx), UT, 0, r->stream>>>(r->acc, r->aux, r->npx, r->budget, (r->p.flags & OM_RENDER_ADAPTIVE) ? 1u : 0u, r->chunk_counts);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(host_counts, r->chunk_counts, chunks * sizeof(uint32_t), cudaMemcpyDeviceToHost, r->stream));
    CU(cudaStreamSynchronize(r->stream));
    return OM_OK;
}
om_status om_bloom_or(uint64_t* dev_dst, const uint64_t* dev_src, size_t nsrc, size_t n, int32_t device) {
    if (!dev_dst || (!dev_src && nsrc)) return fail(OM_ERR_INVALID, "om_bloom_or: NULL argument");
    int dev = 0;
    om_status s = pick_device(device, dev);
    if (s != OM_OK) return s;
    CU(cudaSetDevice(dev));
    k_bloom_or<<<ublocks(n), UT>>>((unsigned long long*)dev_dst, (const unsigned long long*)dev_src, (uint32_t)nsrc, n);
    CU(cudaGetLastError());
    CU(cudaDeviceSynchronize());
    return OM_OK;
}

om_status om_render_bloom_or(om_render* r, uint64_t* dev_dst, const uint64_t* dev_src, size_t nsrc) {
    if (!r || !dev_dst || (!dev_src && nsrc)) return fail(OM_ERR_INVALID, "om_render_bloom_or: NULL argument");
    CU(cudaSetDevice(r->dev));
    k_bloom_or<<<ublocks(r->npx), UT, 0, r->stream>>>((unsigned long long*)dev_dst, (const unsigned long long*)dev_src, (uint32_t)nsrc, (size_t)r->npx);
    CU(cudaGetLastError());
    return OM_OK;
}

om_status om_trace_rays(const om_scene* frozen, const float* origins, const float* dirs, size_t n, float tmin, float tmax, uint32_t flags, int32_t device,
                        uint64_t* ids, float* t, float* points, float* normals) {
    if (!frozen || !frozen->frozen) return fail(OM_ERR_INVALID, "om_trace_rays: scene is not frozen");
    if ((!origins || !dirs) && n) return fail(OM_ERR_INVALID, "om_trace_rays: NULL rays");
    if (n > 0x7FFFFFFFull / 3) return fail(OM_ERR_INVALID, "om_trace_rays: too many rays");
    int dev = 0;
    om_status s = pick_device(device, dev);
    if (s != OM_OK) return s;
    if (n == 0) return OM_OK;
    CU(cudaSetDevice(dev));
    DeviceProps props;
    s = device_props(dev, props);
    if (s != OM_OK) return s;
    DevScene ds;
    s = scene_on_device(const_cast<om_scene*>(frozen), dev, ds);
    if (s != OM_OK) return s;
    char* buf = nullptr;
    const size_t b_rays = n * 12u, b_ids = n * 8u, b_t = n * 4u;
    CU(cudaMalloc(&buf, 2 * b_rays + b_ids + b_t + 2 * b_rays));
    omk::TraceArgs a;
    memset(&a, 0, sizeof a);
    a.scene = frozen->fz.hdr; a.hot_g = ds.hot; a.cold_g = ds.cold;
    a.ids = (unsigned long long*)buf;
    float* d_o = (float*)(buf + b_ids);
    float* d_d = d_o + 3 * n;
    a.origins = d_o; a.dirs = d_d;
    a.t = d_d + 3 * n;
    a.points = a.t + n;
    a.normals = a.points + 3 * n;
    a.n = (uint32_t)n; a.tmin = tmin; a.tmax = tmax; a.flags = flags;
    auto done = [&](om_status st) { cudaFree(buf); return st; };
    cudaError_t e = cudaMemcpy(d_o, origins, b_rays, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_d, dirs, b_rays, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return done(fail(OM_ERR_CUDA, "om_trace_rays: upload failed: %s", cudaGetErrorString(e)));
    const size_t hot_bytes = size_t(frozen->fz.hdr.hot_units) * 16u;
    const size_t smem_bytes = hot_bytes + om_exact::hook_stack_bytes(frozen->fz.hdr.stack_need);  // scene image + traversal stacks
    const bool smem = !(flags & OM_RENDER_NO_SMEM_SCENE) && frozen->fz.hdr.dyn_pad == 0.f && frozen->fz.hdr.n_nodes <= 16384u && smem_bytes + 1024u <= props.smem_optin / 2;
    const int grid = std::min(ublocks(n), props.sms * 8);
    e = !(flags & OM_RENDER_FAST_MATH) ? om_exact::launch_trace(a, smem, grid, smem ? smem_bytes : 0, 0) : om_fast::launch_trace(a, smem, grid, smem ? smem_bytes : 0, 0);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return done(fail(OM_ERR_CUDA, "k_trace_rays failed: %s", cudaGetErrorString(e)));
    if (ids) e = cudaMemcpy(ids, a.ids, b_ids, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && t) e = cudaMemcpy(t, a.t, b_t, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && points) e = cudaMemcpy(points, a.points, b_rays, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && normals) e = cudaMemcpy(normals, a.normals, b_rays, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return done(fail(OM_ERR_CUDA, "om_trace_rays: download failed: %s", cudaGetErrorString(e)));
    return done(OM_OK);
}

om_status om_bench_fp32_peak(int32_t device, float* tflops, float* ms_out) {
    int dev = 0;
    om_status s = pick_device(device, dev);
    if (s != OM_OK) return s;
    CU(cudaSetDevice(dev));
    DeviceProps props;
    s = device_props(dev, props);
    if (s != OM_OK) return s;
    float* d = nullptr;
    CU(cudaMalloc(&d, 4));
    cudaEvent_t a, b;
    CU(cudaEventCreate(&a)); CU(cudaEventCreate(&b));
    const int iters = 1 << 16, grid = props.sms * 8, threads = 256;
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
        CU(cudaEventRecord(a));
        k_fma_peak<<<grid, threads>>>(d, iters);
        CU(cudaEventRecord(b));
        CU(cudaEventSynchronize(b));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, a, b));
        if (rep > 0 && ms < best) best = ms;
    }
    cudaEventDestroy(a); cudaEventDestroy(b); cudaFree(d);
    const double flops = 2.0 * 8.0 * double(iters) * double(grid) * double(threads);
    if (tflops) *tflops = float(flops / (double(best) * 1e-3) / 1e12);
    if (ms_out) *ms_out = best;
    return OM_OK;
}

}  // extern "C"
