// CTA-resident wavefront ("path pool") pipeline of the B200 renderer.  Included after om_device.cuh and om_wavefront.cuh
// inside namespace OM_NS (both builds).
//
// The persistent kernel (k_render) ties a path to a lane: at any moment a third of a warp's lanes are ending a sample and
// starting the next one, two thirds are shading a hit, a few are marching, so everything but the traversal runs at ~10 of 32
// lanes (ncu, round 2).  The global wavefront pipeline (om_wavefront.cuh) sorts the work but pays for it with L2 round trips
// and a kernel launch per phase.  This kernel keeps the sorting and drops both costs: every CTA owns a POOL of path slots
// whose state lives in SHARED memory next to the scene image, and alternates two phases separated by __syncthreads:
//   phase E (extend)  the slots that hold a ray are taken 32 at a time, compacted, and traversed (closest_traced_warp: the
//                     same warp-collective BVH walk as k_render); each resolved ray is filed by what it needs next
//   phase S (shade)   the files are taken 32 at a time, longest kind first: marching rays (a bounded number of SDF steps
//                     per phase, the rest next time), hits on other primitives, dielectric / metal / Lambertian spheres,
//                     misses.  A batch holds one kind, so Material::scatter, Stats::add and the next sample's camera ray
//                     run with all 32 lanes.  Paths that go on are queued for the next phase E.
// A slot is bound to one pixel until that pixel has its samples of the launch (Stats::add is order dependent), then takes
// the next unstarted pixel by a global ticket.  Stats live in the framebuffer (read-modify-write per sample, L2).
// Same arithmetic per path as k_render, in the same order: bit-identical output (tests/test_gpu_parity.py).
#pragma once

#ifndef OM_POOL_THREADS
#define OM_POOL_THREADS 896
#endif
#ifndef OM_POOL_MARCH_STEPS
#define OM_POOL_MARCH_STEPS 4
#endif
constexpr int POOL_T = OM_POOL_THREADS;
#ifndef OM_POOL_SLOTS
#define OM_POOL_SLOTS 896
#endif
constexpr uint32_t POOL_N = OM_POOL_SLOTS;  // slots per CTA (>= POOL_T; more slots than threads = several batches per warp and phase)
static_assert(POOL_N >= (uint32_t)POOL_T && POOL_N < 65535u, "slot ids are 16 bits; every thread binds at least one slot");
constexpr uint32_t POOL_FIELDS = 6u;           // float4 per slot
// phase S work kinds in the order they are handed out (longest batches first)
enum : uint32_t { PQ_MARCH = 0, PQ_OTHER = 1, PQ_DIEL = 2, PQ_METAL = 3, PQ_LAMB = 4, PQ_SKY = 5, PQ_COUNT = 6 };
constexpr uint32_t POOL_QUEUES = 2u + 2u * PQ_COUNT;  // ray queue x 2, every phase S file x 2 (phase S reads one parity while marches that end re-file into the other)

struct PoolCtl {
    uint32_t nE[2], tE[2];         // ray queue of each parity: entries, batch ticket
    uint32_t nQ[2][PQ_COUNT];      // phase S files of each parity: entries
    uint32_t tS[2];                // phase S batch ticket
};

// shared-memory bytes behind the scene image: traversal stacks, slot fields, queues
__host__ __device__ inline size_t pool_extra_bytes(uint32_t stack_need) {
    return size_t(POOL_T) * stack_need * 4u + 32u + size_t(POOL_N) * POOL_FIELDS * 16u + size_t(POOL_QUEUES) * POOL_N * 2u;
}

struct PoolPath {
    Ray ray;
    V3 color;
    float first_t, hit_t;
    uint32_t hit_ref, pxl, seg, draw, todo, b0, b1, b2, b3, first_lo, first_hi, n;
    float mt;          // march: parametric distance reached
    uint32_t mleft;    // march: iterations left of the reference's 1024; 0xFFFFFFFF = not started (unstuck first)
};
__device__ __forceinline__ void pool_load(const float4* sl, uint32_t s, PoolPath& P) {
    const float4 f0 = sl[s], f1 = sl[POOL_N + s], f2 = sl[2u * POOL_N + s], f3 = sl[3u * POOL_N + s], f4 = sl[4u * POOL_N + s], f5 = sl[5u * POOL_N + s];
    P.ray.o = v3(f0.x, f0.y, f0.z); P.ray.d = v3(f0.w, f1.x, f1.y); P.color = v3(f1.z, f1.w, f2.x);
    P.first_t = f2.y; P.hit_t = f2.z; P.hit_ref = __float_as_uint(f2.w);
    // seg and the march iteration count share a word (the host only picks this pipeline for max_depth < 65536)
    P.pxl = __float_as_uint(f3.x); P.seg = __float_as_uint(f3.y) & 0xFFFFu; P.draw = __float_as_uint(f3.z); P.todo = __float_as_uint(f3.w);
    { const uint32_t m = __float_as_uint(f3.y) >> 16; P.mleft = m == 0xFFFFu ? 0xFFFFFFFFu : m; }
    P.b0 = __float_as_uint(f4.x); P.b1 = __float_as_uint(f4.y); P.b2 = __float_as_uint(f4.z); P.b3 = __float_as_uint(f4.w);
    P.first_lo = __float_as_uint(f5.x); P.first_hi = __float_as_uint(f5.y); P.n = __float_as_uint(f5.z); P.mt = f5.w;
}
__device__ __forceinline__ void pool_store(float4* sl, uint32_t s, const PoolPath& P) {
    sl[s] = make_float4(P.ray.o.x, P.ray.o.y, P.ray.o.z, P.ray.d.x);
    sl[POOL_N + s] = make_float4(P.ray.d.y, P.ray.d.z, P.color.x, P.color.y);
    sl[2u * POOL_N + s] = make_float4(P.color.z, P.first_t, P.hit_t, __uint_as_float(P.hit_ref));
    sl[3u * POOL_N + s] = make_float4(__uint_as_float(P.pxl), __uint_as_float((P.seg & 0xFFFFu) | ((P.mleft == 0xFFFFFFFFu ? 0xFFFFu : P.mleft) << 16)), __uint_as_float(P.draw), __uint_as_float(P.todo));
    sl[4u * POOL_N + s] = make_float4(__uint_as_float(P.b0), __uint_as_float(P.b1), __uint_as_float(P.b2), __uint_as_float(P.b3));
    sl[5u * POOL_N + s] = make_float4(__uint_as_float(P.first_lo), __uint_as_float(P.first_hi), __uint_as_float(P.n), P.mt);
}

// warp-aggregated append of `s` to the queue `q` counted by `*n` for the lanes with `pred`; all 32 lanes call
__device__ __forceinline__ void pool_push(uint16_t* q, uint32_t* n, bool pred, uint32_t s, uint32_t lane) {
    const unsigned m = __ballot_sync(FULL, pred);
    if (m == 0u) return;
    uint32_t base = 0;
    const uint32_t leader = (uint32_t)__ffs(m) - 1u;
    if (lane == leader) base = atomicAdd(n, (uint32_t)__popc(m));
    base = __shfl_sync(FULL, base, leader);
    if (pred) q[base + (uint32_t)__popc(m & ((1u << lane) - 1u))] = (uint16_t)s;
}

// ============================================================================================================================
// k_render_pool<COUNT, MA>: MA as in k_render (0 = no marched structures / dyn objects, 1 = candidate grid or distance BVH,
// 3 = dyn marched objects).  The scene image is always staged (the host only picks this kernel when everything fits).
// ============================================================================================================================
template <bool COUNT, int MA>
__global__ void __launch_bounds__(POOL_T, 1) k_render_pool(const __grid_constant__ RenderArgs A) {
    stage_scene_smem(A.hot_g, A.scene.hot_units);
    SceneView<true> sv;
    sv.hot_g = A.hot_g;
    sv.cold_g = A.cold_g;
    const SceneHeader& S = A.scene;
    const bool brute = (A.flags & OM_RENDER_BRUTE_FORCE) != 0;
    const bool adaptive = (A.flags & OM_RENDER_ADAPTIVE) != 0;
    const uint32_t nseg = A.max_depth > 1u ? A.max_depth : 1u;
    const uint32_t lane = threadIdx.x & 31u;
    TravStack<true, POOL_T> stack;
    stack.init(S.hot_units);
    // shared memory behind the scene image and the stacks: slot fields, then the queues
    float4* sl = om_smem + S.hot_units + ((uint32_t)POOL_T * S.stack_need + 3u) / 4u + 1u;
    uint16_t* qbase = reinterpret_cast<uint16_t*>(sl + POOL_FIELDS * POOL_N);
    auto q_ray = [&](uint32_t par) { return qbase + par * POOL_N; };
    auto q_kind = [&](uint32_t par, uint32_t k) { return qbase + (2u + par * PQ_COUNT + k) * POOL_N; };
    __shared__ PoolCtl ctl;
    if (threadIdx.x == 0) {
        for (int p = 0; p < 2; p++) { ctl.nE[p] = 0; ctl.tE[p] = 0; ctl.tS[p] = 0; for (uint32_t k = 0; k < PQ_COUNT; k++) ctl.nQ[p][k] = 0; }
    }
    __syncthreads();
    Counters cn = {0, 0, 0, 0, 0, 0, 0};
    unsigned long long credited = 0, traced = 0;

    // a slot takes the next pixel that still wants samples; false = none left (the slot stays empty)
    auto bind_pixel = [&](PoolPath& P, uint32_t k_first, bool have_first) -> bool {
        for (;;) {
            const uint32_t k = have_first ? k_first : (uint32_t)atomicAdd(&A.counters[CTR_TICKET], 1ull);
            have_first = false;
            if (k >= A.n_work) { P.pxl = 0xFFFFFFFFu; P.todo = 0u; return false; }
            const uint32_t pxl = work_to_pixel(A, k);
            if (pxl == 0xFFFFFFFFu) continue;
            const uint32_t n = __float_as_uint(A.acc[pxl].w);
            uint32_t todo = A.spp_budget > n ? min(A.spp_budget - n, A.spp_pass) : 0u;
            if (A.n_target) { const uint32_t tg = min(A.spp_budget, A.n_target); todo = tg > n ? tg - n : 0u; }
            if (adaptive && A.aux[pxl].y >= 5u) todo = 0u;
            if (todo == 0u) continue;
            P.pxl = pxl; P.n = n; P.todo = todo;
            return true;
        }
    };
    auto new_sample = [&](PoolPath& P) {  // render_thread.rs:183-192
        Rng g;
        const uint32_t sample = A.sample_base + P.n;
        rng_init(g, A.rk, P.pxl, sample);
        g.b0 = g.b1 = g.b2 = g.b3 = 0u;
        P.ray = primary_ray(A, P.pxl, sample, g);
        P.color = v3(1.f, 1.f, 1.f);
        P.seg = 0u; P.first_t = kInf; P.first_lo = 0u; P.first_hi = 0u;
        P.draw = g.draw; P.b0 = g.b0; P.b1 = g.b1; P.b2 = g.b2; P.b3 = g.b3;
        P.mleft = 0xFFFFFFFFu; P.mt = 0.f;
    };

    // ---- every slot takes its first pixel: a warp's 32 slots get one 8x4 tile (coherent camera rays) ----
    for (uint32_t s = threadIdx.x; s < ((POOL_N + 31u) & ~31u); s += POOL_T) {  // (warp-uniform trip count)
        uint32_t k0 = 0;
        if (lane == 0u) k0 = (uint32_t)atomicAdd(&A.counters[CTR_TICKET], 32ull);
        k0 = __shfl_sync(FULL, k0, 0) + lane;
        PoolPath P;
        P.ray.o = v3(0.f, 0.f, 0.f); P.ray.d = v3(0.f, 0.f, 1.f); P.color = v3(1.f, 1.f, 1.f);
        P.first_t = kInf; P.hit_t = 0.f; P.hit_ref = omd::REF_NONE; P.seg = 0u; P.draw = 0u; P.b0 = P.b1 = P.b2 = P.b3 = 0u;
        P.first_lo = P.first_hi = 0u; P.n = 0u; P.mt = 0.f; P.mleft = 0xFFFFFFFFu;
        const bool live = s < POOL_N && bind_pixel(P, k0, true);
        if (live) new_sample(P);
        if (s < POOL_N) pool_store(sl, s, P);
        pool_push(q_ray(0u), &ctl.nE[0], live, s, lane);
    }
    __syncthreads();

    for (uint32_t it = 0;; it++) {
        const uint32_t cur = it & 1u, nxt = cur ^ 1u;
        const uint32_t nE = ctl.nE[cur];
        {   // no ray, no march in progress, nothing re-filed by a march that just ended: every path of this CTA is finished
            uint32_t waiting = nE;
#pragma unroll
            for (uint32_t k = 0; k < PQ_COUNT; k++) waiting += ctl.nQ[cur][k];
            if (waiting == 0u) break;
        }
        // ---- phase E: closest traced hit of every queued ray (hits.rs:270-286) ----
        if (threadIdx.x == 0) {  // the other parity's phase S counters are free now
            ctl.tS[nxt] = 0u;
            for (uint32_t k = 0; k < PQ_COUNT; k++) if (k != PQ_MARCH) ctl.nQ[nxt][k] = 0u;
        }
        for (;;) {
            uint32_t b = 0;
            if (lane == 0u) b = atomicAdd(&ctl.tE[cur], 32u);
            b = __shfl_sync(FULL, b, 0);
            if (b >= nE) break;
            const uint32_t i = b + lane;
            const bool alive = i < nE;
            const uint32_t s = alive ? (uint32_t)q_ray(cur)[i] : 0u;
            Ray ray;
            ray.o = v3(0.f, 0.f, 0.f); ray.d = v3(0.f, 0.f, 1.f);
            if (alive) {
                const float4 f0 = sl[s], f1 = sl[POOL_N + s];
                ray.o = v3(f0.x, f0.y, f0.z); ray.d = v3(f0.w, f1.x, f1.y);
            }
            if (COUNT && alive) cn.segments++;
            const Best best = closest_traced_warp<true, COUNT>(sv, S, ray, A.tmin, A.tmax, brute, alive, cn, stack);
            uint32_t kind = 0xFFFFFFFFu;
            if (alive) {
                bool marches = false;
                if (S.n_marched != 0u) {  // see closest_hit_warp: a ray that misses the marched objects' bounds cannot meet one
                    const RayInv q = ray_inv(ray);
                    marches = slab(S.march_lo[0], S.march_lo[1], S.march_lo[2], S.march_hi[0], S.march_hi[1], S.march_hi[2], q, A.tmin, fminf(A.tmax, best.t)) < kInf;
                }
                if (marches) kind = PQ_MARCH;
                else {
                    const uint32_t c = shade_class(sv, S, best.ref);
                    kind = c == wfk::CLS_SKY ? PQ_SKY : (c == wfk::CLS_LAMBERTIAN ? PQ_LAMB : (c == wfk::CLS_METAL ? PQ_METAL : (c == wfk::CLS_DIELECTRIC ? PQ_DIEL : PQ_OTHER)));
                }
                reinterpret_cast<float2*>(&sl[2u * POOL_N + s])[1] = make_float2(best.t, __uint_as_float(best.ref));
            }
            const unsigned mine = __match_any_sync(FULL, kind);
            const uint32_t leader = (uint32_t)__ffs(mine) - 1u;
            uint32_t base = 0;
            if (kind != 0xFFFFFFFFu && lane == leader) base = atomicAdd(&ctl.nQ[cur][kind], (uint32_t)__popc(mine));
            base = __shfl_sync(FULL, base, leader);
            if (kind != 0xFFFFFFFFu) q_kind(cur, kind)[base + (uint32_t)__popc(mine & ((1u << lane) - 1u))] = (uint16_t)s;
        }
        __syncthreads();
        // ---- phase S: marching, scatter, Stats::add, next camera ray — one kind of work per batch ----
        if (threadIdx.x == 0) { ctl.nE[cur] = 0u; ctl.tE[cur] = 0u; }  // this parity's ray queue is consumed
        uint32_t cnts[PQ_COUNT], total = 0;
#pragma unroll
        for (uint32_t k = 0; k < PQ_COUNT; k++) { cnts[k] = ctl.nQ[cur][k]; total += (cnts[k] + 31u) >> 5; }
        for (;;) {
            uint32_t t = 0;
            if (lane == 0u) t = atomicAdd(&ctl.tS[cur], 1u);
            t = __shfl_sync(FULL, t, 0);
            if (t >= total) break;
            uint32_t kind = PQ_COUNT, n_kind = 0;  // ticket -> (kind, batch within the kind)
#pragma unroll
            for (uint32_t k = 0; k < PQ_COUNT; k++) {
                const uint32_t nb = (cnts[k] + 31u) >> 5;
                if (kind == PQ_COUNT) { if (t < nb) { kind = k; n_kind = cnts[k]; } else t -= nb; }
            }
            const uint32_t i = t * 32u + lane;
            const bool alive = i < n_kind;
            const uint32_t s = alive ? (uint32_t)q_kind(cur, kind)[i] : 0u;
            PoolPath P;
            pool_load(sl, s, P);
            Hit h;
            h.t = P.hit_t; h.ref = P.hit_ref; h.point = v3(0.f, 0.f, 0.f); h.normal = v3(0.f, 0.f, 0.f);
            if (kind == PQ_MARCH) {
                // A marching batch only marches (a bounded number of steps: the phase is as long as its longest batch).  A march
                // that ends re-files its slot by what the segment needs now - the marched hit (normal recomputed by finish_hit, as
                // k_render does) or the traced record that stands - for the NEXT phase S; one that does not stays in the march file.
                uint32_t refile = 0xFFFFFFFFu;
                bool still_marching = false;
                if (alive) {
                    Hit m;
                    float mt = P.mt;
                    uint32_t mleft = P.mleft;
                    bool over;
                    if constexpr (MA == 1) over = march_some<true, COUNT, true, false>(sv, S, P.ray, A.tmin, A.tmax, P.hit_t, mt, mleft, OM_POOL_MARCH_STEPS, m, &cn);
                    else if constexpr (MA == 3) over = march_some<true, COUNT, false, true>(sv, S, P.ray, A.tmin, A.tmax, P.hit_t, mt, mleft, OM_POOL_MARCH_STEPS, m, &cn);
                    else over = march_some<true, COUNT, false, false>(sv, S, P.ray, A.tmin, A.tmax, P.hit_t, mt, mleft, OM_POOL_MARCH_STEPS, m, &cn);
                    if (!over) {
                        still_marching = true;
                        sl[5u * POOL_N + s].w = mt;
                        reinterpret_cast<uint32_t*>(&sl[3u * POOL_N + s])[1] = (P.seg & 0xFFFFu) | (mleft << 16);
                    } else {
                        if (m.ref != omd::REF_NONE) { reinterpret_cast<float2*>(&sl[2u * POOL_N + s])[1] = make_float2(m.t, __uint_as_float(m.ref)); refile = PQ_OTHER; }
                        else {
                            const uint32_t c = shade_class(sv, S, P.hit_ref);
                            refile = c == wfk::CLS_SKY ? PQ_SKY : (c == wfk::CLS_LAMBERTIAN ? PQ_LAMB : (c == wfk::CLS_METAL ? PQ_METAL : (c == wfk::CLS_DIELECTRIC ? PQ_DIEL : PQ_OTHER)));
                        }
                    }
                }
                pool_push(q_kind(nxt, PQ_MARCH), &ctl.nQ[nxt][PQ_MARCH], still_marching, s, lane);
#pragma unroll
                for (uint32_t k = 1; k < PQ_COUNT; k++) pool_push(q_kind(nxt, k), &ctl.nQ[nxt][k], refile == k, s, lane);
                continue;
            }
            const bool shade = alive;
            bool has_ray = false;
            if (shade) {
                Rng g;
                g.rk = A.rk; g.a = P.pxl; g.b = A.sample_base + P.n; g.draw = P.draw;
                g.b0 = P.b0; g.b1 = P.b1; g.b2 = P.b2; g.b3 = P.b3;
                bool finished;
                V3 result = P.color;
                if (h.ref != omd::REF_NONE) {  // handle_hit Some(hr), render_thread.rs:110-116
                    finish_hit<true, MA == 3>(sv, S, P.ray, A.tmin, h);
                    const Material m = load_material(sv, S, h.ref);
                    if (P.seg == 0u) { P.first_t = h.t; P.first_lo = m.bloom_lo; P.first_hi = m.bloom_hi; }
                    P.color = P.color * scatter(m, P.ray, h.point, h.normal, g);
                    P.seg++;
                    finished = P.seg >= nseg;
                    if (finished) result = -v3(0.f, 0.f, 0.f);  // out of depth: -black (:142)
                } else {  // None: sky (:117-123)
                    const float tt = 0.5f * (P.ray.d.y + 1.0f);
                    const V3 sky = (1.0f - tt) * v3(1.f, 1.f, 1.f) + tt * v3(0.5f, 0.7f, 1.0f);
                    result = P.color * sky;
                    finished = true;
                }
                P.draw = g.draw; P.b0 = g.b0; P.b1 = g.b1; P.b2 = g.b2; P.b3 = g.b3;
                P.mleft = 0xFFFFFFFFu; P.mt = 0.f;
                has_ray = true;
                if (finished) {  // pixel.stats.add(color, depth, obj_id), render_thread.rs:194
                    const float4 a = A.acc[P.pxl];
                    const uint4 x = A.aux[P.pxl];
                    Stats st;
                    st.sum = v3(a.x, a.y, a.z); st.n = __float_as_uint(a.w);
                    st.avg_depth = __uint_as_float(x.x); st.bad_avgs = x.y; st.bloom_lo = x.z; st.bloom_hi = x.w;
                    st.color = stats_color_of(st.sum, st.n);
                    bool done = stats_add(st, result, P.first_t, P.first_lo, P.first_hi);
                    if (!adaptive) done = false;
                    A.acc[P.pxl] = make_float4(st.sum.x, st.sum.y, st.sum.z, __uint_as_float(st.n));
                    A.aux[P.pxl] = make_uint4(__float_as_uint(st.avg_depth), st.bad_avgs, st.bloom_lo, st.bloom_hi);
                    traced++;
                    P.todo--;
                    P.n = st.n;
                    const uint32_t left = A.spp_budget - st.n;
                    credited += (done ? (unsigned long long)left : 0ull) + 1ull;  // render_thread.rs:196-198
                    if (done) P.todo = 0u;
                    if (P.todo == 0u) has_ray = bind_pixel(P, 0u, false);
                    if (has_ray) new_sample(P);
                }
            }
            if (alive) pool_store(sl, s, P);
            pool_push(q_ray(nxt), &ctl.nE[nxt], has_ray, s, lane);
        }
        __syncthreads();
        if (threadIdx.x == 0) ctl.nQ[cur][PQ_MARCH] = 0u;  // consumed; the next writer of this parity's march file is phase S of it + 1 ... after a barrier
    }

    // ---- progress counters: warp-aggregate, one atomic per warp ----
    for (int o = 16; o > 0; o >>= 1) {
        credited += __shfl_xor_sync(FULL, credited, o);
        traced += __shfl_xor_sync(FULL, traced, o);
    }
    if (lane == 0u) {
        if (credited) atomicAdd(&A.counters[CTR_CREDITED], credited);
        if (traced) atomicAdd(&A.counters[CTR_TRACED], traced);
    }
    if (COUNT) {
        unsigned long long v[7] = {cn.segments, cn.node_visits, cn.prim_tests, cn.sdf_evals, cn.march_steps, cn.normal_evals, cn.march_entered};
        for (int i = 0; i < 7; i++) {
            for (int o = 16; o > 0; o >>= 1) v[i] += __shfl_xor_sync(FULL, v[i], o);
            if (lane == 0u && v[i]) atomicAdd(&A.counters[CTR_SEGMENTS + i], v[i]);
        }
    }
}
