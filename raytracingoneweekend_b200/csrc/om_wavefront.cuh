// Wavefront pipeline of the B200 renderer: the recursion of ray_color (render_thread.rs:128-143)
// as iterative queues.  Included after om_device.cuh inside namespace OM_NS (both builds).
//
// One ITERATION advances every live path by one segment with three small kernels:
//   k_wf_extend  FrozenHittableList::hit, traced part (hits.rs:270-286): BVH in shared memory, warps
//                refill finished lanes from the ray queue (persistent, voting inner/leaf/fetch)
//   k_wf_march   hits.rs:288-365 for the rays that enter the marched objects' bounds, a bounded number
//                of steps per launch
//   k_wf_shade   handle_hit / Material::scatter / sky (render_thread.rs:105-126, materials.rs:39-95),
//                Stats::add on termination, next sample or next pixel (camera ray), and the
//                ballot+popc compaction of the live slots into the next iteration's ray queue
// Resolved segments are queued by SHADE CLASS (sky / Lambertian / metal / dielectric sphere / other
// primitive) so that a warp of k_wf_shade only ever sees one kind of work.
// Why not one kernel: the megakernel is 70-78 KB of SASS against a 32 KB L1.5 instruction cache
// (40-48 % of its stall samples are `no_instruction`) and its shading code diverges inside the warp;
// here each kernel is a few KB.
//
// A SLOT carries one path and is bound to one pixel until that pixel has all its samples of this
// launch (so Stats::add sees the samples in order: it is order-dependent), then takes the next
// unstarted pixel.  The pool of slots is sized to stay resident in L2.  Slot record, 96 bytes:
//   ray[2*slot+0] = {origin.xyz, sample index}      ray[2*slot+1] = {direction.xyz, pixel}
//   st[4*slot+0]  = {throughput.rgb, primary-hit t} st[4*slot+1]  = {segment, Philox draw, bloom lo, hi}
//   st[4*slot+2]  = {hit t, hit ref, march t, march iterations left}      st[4*slot+3] unused
// so that every kernel reads and writes whole 32-byte sectors whatever order the queues are in.
#pragma once

namespace wfk {
using namespace omk::wf;
}

#ifndef OM_WF_EXT_MIN_CTAS
#define OM_WF_EXT_MIN_CTAS 2  // CTAs per SM the extend kernel is compiled for (64 registers at 2)
#endif
#ifndef OM_WF_EXT_THREADS
#define OM_WF_EXT_THREADS 512
#endif
constexpr int WF_EXT_THREADS = OM_WF_EXT_THREADS;
constexpr int WF_THREADS = 256;
constexpr uint32_t WF_DESTS = (uint32_t)omk::wf::N_CLS + 1u;                 // shade classes + the march queue
constexpr size_t WF_STAGE_BYTES = size_t(WF_EXT_THREADS / 32) * WF_DESTS * 32u * 4u;  // k_wf_extend: per-warp staging of retired slots
enum : uint32_t { WS_DONE = 0u, WS_FETCH = 1u, WS_INNER = 2u, WS_LEAF = 3u };
constexpr uint32_t WF_NONE = 0xFFFFFFFFu;

__device__ __forceinline__ uint32_t* wf_set(const omk::WfArgs& W, uint32_t it) { return W.ctr + (it % 3u) * wfk::C_SET_WORDS; }

// shade class of a resolved segment (wf::CLS_*)
template <bool SMEM>
__device__ __forceinline__ uint32_t shade_class(const SceneView<SMEM>& sv, const SceneHeader& S, uint32_t ref) {
    if (ref == omd::REF_NONE) return wfk::CLS_SKY;
    if (omd::ref_type(ref) != omd::PT_SPHERE) return wfk::CLS_OTHER;
    const uint32_t mt = sv.hot_byte(S.off_sph_mat, omd::ref_index(ref));
    return wfk::CLS_LAMBERTIAN + (mt > 2u ? 0u : mt);
}
// destination d of a resolved slot: d < N_CLS = shade class queue d, d == N_CLS = march queue
__device__ __forceinline__ uint32_t* wf_dest_counter(uint32_t* set, uint32_t d) { return set + (d < (uint32_t)wfk::N_CLS ? wfk::C_N_SHADE + wfk::C_STRIDE * d : wfk::C_N_MARCH); }
__device__ __forceinline__ uint32_t* wf_dest_queue(const omk::WfArgs& W, uint32_t parity, uint32_t d) {
    return d < (uint32_t)wfk::N_CLS ? W.q_shade[parity] + size_t(d) * W.n_slots : W.q_march[parity];
}
// Warp-aggregated append of `slot` to destination `dest` (WF_NONE = nothing) of iteration parity `parity`,
// counters in `set`.  Lanes with the same destination find each other with match.any; the lowest of each
// group does the group's single atomicAdd, and all groups' atomics are in flight together.  All 32 lanes call.
__device__ __forceinline__ void wf_push(const omk::WfArgs& W, uint32_t* set, uint32_t parity, uint32_t dest, uint32_t slot, uint32_t lane) {
    const unsigned mine = __match_any_sync(FULL, dest);
    const uint32_t leader = (uint32_t)__ffs(mine) - 1u;
    uint32_t b = 0;
    if (dest != WF_NONE && lane == leader) b = atomicAdd(wf_dest_counter(set, dest), (uint32_t)__popc(mine));
    b = __shfl_sync(FULL, b, leader);
    if (dest != WF_NONE) wf_dest_queue(W, parity, dest)[b + __popc(mine & ((1u << lane) - 1u))] = slot;
}

// ---------------------------------------------------------------------------------------------
// k_wf_extend
// ---------------------------------------------------------------------------------------------
template <bool SMEM, bool COUNT>
__global__ void __launch_bounds__(WF_EXT_THREADS, OM_WF_EXT_MIN_CTAS) k_wf_extend(const __grid_constant__ RenderArgs A, const __grid_constant__ omk::WfArgs W) {
    uint32_t* cs = wf_set(W, W.it);
    const uint32_t n_ray = cs[wfk::C_N_RAY];
    if (n_ray == 0u) return;
    if (SMEM) stage_scene_smem(A.hot_g, A.scene.hot_units);
    SceneView<SMEM> sv;
    sv.hot_g = A.hot_g;
    sv.cold_g = A.cold_g;
    const SceneHeader& S = A.scene;
    const bool brute = (A.flags & OM_RENDER_BRUTE_FORCE) != 0;
    const uint32_t* __restrict__ q_in = W.q_ray[W.it & 1u];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t n_lin = brute ? (S.n[0] + S.n[1] + S.n[2] + S.n[3] + S.n[4] + S.n[omd::PT_TDYN]) : S.n_large;
    const bool have_bvh = !brute && S.n_nodes > 0u;
    const bool have_marched = S.n_marched != 0u;
    Counters cn = {0, 0, 0, 0, 0, 0, 0};
    TravStack<false, WF_EXT_THREADS> stack;  // local memory: two CTAs per SM leave no room for a shared stack beside the scene

    uint32_t state = WS_FETCH;
    uint32_t slot = WF_NONE;  // the ray this lane is tracing
    Ray ray;
    ray.o = v3(0.f, 0.f, 0.f); ray.d = v3(0.f, 0.f, 1.f);
    RayInv q = ray_inv(ray);
    Best best;
    best.t = A.tmax; best.ref = omd::REF_NONE;
    uint32_t k = 0;
    int sp = 0, cur = 0;
    float te_march = kInf;
    // the warp takes ray-queue entries 32 at a time: lane j holds entry j of the chunk (WF_NONE past the end)
    uint32_t chunk_slot = WF_NONE, chunk_used = 32u;
    // Retired slots are staged per destination in shared memory and leave 32 at a time: one atomicAdd and one
    // 128-byte store per 32 slots (a counter per retiring group cost more than the traversal: atomics on one
    // address serialise in L2).  Lane d < WF_DESTS counts the entries staged for destination d.
    uint32_t* stage = reinterpret_cast<uint32_t*>(om_smem + (SMEM ? S.hot_units : 0u)) + (threadIdx.x >> 5) * (WF_DESTS * 32u);
    uint32_t my_cnt = 0;
    const uint32_t parity = W.it & 1u;

    auto pop = [&]() {
        while (sp > 0) {
            sp--;
            if (stack.pop(sp, best.t, sv, S, cur)) { state = cur >= 0 ? WS_INNER : WS_LEAF; return; }
        }
        state = WS_FETCH;  // traversal over: retire the result, take the next ray
    };
    auto next_item = [&]() {
        if (k < n_lin) { cur = ~(int)linear_ref(sv, S, brute, k); k++; state = WS_LEAF; return; }
        if (k == n_lin) { k = n_lin + 1u; if (have_bvh) { cur = 0; state = WS_INNER; return; } }
        pop();
    };

    for (;;) {
        const unsigned m_in = __ballot_sync(FULL, state == WS_INNER);
        const unsigned m_lf = __ballot_sync(FULL, state == WS_LEAF);
        const unsigned m_f = __ballot_sync(FULL, state == WS_FETCH);
        if ((m_in | m_lf | m_f) == 0u) break;
        const int c_in = __popc(m_in), c_lf = __popc(m_lf), c_f = __popc(m_f);
        if (c_f >= 6 || (c_in | c_lf) == 0) {
            // ---- the next rays: the f-th fetching lane takes entry chunk_used + f; beyond the chunk it waits ----
            if (chunk_used >= 32u) {
                uint32_t b = 0;
                if (lane == 0u) b = atomicAdd(&cs[wfk::C_CUR_RAY], 32u);
                b = __shfl_sync(FULL, b, 0);
                chunk_slot = (b + lane < n_ray) ? q_in[b + lane] : WF_NONE;
                chunk_used = 0u;
            }
            const uint32_t rank = chunk_used + (uint32_t)__popc(m_f & lt_mask);
            const uint32_t new_slot = __shfl_sync(FULL, chunk_slot, rank & 31u);
            const bool fetching = state == WS_FETCH && rank < 32u;
            float4 nro = make_float4(0.f, 0.f, 0.f, 0.f), nrd = make_float4(0.f, 0.f, 1.f, 0.f);
            if (fetching && new_slot != WF_NONE) { nro = W.s_ray[2u * new_slot]; nrd = W.s_ray[2u * new_slot + 1u]; }
            chunk_used = min(32u, chunk_used + (uint32_t)c_f);
            // ---- retire the finished rays: hit record, then queue by what the segment needs next ----
            const bool had = state == WS_FETCH && slot != WF_NONE;
            uint32_t dest = WF_NONE;
            if (had) {
                // A ray whose [t_min, min(t_max, closest)] span misses the padded union of the marched bounds can
                // neither start within HIT_SIZE of a marched surface nor reach one (hits.rs:288-332 would end
                // without touching `rec`); the others go through k_wf_march before they are shaded.
                W.s_st[4u * slot + 2u] = make_float4(best.t, __uint_as_float(best.ref), 0.f, __uint_as_float(WF_NONE));
                dest = (te_march <= best.t) ? (uint32_t)wfk::N_CLS : shade_class(sv, S, best.ref);
            }
            {
                const unsigned mine = __match_any_sync(FULL, dest);
                const uint32_t pos = __shfl_sync(FULL, my_cnt, dest < WF_DESTS ? dest : 0u) + (uint32_t)__popc(mine & lt_mask);
                if (dest != WF_NONE && pos < 32u) stage[dest * 32u + pos] = slot;
                __syncwarp();
#pragma unroll
                for (uint32_t d = 0; d < WF_DESTS; d++) {
                    uint32_t tot = __shfl_sync(FULL, my_cnt, d) + (uint32_t)__popc(__ballot_sync(FULL, dest == d));
                    if (tot >= 32u) {  // warp-uniform: a full line leaves
                        uint32_t b = 0;
                        if (lane == 0u) b = atomicAdd(wf_dest_counter(cs, d), 32u);
                        b = __shfl_sync(FULL, b, 0);
                        wf_dest_queue(W, parity, d)[b + lane] = stage[d * 32u + lane];
                        __syncwarp();
                        if (dest == d && pos >= 32u) stage[d * 32u + pos - 32u] = slot;
                        tot -= 32u;
                    }
                    if (lane == d) my_cnt = tot;
                }
                __syncwarp();
            }
            if (had) slot = WF_NONE;
            // ---- start the new rays ----
            if (fetching) {
                if (new_slot != WF_NONE) {
                    slot = new_slot;
                    ray.o = v3(nro.x, nro.y, nro.z); ray.d = v3(nrd.x, nrd.y, nrd.z);
                    if (COUNT) cn.segments++;
                    q = ray_inv(ray);
                    best.t = A.tmax; best.ref = omd::REF_NONE;
                    k = 0; sp = 0;
                    te_march = have_marched ? slab(S.march_lo[0], S.march_lo[1], S.march_lo[2], S.march_hi[0], S.march_hi[1], S.march_hi[2], q, A.tmin, A.tmax) : kInf;
                    next_item();  // (a scene without traced objects leaves WS_FETCH: the next refill retires the ray)
                } else state = WS_DONE;  // the queue is empty
            }
        } else if (c_in >= c_lf) {
            if (state == WS_INNER) {
                if (COUNT) cn.node_visits++;
                if (visit4(sv, S, (uint32_t)cur, q, ray.o, S.dyn_pad, A.tmin, best.t, stack, sp, cur)) state = cur >= 0 ? WS_INNER : WS_LEAF;
                else pop();
            }
        } else {
            if (state == WS_LEAF) {
                if (COUNT) cn.prim_tests++;
                test_prim(sv, S, (uint32_t)(~cur), ray, A.tmin, best);
                next_item();
            }
        }
    }
    __syncwarp();
#pragma unroll
    for (uint32_t d = 0; d < WF_DESTS; d++) {  // what is left in the staging lines
        const uint32_t cnt = __shfl_sync(FULL, my_cnt, d);
        if (cnt == 0u) continue;
        uint32_t b = 0;
        if (lane == 0u) b = atomicAdd(wf_dest_counter(cs, d), cnt);
        b = __shfl_sync(FULL, b, 0);
        if (lane < cnt) wf_dest_queue(W, parity, d)[b + lane] = stage[d * 32u + lane];
    }
    if (COUNT) {
        unsigned long long v[3] = {cn.segments, cn.node_visits, cn.prim_tests};
        for (int i = 0; i < 3; i++) {
            for (int o = 16; o > 0; o >>= 1) v[i] += __shfl_xor_sync(FULL, v[i], o);
            if (lane == 0u && v[i]) atomicAdd(&A.counters[CTR_SEGMENTS + i], v[i]);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// wf_march_warp: unstuck + sphere tracing (hits.rs:288-365) for 32 queued slots, at most `steps` SDF
// steps per iteration: a grazing ray takes up to 1024 dependent steps and a launch is as slow as its
// slowest lane, so an unfinished march keeps its (t, iterations left) and continues next iteration
// while the rest of the pool moves on.  Runs inside k_wf_shade (the march warps hide behind the
// shading warps' memory latency).  A finished march queues its slot for the NEXT iteration's shading.
// ---------------------------------------------------------------------------------------------
template <bool COUNT>
__device__ __noinline__ void wf_march_warp(const RenderArgs& A, const omk::WfArgs& W, uint32_t i, uint32_t n, uint32_t lane, Counters& cn) {
    SceneView<false> sv;
    sv.hot_g = A.hot_g;
    sv.cold_g = A.cold_g;
    const SceneHeader& S = A.scene;
    uint32_t* ns = wf_set(W, W.it + 1u);
    const int steps = (int)min(64u, max(8u, 512u / S.n_marched));
    uint32_t dest = WF_NONE, slot = 0;
    int2 mstk[MSTACK];  // scratch of nearest_marched's distance-BVH walk
    if (i < n) {
        slot = W.q_march[W.it & 1u][i];
        const float4 o = W.s_ray[2u * slot], d = W.s_ray[2u * slot + 1u];
        Ray ray;
        ray.o = v3(o.x, o.y, o.z); ray.d = v3(d.x, d.y, d.z);
        const float4 hm = W.s_st[4u * slot + 2u];
        const float closest = hm.x;  // the traced `closest_so_far` (or t_max)
        uint32_t ref = __float_as_uint(hm.y);
        float t_m = hm.z;
        int iter = 0;
        bool finished = false;
        // hm.w: WF_NONE = not started; 0x80000000 | object = unstuck (hits.rs:336-365) in progress on that marched
        // object, at t = hm.z; else the march iterations left, at t = hm.z
        int budget = steps;
        uint32_t code = __float_as_uint(hm.w);
        if (code == WF_NONE) {
            const Near n0 = nearest_marched(sv, S, ray_at(ray, A.tmin), mstk);
            if (COUNT) { cn.march_entered++; cn.sdf_evals += n0.evals; }
            if (n0.ref == omd::REF_NONE) finished = true;  // `marched.is_none()` -> INF -> return rec (also every NaN ray)
            else if (n0.d < HIT_SIZE) { code = 0x80000000u | n0.ref; t_m = A.tmin; }
            else { code = 1024u; t_m = A.tmin; }
        }
        if (!finished && (code & 0x80000000u)) {
            const float MIN_STEP_SIZE = HIT_SIZE / 2.f;
            const uint32_t uref = code & 0x7FFFFFFFu;
            float new_t = t_m;
            bool out = false;
            while (budget > 0) {  // `while aux < HIT_SIZE { new_t += MIN_STEP_SIZE; aux = |sdf(r.at(new_t))| }`
                new_t += MIN_STEP_SIZE;
                const float aux = abs_sdf_of(sv, S, uref, ray_at(ray, new_t));
                budget--;
                if (COUNT) cn.sdf_evals++;
                if (!(aux < HIT_SIZE) || new_t > A.tmin + 525.f) { out = true; break; }  // (the reference has no cap; the persistent kernel stops after 2^20 steps = 524 units)
            }
            t_m = new_t;
            if (out) code = 1024u;
        }
        iter = (int)code;
        float hit_t = closest;
        if (!finished && !(code & 0x80000000u)) {
#pragma unroll 1
            for (; budget > 0; budget--) {  // hits.rs:294-332
                if (!(t_m < A.tmax && t_m < closest && iter > 0)) { finished = true; break; }
                iter--;
                const Near nn = nearest_marched(sv, S, ray_at(ray, t_m), mstk);
                if (COUNT) { cn.march_steps++; cn.sdf_evals += nn.evals; }
                if (nn.ref == omd::REF_NONE) { finished = true; break; }  // `material.is_none()`
                if (nn.d < HIT_SIZE) {
                    // the reference evaluates the nearest object's normal at every step (hits.rs:306); only the
                    // value at the accepted point survives, and k_wf_shade computes it there
                    ref = nn.ref; hit_t = t_m;
                    if (COUNT) cn.normal_evals++;
                    finished = true;
                    break;
                }
                t_m += nn.d;
            }
        }
        if (finished) { W.s_st[4u * slot + 2u] = make_float4(hit_t, __uint_as_float(ref), 0.f, 0.f); dest = shade_class(sv, S, ref); }
        else { W.s_st[4u * slot + 2u] = make_float4(closest, __uint_as_float(ref), t_m, __uint_as_float((uint32_t)iter)); dest = wfk::N_CLS; }
    }
    wf_push(W, ns, (W.it + 1u) & 1u, dest, slot, lane);
}

// ---------------------------------------------------------------------------------------------
// k_wf_shade
// ---------------------------------------------------------------------------------------------
template <bool COUNT>
__global__ void __launch_bounds__(WF_THREADS, 4) k_wf_shade(const __grid_constant__ RenderArgs A, const __grid_constant__ omk::WfArgs W) {
    uint32_t* cs = wf_set(W, W.it);
    // work = the class queues and the march queue back to back, each padded to whole warps: a warp only ever
    // sees one kind of work
    uint32_t n0 = 0, n1 = 0, n2 = 0, n3 = 0, n4 = 0, n5 = 0;
    if (!W.init) {
        n0 = cs[wfk::C_N_SHADE]; n1 = cs[wfk::C_N_SHADE + wfk::C_STRIDE]; n2 = cs[wfk::C_N_SHADE + 2 * wfk::C_STRIDE];
        n3 = cs[wfk::C_N_SHADE + 3 * wfk::C_STRIDE]; n4 = cs[wfk::C_N_SHADE + 4 * wfk::C_STRIDE]; n5 = cs[wfk::C_N_MARCH];
    }
    const uint32_t e0 = (n0 + 31u) & ~31u, e1 = e0 + ((n1 + 31u) & ~31u), e2 = e1 + ((n2 + 31u) & ~31u), e3 = e2 + ((n3 + 31u) & ~31u), e4 = e3 + ((n4 + 31u) & ~31u);
    const uint32_t e5 = e4 + ((n5 + 31u) & ~31u);
    const uint32_t total = W.init ? ((W.n_slots + 31u) & ~31u) : e5;
    if (!W.init && blockIdx.x == 0) for (uint32_t w = threadIdx.x; w < (uint32_t)wfk::C_SET_WORDS; w += WF_THREADS) wf_set(W, W.it + 2u)[w] = 0u;
    if (total == 0u) return;
    SceneView<false> sv;
    sv.hot_g = A.hot_g;
    sv.cold_g = A.cold_g;
    const SceneHeader& S = A.scene;
    const bool adaptive = (A.flags & OM_RENDER_ADAPTIVE) != 0;
    const uint32_t nseg = A.max_depth > 1u ? A.max_depth : 1u;
    const uint32_t target = min(A.spp_budget, W.n_target);
    const uint32_t lane = threadIdx.x & 31u;
    unsigned long long credited = 0, traced = 0;
    Counters cn = {0, 0, 0, 0, 0, 0, 0};

    for (uint32_t base = blockIdx.x * WF_THREADS + (threadIdx.x & ~31u); base < total; base += gridDim.x * WF_THREADS) {  // per warp
        uint32_t cls = 0, off = base, cnt = W.n_slots;
        if (!W.init) {
            if (base < e0) { cls = 0; off = base; cnt = n0; }
            else if (base < e1) { cls = 1; off = base - e0; cnt = n1; }
            else if (base < e2) { cls = 2; off = base - e1; cnt = n2; }
            else if (base < e3) { cls = 3; off = base - e2; cnt = n3; }
            else if (base < e4) { cls = 4; off = base - e3; cnt = n4; }
            else {  // the march queue (warp-uniform branch)
                wf_march_warp<COUNT>(A, W, base - e4 + lane, n5, lane, cn);
                continue;
            }
        }
        const uint32_t i = off + lane;
        const bool have = i < cnt;
        uint32_t slot = 0, pix = WF_NONE, sample = 0;
        bool live = false, fresh = false, need_pixel = false;
        Ray ray;
        ray.o = v3(0.f, 0.f, 0.f); ray.d = v3(0.f, 0.f, 1.f);
        V3 color = v3(1.f, 1.f, 1.f);
        float first_t = kInf;
        uint32_t seg = 0, first_lo = 0, first_hi = 0;
        Rng g;
        rng_init(g, A.rk, 0u, 0u);
        g.b0 = g.b1 = g.b2 = g.b3 = 0u;
        if (have && W.init) { slot = i; need_pixel = true; }
        if (have && !W.init) {
            slot = W.q_shade[W.it & 1u][size_t(cls) * W.n_slots + i];
            const float4 o = W.s_ray[2u * slot], d = W.s_ray[2u * slot + 1u];
            const float4 th = W.s_st[4u * slot], mi = W.s_st[4u * slot + 1u], hm = W.s_st[4u * slot + 2u];
            ray.o = v3(o.x, o.y, o.z); ray.d = v3(d.x, d.y, d.z);
            sample = __float_as_uint(o.w); pix = __float_as_uint(d.w);
            color = v3(th.x, th.y, th.z); first_t = th.w;
            seg = __float_as_uint(mi.x); first_lo = __float_as_uint(mi.z); first_hi = __float_as_uint(mi.w);
            rng_init(g, A.rk, pix, sample);
            g.draw = __float_as_uint(mi.y);
            const uint32_t href = __float_as_uint(hm.y);
            bool finished;
            V3 result = color;
            if (href != omd::REF_NONE) {  // handle_hit Some(hr), render_thread.rs:110-116
                Hit h;
                h.t = hm.x; h.ref = href;
                h.point = v3(0.f, 0.f, 0.f); h.normal = v3(0.f, 0.f, 0.f);
                finish_hit(sv, S, ray, A.tmin, h);
                const Material m = load_material(sv, S, h.ref);
                if (seg == 0u) { first_t = h.t; first_lo = m.bloom_lo; first_hi = m.bloom_hi; }
                if ((g.draw & 3u) != 0u) {  // the block the stream is in the middle of (not carried between kernels)
                    const uint4 b = philox4x32_10(g.draw >> 2, g.a, g.b, g.rk);
                    g.b0 = b.x; g.b1 = b.y; g.b2 = b.z; g.b3 = b.w;
                }
                color = color * scatter(m, ray, h.point, h.normal, g);
                seg++;
                finished = seg >= nseg;
                if (finished) result = -v3(0.f, 0.f, 0.f);  // out of depth: -black (:142)
            } else {  // None: sky (:117-123)
                const float t = 0.5f * (ray.d.y + 1.0f);
                const V3 sky = (1.0f - t) * v3(1.f, 1.f, 1.f) + t * v3(0.5f, 0.7f, 1.0f);
                result = color * sky;
                finished = true;
            }
            if (finished) {  // pixel.stats.add(color, depth, obj_id), render_thread.rs:194
                const float4 a = A.acc[pix];
                const uint4 x = A.aux[pix];
                Stats st;
                st.sum = v3(a.x, a.y, a.z); st.n = __float_as_uint(a.w);
                st.avg_depth = __uint_as_float(x.x); st.bad_avgs = x.y; st.bloom_lo = x.z; st.bloom_hi = x.w;
                st.color = stats_color_of(st.sum, st.n);
                bool done = stats_add(st, result, first_t, first_lo, first_hi);
                if (!adaptive) done = false;
                A.acc[pix] = make_float4(st.sum.x, st.sum.y, st.sum.z, __uint_as_float(st.n));
                A.aux[pix] = make_uint4(__float_as_uint(st.avg_depth), st.bad_avgs, st.bloom_lo, st.bloom_hi);
                traced++;
                const uint32_t left = A.spp_budget - st.n;
                credited += (done ? (unsigned long long)left : 0ull) + 1ull;  // render_thread.rs:196-198
                if (!done && st.n < target) { sample = A.sample_base + st.n; fresh = true; }
                else need_pixel = true;
            } else live = true;
        }
        // ---- the slot's pixel is complete: take the next unstarted pixel that still wants samples ----
        while (need_pixel) {
            const uint32_t kk = atomicAdd(&W.ctr[wfk::G_NEXT_WORK], 1u);
            if (kk >= A.n_work) { pix = WF_NONE; break; }
            const uint32_t p = work_to_pixel(A, kk);
            if (p == WF_NONE) continue;
            const uint32_t pn = __float_as_uint(A.acc[p].w);
            if (pn < target && !(adaptive && A.aux[p].y >= 5u)) { pix = p; sample = A.sample_base + pn; fresh = true; need_pixel = false; }
        }
        if (fresh) {  // render_thread.rs:183-192: jittered (u,v), camera ray
            rng_init(g, A.rk, pix, sample);
            ray = primary_ray(A, pix, sample, g);
            color = v3(1.f, 1.f, 1.f);
            seg = 0; first_t = kInf; first_lo = 0; first_hi = 0;
            live = true;
        }
        if (have && live) {
            W.s_ray[2u * slot] = make_float4(ray.o.x, ray.o.y, ray.o.z, __uint_as_float(sample));
            W.s_ray[2u * slot + 1u] = make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(pix));
            W.s_st[4u * slot] = make_float4(color.x, color.y, color.z, first_t);
            W.s_st[4u * slot + 1u] = make_float4(__uint_as_float(seg), __uint_as_float(g.draw), __uint_as_float(first_lo), __uint_as_float(first_hi));
        }
        // the slot has a ray for the next iteration: k_wf_compact gathers the flags into the ray queue in SLOT
        // order (neighbouring slots hold neighbouring pixels: their rays traverse the BVH together)
        if (have && live) W.s_live[slot] = 1u;
    }
    for (int o = 16; o > 0; o >>= 1) {
        credited += __shfl_xor_sync(FULL, credited, o);
        traced += __shfl_xor_sync(FULL, traced, o);
    }
    if (lane == 0u) {
        if (credited) atomicAdd(&A.counters[CTR_CREDITED], credited);
        if (traced) atomicAdd(&A.counters[CTR_TRACED], traced);
    }
    if (COUNT) {
        unsigned long long v[4] = {cn.sdf_evals, cn.march_steps, cn.normal_evals, cn.march_entered};
        for (int i = 0; i < 4; i++)
            if (v[i]) atomicAdd(&A.counters[CTR_SDF_EVALS + i], v[i]);
    }
}

// ---------------------------------------------------------------------------------------------
// k_wf_compact: live-ray compaction.  Gathers the slots that k_wf_shade flagged into the next
// iteration's ray queue in slot order: 16 flags per lane, warp prefix sum, one atomicAdd per 512 slots.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(WF_THREADS) k_wf_compact(const __grid_constant__ omk::WfArgs W) {
    uint32_t* ns = W.init ? wf_set(W, W.it) : wf_set(W, W.it + 1u);
    uint32_t* __restrict__ q_out = W.q_ray[W.init ? 0u : ((W.it + 1u) & 1u)];
    const uint32_t lane = threadIdx.x & 31u;
    uint4* flags = reinterpret_cast<uint4*>(W.s_live);
    const uint32_t n16 = (W.n_slots + 15u) >> 4, n16w = (n16 + 31u) & ~31u;  // s_live is allocated in whole 512-slot groups
    for (uint32_t i = blockIdx.x * WF_THREADS + threadIdx.x; i < n16w; i += gridDim.x * WF_THREADS) {
        uint4 f = make_uint4(0u, 0u, 0u, 0u);
        if (i < n16) f = flags[i];
        const bool any = (f.x | f.y | f.z | f.w) != 0u;
        if (!__any_sync(FULL, any)) continue;
        if (any) flags[i] = make_uint4(0u, 0u, 0u, 0u);
        const uint32_t w[4] = {f.x, f.y, f.z, f.w};
        uint32_t cnt = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) cnt += __popc(w[k] & 0x01010101u);
        uint32_t incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, incl, o); if (lane >= (uint32_t)o) incl += t; }
        uint32_t b = 0;
        if (lane == 31u) b = atomicAdd(&ns[wfk::C_N_RAY], incl);
        b = __shfl_sync(FULL, b, 31) + incl - cnt;
#pragma unroll
        for (int k = 0; k < 16; k++)
            if ((w[k >> 2] >> (8 * (k & 3))) & 1u) q_out[b++] = 16u * i + (uint32_t)k;
    }
}
