// rt_bvh_render.cuh -- persistent bounce loop over the GPU-built LBVH with a RESUMABLE traversal and a cooperative end game.
// Included by rt_device.cuh: rtx (parity shading, rt_kernels_exact.cu) and rtb (throughput shading, rt_kernels_bvh.cu -- compiled
// without FMA contraction: the kernel shades at two sites, and they must be the same arithmetic).
//
// Traversal lengths are heavy tailed (sky rays leave after a handful of nodes, rays that skim the sphere field visit
// hundreds), so "every lane traverses its ray to the end, then the warp shades" keeps 6 of 32 lanes busy (ncu,
// profiles/r01_ncu_k_render_lbvh.json).  Here the traversal state (node, stack, running closest hit) lives across
// passes: the warp steps all traversals one node at a time and leaves the loop as soon as RT_BVH_IDLE lanes have
// finished; those lanes are shaded / regenerated (the others "hold"), start their next traversal, and the loop resumes.
// The stack lives in shared memory ([depth][thread], conflict free); depths beyond RT_BVH_SMEM_DEPTH (never seen on
// Morton trees of the test scenes) go to a local-memory overflow so the result never depends on the stack size.
// Semantics are those of closest_hit_lbvh (conservative boxes, reference test at the leaves, (t, index) order).
// Once a warp's work pool is dry and only a few lanes still traverse, the WHOLE warp finishes each remaining query
// (bvh_coop_query, bvh_straggler_loop below): single paths of 10^5 node visits were the end of every multi-GPU frame.
#pragma once

#ifndef RT_BVH_IDLE
#define RT_BVH_IDLE 24           // leave the traversal loop when this many lanes wait for shading
#endif
#ifndef RT_BVH_CTAS
#define RT_BVH_CTAS 4
#endif
#ifndef RT_BVH_STEPS
#define RT_BVH_STEPS 2              // node steps between two exit votes
#endif
#define RT_BVH_OVF_DEPTH 64
#ifndef RT_BVH_BATCH
#define RT_BVH_BATCH RT_BATCH       // work items a warp takes at once.  32 (one per lane) was measured against 256 on 2 GPUs for the
                                    // 1M-sphere frame: no change (382 vs 386 ms) -- the multi-GPU tail of that frame is single PATHS
                                    // that run for tens of ms (profiles/r02_trace_c5.log), not batches held by one warp
#endif
#ifndef RT_BVH_PATHSTAT
#define RT_BVH_PATHSTAT 0          // measurement build: > 0 prints every path with more node visits than this (device printf)
#endif
#ifndef RT_BVH_BRANCHFREE
#define RT_BVH_BRANCHFREE 0        // 1: push / pop as single predicated shared-memory instructions (measured slower: 795 vs 809 Msamples/s on 1M spheres)
#endif

#if RT_BVH_PATHSTAT
__device__ __forceinline__ unsigned long long globaltimer_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#endif

// RT_BVH_COOP / RT_BVH_COOP_AFTER / RT_BVH_COOP_CAP: rt_types.h (the host passes them in RtRenderArgs; RT_FLAG_TEST_STRAGGLERS overrides them)
#ifndef RT_BVH_COOP_WIDE
#define RT_BVH_COOP_WIDE 1         // nodes a lane takes from the work list per round (bvh_coop_query).  2 (the loads of both in flight
                                   // together) was measured on the 1M-sphere frame on 8 GPUs: 94.3 vs 87.8 ms -- most queries of a frame's
                                   // tail are short, and for them the second slot is only more code
#endif

// One ray, 32 lanes (bvh_coop_query).  Why: traversal cost is heavy tailed beyond anything a schedule can absorb -- 30 of the
// 5.3e8 paths of the 1M-sphere frame visit more than 20 000 nodes, the worst 301 812 in four segments (rays that leave the far
// ground and skim the whole sphere field inside the conservative padding of its boxes), and such a path, alone in its warp at
// one dependent L2 round trip per node, runs for 389 ms of a 620 ms frame (profiles/r02_pathstat.log); on 8 GPUs the frame
// ends when the unluckiest GPU's last such path does.
// How: the warp's 32 stack columns become ONE work list of node indices.  Every round the top min(32, count) nodes are taken,
// one per lane; a lane tests both children of its node (the same conservative box test, the same leaf test as the sequential
// traversal), the closest hit so far is merged over the warp as the (t, index) minimum -- the closest hit IS that minimum
// over all spheres the reference test accepts, so the order in which candidates are met cannot change it, and a box is only
// ever culled against a bound that is no smaller than the final one -- and the surviving inner children are appended by
// ballot + popc prefix, far child below near child.  The query restarts from the root with the lane's best hit so far as
// the bound (what the lane had on its own stack is dropped: the list overwrites the columns).  A list that would overflow
// gives up (returns false): the caller restarts the sequential traversal from the root, again with the best so far.
struct BvhCoopResult { float t; int i; uint32_t nodes, leaves; bool fits; };
// (everything by value: a reference to the kernel's long-lived state would pin that state to local memory)
__device__ __noinline__ BvhCoopResult bvh_coop_query(const RtDeviceScene &sc, uint32_t *list, uint32_t CAP, float ox, float oy, float oz, float dx,
                                                     float dy, float dz, float a, float ix, float iy, float iz, float best_t, int best_i)
{
    BvhCounters cnt = {0u, 0u};
    // CAP: entries of the work list; entry i lives at list[(i >> 5) * RT_THREADS + (i & 31)]
    const uint32_t lane = threadIdx.x & 31u, below = (1u << lane) - 1u;
    float bt = best_t; int bi = best_i;
    uint32_t count = 1;
    if (lane == 0) list[0] = sc.bvh_root;
    __syncwarp();
    bool fits = true;
    while (count) {
        // RT_BVH_COOP_WIDE nodes per lane and round (the loads of all of them are in flight together)
        constexpr int K = RT_BVH_COOP_WIDE;
        const uint32_t w = min(count, 32u * K);
        uint32_t node[K];
        bool mine[K];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            mine[k] = lane + 32u * k < w;
            node[k] = 0u;
            if (mine[k]) { const uint32_t i = count - 1u - lane - 32u * k; node[k] = list[(i >> 5) * RT_THREADS + (i & 31u)]; }
        }
        count -= w;
        __syncwarp();                                       // every lane has its nodes before anyone appends
        float4 n0[K], n1[K], n2[K], n3[K];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const float4 *np = sc.bvh_nodes + 4 * (size_t)node[k];          // (node 0 for an empty slot: a valid address, never used)
            n0[k] = __ldg(np); n1[k] = __ldg(np + 1); n2[k] = __ldg(np + 2); n3[k] = __ldg(np + 3);
        }
        float lt = bt; int li = bi;
        bool pl[K], pr[K], left_first[K];
        uint2 ch[K];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            pl[k] = false; pr[k] = false; left_first[k] = true;
            ch[k] = make_uint2(__float_as_uint(n3[k].x), __float_as_uint(n3[k].y));
            if (mine[k]) {
                ++cnt.nodes;
                float tl, tr;
                const bool hl = RT_BVH_BOX_L(n0[k], n1[k], n2[k], n3[k], ox, oy, oz, ix, iy, iz, sc.bvh_pad0, lt, tl);
                const bool hr = RT_BVH_BOX_R(n0[k], n1[k], n2[k], n3[k], ox, oy, oz, ix, iy, iz, sc.bvh_pad0, lt, tr);
                if (hl && (ch[k].x >> 31)) RT_BVH_LEAF_L(n0[k], n1[k], n2[k], ch[k].x, ox, oy, oz, dx, dy, dz, a, lt, li, cnt);
                if (hr && (ch[k].y >> 31)) RT_BVH_LEAF_R(n0[k], n1[k], n2[k], ch[k].y, ox, oy, oz, dx, dy, dz, a, lt, li, cnt);
                pl[k] = hl && !(ch[k].x >> 31); pr[k] = hr && !(ch[k].y >> 31);
                left_first[k] = tl <= tr;
            }
        }
        if (__any_sync(RT_FULL, lt < bt || li != bi)) {      // (t, index) minimum over the warp; t > 0, so its bits order like t
            const uint32_t tb = __float_as_uint(lt);
            const uint32_t tmin = __reduce_min_sync(RT_FULL, tb);
            bi = (int)__reduce_min_sync(RT_FULL, tb == tmin ? (uint32_t)li : 0xffffffffu);
            bt = __uint_as_float(tmin);
        }
        // append the surviving inner children: per lane slot K-1 first ... slot 0 last (on top), far child below near child
        uint32_t add = 0, at = count;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const uint32_t bl = __ballot_sync(RT_FULL, pl[k]), br = __ballot_sync(RT_FULL, pr[k]);
            add += (uint32_t)(__popc(bl) + __popc(br));
            at += (uint32_t)(__popc(bl & below) + __popc(br & below));
        }
        if (count + add > CAP) { fits = false; break; }
#pragma unroll
        for (int k = K - 1; k >= 0; --k) {
            if (pl[k] && pr[k]) {
                list[(at >> 5) * RT_THREADS + (at & 31u)] = left_first[k] ? ch[k].y : ch[k].x;
                ++at;
                list[(at >> 5) * RT_THREADS + (at & 31u)] = left_first[k] ? ch[k].x : ch[k].y;
                ++at;
            } else if (pl[k] || pr[k]) {
                list[(at >> 5) * RT_THREADS + (at & 31u)] = pl[k] ? ch[k].x : ch[k].y;
                ++at;
            }
        }
        count += add;
        __syncwarp();
    }
    BvhCoopResult r;
    r.t = bt; r.i = bi; r.nodes = cnt.nodes; r.leaves = cnt.leaves; r.fits = fits;
    return r;
}

struct BvhTrav {
    uint32_t node;
    int sp;
    float best_t; int best_i;
    float a, ix, iy, iz;
    bool active;
};

#if RT_BVH_COOP
// per-warp statistics into the launch's counters (the tail of both k_render_bvh and bvh_straggler_loop)
__device__ __forceinline__ void bvh_flush_stats(const RtRenderArgs &A, const ThreadStats &ts)
{
    uint32_t v[7] = {ts.samples, ts.segments, ts.sky, ts.absorbed, ts.capped, ts.bvh.nodes, ts.bvh.leaves};
#pragma unroll
    for (int k = 0; k < 7; ++k) {
        uint32_t x = v[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(RT_FULL, x, o);
        if ((threadIdx.x & 31u) == 0 && x) atomicAdd(A.stats + k, (unsigned long long)x);
    }
}

// Straggler loop: the rest of a warp's paths once its work pool is dry and only a few long traversals are left -- every
// closest-hit query by the whole warp.  Out of line and entered once, at the end of the kernel, with the warp's state BY VALUE, so
// that nothing it needs lengthens a live range in the traversal loop of k_render_bvh (as an inline block it put 6 local-memory
// loads into every node step).
template <bool QUEUE>
__device__ __noinline__ void bvh_straggler_loop(const RtRenderArgs &A, uint32_t *list, PathSlot S, BvhTrav T, WarpPool pool, ThreadStats ts)
{
    const RtDeviceScene &sc = A.sc;
    const uint32_t lane = threadIdx.x & 31u;
    // entries of the work list: all of the warp's stack columns; RT_FLAG_TEST_STRAGGLERS gives the odd warps of a CTA 8 (every query overflows)
    const uint32_t cap = min(((threadIdx.x >> 5) & 1u) ? A.coop_cap_odd : (uint32_t)RT_BVH_COOP_CAP, (uint32_t)(RT_BVH_SMEM_DEPTH * 32u));
    for (;;) {
        for (uint32_t m = __ballot_sync(RT_FULL, T.active); m; m &= m - 1u) {
            const int src = __ffs(m) - 1;
            const float ox = __shfl_sync(RT_FULL, S.ox, src), oy = __shfl_sync(RT_FULL, S.oy, src), oz = __shfl_sync(RT_FULL, S.oz, src);
            const float dx = __shfl_sync(RT_FULL, S.dx, src), dy = __shfl_sync(RT_FULL, S.dy, src), dz = __shfl_sync(RT_FULL, S.dz, src);
            const BvhCoopResult r = bvh_coop_query(sc, list, cap, ox, oy, oz, dx, dy, dz, __shfl_sync(RT_FULL, T.a, src),
                                                   __shfl_sync(RT_FULL, T.ix, src), __shfl_sync(RT_FULL, T.iy, src),
                                                   __shfl_sync(RT_FULL, T.iz, src), __shfl_sync(RT_FULL, T.best_t, src),
                                                   __shfl_sync(RT_FULL, T.best_i, src));
            ts.bvh.nodes += r.nodes; ts.bvh.leaves += r.leaves;
            if ((int)lane == src) {
                float bt = r.t; int bi = r.i;
                if (!r.fits) {
                    // a work list that overflowed: this lane runs the plain traversal to the end, with the best hit so far as its bound
                    BvhCounters c2 = {0u, 0u};
                    closest_hit_lbvh(sc, true, ox, oy, oz, dx, dy, dz, bt, bi, c2, true);
                    ts.bvh.nodes += c2.nodes; ts.bvh.leaves += c2.leaves;
                }
                T.best_t = bt; T.best_i = bi; T.active = false;
            }
            __syncwarp();
        }
        shade_slot<true, QUEUE, RT_PAR_DEFAULT, RT_BVH_BATCH>(A, nullptr, S, T.best_t, T.best_i, pool, ts, false);
        T.best_t = FLT_MAX; T.best_i = -1;
        if (S.st == ST_ACTIVE) {
            T.a = ray_a(S.dx, S.dy, S.dz);
            if (T.a > 0.f && T.a < 3.0e38f) {       // d = 0 or non-finite: the reference test finds nothing
                T.ix = 1.f / S.dx; T.iy = 1.f / S.dy; T.iz = 1.f / S.dz;
                T.active = true;
            }
        }
        if (!__any_sync(RT_FULL, S.st != ST_EMPTY)) break;
    }
    bvh_flush_stats(A, ts);
}
#endif

template <bool QUEUE>
__global__ void __launch_bounds__(RT_THREADS, RT_BVH_CTAS) k_render_bvh(const __grid_constant__ RtRenderArgs A)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint32_t *s_stack = reinterpret_cast<uint32_t *>(smem_raw);           // [RT_BVH_SMEM_DEPTH][RT_THREADS]
    const RtDeviceScene &sc = A.sc;
    const uint32_t tid = threadIdx.x;
#if RT_BVH_BRANCHFREE
    const uint32_t s_mine = smem_u32(s_stack) + tid * 4u;                  // this thread's column of the stack (shared-window address)
#endif

    PathSlot S;
    S.st = ST_EMPTY; S.pix = 0; S.smp = 0; S.bounce = 0;
    S.ox = S.oy = S.oz = 0.f; S.dx = S.dy = S.dz = 0.f;
    S.ar = S.ag = S.ab = 1.f;
    WarpPool pool; pool.next = 0; pool.end = 0; pool.exhausted = false;
    ThreadStats ts = {0, 0, 0, 0, 0, {0, 0}};
    BvhTrav T;
    T.node = 0; T.sp = 0; T.best_t = FLT_MAX; T.best_i = -1; T.a = 0.f; T.ix = T.iy = T.iz = 0.f; T.active = false;
    uint32_t ovf[RT_BVH_OVF_DEPTH];

#if RT_BVH_PATHSTAT
    // measurement build: report every path that visited more than RT_BVH_PATHSTAT nodes (pixel, sample, node visits, bounces,
    // microseconds from its first to its last segment, end time) -- the heavy tail behind the multi-GPU frame end
    uint32_t ps_id = 0xffffffffu, ps_nodes0 = 0, ps_bounce = 0;
    unsigned long long ps_t0 = 0;
    if (blockIdx.x == 0 && tid == 0) printf("PATHSTAT start %llu\n", globaltimer_ns() / 1000ull);
#endif
#if RT_BVH_COOP
    bool tail = false;                     // warp-uniform: the work pool is dry and only a few long traversals are left
#endif
    for (;;) {
        // ---- shade the lanes whose query finished, hand out new work, start their next query
        const bool hold = T.active;
        shade_slot<true, QUEUE, RT_PAR_DEFAULT, RT_BVH_BATCH>(A, nullptr, S, T.best_t, T.best_i, pool, ts, hold);
#if RT_BVH_PATHSTAT
        {
            const uint32_t id = S.st == ST_EMPTY ? 0xffffffffu : S.pix * 1024u + S.smp;
            if (id != ps_id) {
                const unsigned long long now = globaltimer_ns();
                if (ps_id != 0xffffffffu && ts.bvh.nodes - ps_nodes0 > (uint32_t)RT_BVH_PATHSTAT)
                    printf("PATHSTAT path pix %u smp %u nodes %u bounces %u us %llu end %llu\n", ps_id >> 10, ps_id & 1023u,
                           ts.bvh.nodes - ps_nodes0, ps_bounce, (now - ps_t0) / 1000ull, now / 1000ull);
                ps_id = id; ps_nodes0 = ts.bvh.nodes; ps_t0 = now;
            }
            ps_bounce = S.bounce;
        }
#endif
        if (!hold) {
            T.best_t = FLT_MAX; T.best_i = -1;
            if (S.st == ST_ACTIVE) {
                T.a = ray_a(S.dx, S.dy, S.dz);
                if (T.a > 0.f && T.a < 3.0e38f) {       // d = 0 or non-finite: the reference test finds nothing
                    T.ix = 1.f / S.dx; T.iy = 1.f / S.dy; T.iz = 1.f / S.dz;
                    T.node = sc.bvh_root; T.sp = 0; T.active = true;
                }
            }
        }
        if (!__any_sync(RT_FULL, S.st != ST_EMPTY)) break;
#if RT_BVH_COOP
        if (tail) break;
#endif

        // ---- step all running traversals, one node per iteration
        // lanes that could do something else once their query ends: a live path, or a free slot while work remains
        const uint32_t refill = __ballot_sync(RT_FULL, S.st != ST_EMPTY || !pool.exhausted);
#if RT_BVH_COOP
        uint32_t lone = 0;                 // consecutive loop iterations with a dry pool and at most RT_BVH_COOP lanes traversing
#endif
        for (;;) {
            const uint32_t act = __ballot_sync(RT_FULL, T.active);
            if (act == 0u || __popc(~act & refill) >= RT_BVH_IDLE) break;
#if RT_BVH_COOP
            if (pool.exhausted && (uint32_t)__popc(act) <= A.coop_lanes) {             // (RT_BVH_COOP / RT_BVH_COOP_AFTER unless a test overrides them)
                if (++lone >= A.coop_after) { tail = true; break; }                    // -> straggler loop below
            } else lone = 0;
#endif
#pragma unroll
            for (int step = 0; step < RT_BVH_STEPS; ++step) {
                if (!T.active) continue;
                const float4 *np = sc.bvh_nodes + 4 * (size_t)T.node;
                const float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2);
#if RT_BVH_CENTER
                const float4 n3 = __ldg(np + 3);
                const uint2 ch = make_uint2(__float_as_uint(n3.x), __float_as_uint(n3.y));
#else
                const float4 n3 = make_float4(0.f, 0.f, 0.f, 0.f);
                const uint2 ch = __ldg(reinterpret_cast<const uint2 *>(np + 3));
#endif
                ++ts.bvh.nodes;
                float tl, tr;
                bool hl = RT_BVH_BOX_L(n0, n1, n2, n3, S.ox, S.oy, S.oz, T.ix, T.iy, T.iz, sc.bvh_pad0, T.best_t, tl);
                bool hr = RT_BVH_BOX_R(n0, n1, n2, n3, S.ox, S.oy, S.oz, T.ix, T.iy, T.iz, sc.bvh_pad0, T.best_t, tr);
                const bool ll = hl && (ch.x >> 31), lr = hr && (ch.y >> 31);
                if (ll || lr) {                                   // one divergent region for both leaf children
                    if (ll) RT_BVH_LEAF_L(n0, n1, n2, ch.x, S.ox, S.oy, S.oz, S.dx, S.dy, S.dz, T.a, T.best_t, T.best_i, ts.bvh);
                    if (lr) RT_BVH_LEAF_R(n0, n1, n2, ch.y, S.ox, S.oy, S.oz, S.dx, S.dy, S.dz, T.a, T.best_t, T.best_i, ts.bvh);
                }
                hl = hl && !(ch.x >> 31); hr = hr && !(ch.y >> 31);
#if RT_BVH_BRANCHFREE
                // next node without branches: near inner child, far one on the stack; nothing to enter -> pop.
                // Push and pop are single predicated shared-memory instructions; only depths beyond the shared-memory
                // stack (never reached on the test scenes) take a branch, to the local-memory overflow.
                const bool left_first = tl <= tr;
                const bool go_left = hl && (!hr || left_first);
                const bool push = hl && hr, enter = hl || hr;
                const uint32_t far = left_first ? ch.y : ch.x;
                if (push) {
                    if (T.sp >= RT_BVH_SMEM_DEPTH) ovf[min(T.sp - RT_BVH_SMEM_DEPTH, RT_BVH_OVF_DEPTH - 1)] = far;
                }
                asm volatile("{\n.reg .pred p;\nsetp.ne.u32 p, %2, 0;\n@p st.shared.u32 [%0], %1;\n}"
                             :: "r"(s_mine + (uint32_t)T.sp * (RT_THREADS * 4u)), "r"(far), "r"((uint32_t)(push && T.sp < RT_BVH_SMEM_DEPTH)) : "memory");
                T.sp += push ? 1 : 0;
                const bool pop = !enter && T.sp > 0;
                T.sp -= pop ? 1 : 0;
                uint32_t popped = 0u;
                asm volatile("{\n.reg .pred p;\nsetp.ne.u32 p, %2, 0;\n@p ld.shared.u32 %0, [%1];\n}"
                             : "+r"(popped) : "r"(s_mine + (uint32_t)T.sp * (RT_THREADS * 4u)), "r"((uint32_t)(pop && T.sp < RT_BVH_SMEM_DEPTH)) : "memory");
                if (pop) {
                    if (T.sp >= RT_BVH_SMEM_DEPTH) popped = ovf[min(T.sp - RT_BVH_SMEM_DEPTH, RT_BVH_OVF_DEPTH - 1)];
                }
                T.node = enter ? (go_left ? ch.x : ch.y) : popped;
                T.active = enter || pop;
#else
                // next node by selects: near inner child, far one on the stack; nothing to enter -> pop
                const bool left_first = tl <= tr;
                const bool go_left = hl && (!hr || left_first);
                if (hl && hr) {
                    const uint32_t far = left_first ? ch.y : ch.x;
                    if (T.sp < RT_BVH_SMEM_DEPTH) s_stack[T.sp * RT_THREADS + tid] = far;
                    else ovf[min(T.sp - RT_BVH_SMEM_DEPTH, RT_BVH_OVF_DEPTH - 1)] = far;
                    ++T.sp;
                }
                if (hl || hr) T.node = go_left ? ch.x : ch.y;
                else if (T.sp == 0) T.active = false;
                else {
                    --T.sp;
                    T.node = (T.sp < RT_BVH_SMEM_DEPTH) ? s_stack[T.sp * RT_THREADS + tid]
                                                         : ovf[min(T.sp - RT_BVH_SMEM_DEPTH, RT_BVH_OVF_DEPTH - 1)];
                }
#endif
            }
        }
    }

#if RT_BVH_COOP
    if (tail) {                            // what is left of this warp's paths: bvh_straggler_loop (it also adds up the statistics)
        bvh_straggler_loop<QUEUE>(A, s_stack + (tid & ~31u), S, T, pool, ts);
        return;
    }
#endif
#if RT_BVH_PATHSTAT
    if (blockIdx.x == 0 && tid == 0) printf("PATHSTAT first_warp_end %llu\n", globaltimer_ns() / 1000ull);
#endif
    uint32_t v[7] = {ts.samples, ts.segments, ts.sky, ts.absorbed, ts.capped, ts.bvh.nodes, ts.bvh.leaves};
#pragma unroll
    for (int k = 0; k < 7; ++k) {
        uint32_t x = v[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(RT_FULL, x, o);
        if ((threadIdx.x & 31u) == 0 && x) atomicAdd(A.stats + k, (unsigned long long)x);
    }
}
