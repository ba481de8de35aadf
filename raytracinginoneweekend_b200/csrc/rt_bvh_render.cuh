// rt_bvh_render.cuh -- persistent bounce loop over the GPU-built LBVH with a RESUMABLE traversal.
// Included by rt_device.cuh in both arithmetic builds.
//
// Traversal lengths are heavy tailed (sky rays leave after a handful of nodes, rays that skim the sphere field visit
// hundreds), so "every lane traverses its ray to the end, then the warp shades" keeps 6 of 32 lanes busy (ncu,
// profiles/r01_ncu_k_render_lbvh.json).  Here the traversal state (node, stack, running closest hit) lives across
// passes: the warp steps all traversals one node at a time and leaves the loop as soon as RT_BVH_IDLE lanes have
// finished; those lanes are shaded / regenerated (the others "hold"), start their next traversal, and the loop resumes.
// The stack lives in shared memory ([depth][thread], conflict free); depths beyond RT_BVH_SMEM_DEPTH (never seen on
// Morton trees of the test scenes) go to a local-memory overflow so the result never depends on the stack size.
// Semantics are those of closest_hit_lbvh (conservative boxes, reference test at the leaves, (t, index) order).
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

struct BvhTrav {
    uint32_t node;
    int sp;
    float best_t; int best_i;
    float a, ix, iy, iz;
    bool active;
};

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

        // ---- step all running traversals, one node per iteration
        // lanes that could do something else once their query ends: a live path, or a free slot while work remains
        const uint32_t refill = __ballot_sync(RT_FULL, S.st != ST_EMPTY || !pool.exhausted);
        for (;;) {
            const uint32_t act = __ballot_sync(RT_FULL, T.active);
            if (act == 0u || __popc(~act & refill) >= RT_BVH_IDLE) break;
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
