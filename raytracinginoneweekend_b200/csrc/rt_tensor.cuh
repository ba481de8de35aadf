// rt_tensor.cuh -- brute-force closest hit with the candidate filter on the 5th-generation tensor cores.
// Included by rt_device.cuh in BOTH translation units: rtf::k_render_tc is the timed kernel (throughput shading),
// rtx::k_render_tc the same kernel with the parity shading (RT_FLAG_EXACT; renders bit-identical to the oracle).
//
// What is computed.  For a ray (o, u = d/|d|) and a sphere (c, r) the sign of the discriminant of
// src/raytracer.hxx:55-61 is the sign of
//     q = (u.(c-o))^2 - |c-o|^2 + r^2
//       = sum_ij u_i u_j c_i c_j  +  p.c  +  (r^2 - |c|^2)  +  k,      p = 2(o - (u.o)u),  k = -|o - (u.o)u|^2,
// a BILINEAR form: (ray features) x (sphere features).  One pass of 128 rays over N spheres is therefore a
// [128 x K] x [K x N] matrix product.  It runs as tcgen05.mma kind::tf32 with FP32-accurate inputs through 3-way
// split products (a_h f_h + a_h f_l + a_l f_h, |a_l| <= 2^-11 |a|): K = 9 terms x 3 + 3 (r^2-|c|^2 in three
// pieces) + 2 (k in two pieces) = 32.  Accumulators live in TMEM; the ray rows are the A operand, written to
// TMEM by the thread that owns the ray (tcgen05.st), the sphere matrix is the B operand in shared memory
// (canonical K-major no-swizzle layout, staged once per CTA by TMA).
//
// The product is only a CONSERVATIVE FILTER: the sphere is inflated by eps (|c|^2 + r^2 + |o|^2), eps = 2^-16 (measured
// error of q: < 2^-20.4 of that scale, tools/proto_tc.cu; the float evaluation of the reference test itself is off by at
// most ~2^-21 of it), so no sphere the reference test accepts is dropped.  The hit that is returned is decided by the
// REFERENCE TEST in IEEE arithmetic (sphere_candidate, rt_device.cuh: the same function the parity kernels run) on the few
// candidates, in increasing sphere index -- index and t are the oracle's bit for bit (RT_FLAG_FILTER_CHECK re-runs that
// test over ALL spheres and counts disagreements; k_hits_tc exposes the same closest hit to rt_primary_hits /
// rt_closest_hits so the parity tests exercise exactly this code).
//
// Per SM: ONE CTA = RT_TC_WGS ray warpgroups (128 rays each, one ray per thread = one TMEM lane) + whole warpgroups of
// MMA-issuer warps (one issuer warp per ray warpgroup; they give their registers to the ray warps: setmaxnreg).
// TMEM (512 columns): per warpgroup 32 columns of ray rows (A) + RT_TC_NBUF accumulator tiles (D) of RT_TC_CHUNK columns.
// Handshakes are mbarriers only: a_ready (ray rows stored) -> issuer; full[b] (tcgen05.commit) -> ray warps;
// empty[b] (accumulators copied to registers) -> issuer.  A tile is 64 spheres = 4 MMAs (M128 N64 K8).
// Epilogue: tcgen05.ld 2 x 32 columns -> 64 registers -> two sign-mask words (16 I2IP.U2 + 1 LOP3 each) -> the non-empty
// mask words go to shared memory -> pass 2 walks them.
// Shapes and handshake variants that were measured and dropped (self-issuing ray warps, vote-free termination, nanosleep
// back-off, path-state stash, inlined pass 2, paired candidates, I2IP.U8 / FMA-pipe / funnel-shift masks) are recorded in
// DESIGN.md section 4 with their numbers; the code keeps only the shipped variant.
#pragma once

// RT_TC_WGS / RT_TC_CHUNK / RT_TC_NBUF (rt_types.h): ray warpgroups per CTA, spheres per accumulator tile, tiles per warpgroup
#define RT_TC_RAY_THREADS (RT_TC_WGS * 128)
#define RT_TC_ISSUER_THREADS (((RT_TC_WGS + 3) / 4) * 128)   // whole warpgroups (setmaxnreg is per warpgroup), one issuer warp per ray warpgroup
#define RT_TC_THREADS (RT_TC_RAY_THREADS + RT_TC_ISSUER_THREADS)
// register hand-over (the pool is the CTA's own allocation: what the issuers free must cover what the ray warps take)
//   4 ray warpgroups: 640 threads x 96 at launch; issuers 24, ray warps 112  (128 x 72 freed >= 512 x 16 taken)
//   5 ray warpgroups: 896 threads x 72 at launch; issuers 24, ray warps 88   (256 x 48 freed >= 640 x 16 taken)
#if RT_TC_WGS == 4
#define RT_TC_REG_RAY "112"
#elif RT_TC_WGS == 5
#define RT_TC_REG_RAY "88"
#endif
#define RT_TC_KQ 8                        // K = 32 tf32 slots = 8 chunks of 16 bytes
#define RT_TC_WGCOLS (32 + RT_TC_NBUF * RT_TC_CHUNK)   // TMEM columns per warpgroup
#define RT_TC_WGBARS (1 + 2 * RT_TC_NBUF)              // a_ready, full[NBUF], empty[NBUF]
static_assert(RT_TC_WGS * RT_TC_WGCOLS <= 512, "TMEM has 512 columns");
static_assert(RT_TC_CHUNK % 64 == 0 && RT_TC_CHUNK >= 64 && RT_TC_CHUNK <= 256, "tile width: pairs of mask words, one MMA N");
static_assert((1 + RT_TC_WGS * RT_TC_WGBARS) * 8 <= 448 && 452 + 4 * RT_TC_WGS <= 512, "barriers, TMEM base and done flags live in the first 512 bytes of shared memory");


namespace tc {

__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes)
{
    // K-major, SWIZZLE_NONE: core matrix = 8 rows x 16 B contiguous; LBO = distance between the two 16-byte
    // k-chunks of one K=8 step, SBO = distance between 8-row groups; bit 46 = descriptor version (sm_100)
    return (uint64_t)((saddr >> 4) & 0x3fffu) | ((uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32) | (1ull << 46);
}
// D[tmem] (+)= A[tmem] * B[smem]^T, M = 128, N = RT_TC_CHUNK, K = 8 (tf32).  Called by a CONVERGED warp with
// warp-uniform operands; one elected lane issues (keeps the operands in uniform registers: no per-issue election loop)
__device__ __forceinline__ void mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n.reg .pred p, pe;\nsetp.ne.b32 p, %4, 0;\nelect.sync _|pe, 0xffffffff;\n"
                 "@pe tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n}"
                 ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void mma_commit(uint64_t *bar)
{
    asm volatile("{\n.reg .pred pe;\nelect.sync _|pe, 0xffffffff;\n"
                 "@pe tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n}" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void bar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// ray-warp wait for an accumulator tile (short: the MMAs of the tile are usually in flight already).
// RT_TC_RAY_HINT > 0 adds that suspend-time hint (ns) to the try_wait; measured against the bare spin in DESIGN section 4.
__device__ __forceinline__ void bar_wait(uint64_t *bar, uint32_t parity)
{
#if RT_TC_RAY_HINT
    asm volatile(
        "{\n.reg .pred P1;\nTC_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, %2;\n"
        "@P1 bra TC_DONE;\nbra TC_WAIT;\nTC_DONE:\n}" ::"r"(smem_u32(bar)), "r"(parity), "r"((uint32_t)RT_TC_RAY_HINT) : "memory");
#else
    asm volatile(
        "{\n.reg .pred P1;\nTC_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra TC_DONE;\nbra TC_WAIT;\nTC_DONE:\n}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
#endif
}
// issuer-side wait: may sleep long (suspend-time hint), so the idle issuer warps do not take issue slots
__device__ __forceinline__ void bar_wait_sleep(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n.reg .pred P1;\nTC_WAITS:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, 0x989680;\n"
        "@P1 bra TC_DONES;\nbra TC_WAITS;\nTC_DONES:\n}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// warpgroup-wide OR over a named barrier (ids 1..RT_TC_WGS)
__device__ __forceinline__ bool wg_any(uint32_t id, bool pred)
{
    uint32_t r;
    asm volatile("{\n.reg .pred p, q;\nsetp.ne.u32 q, %1, 0;\nbar.red.or.pred p, %2, 128, q;\nselp.u32 %0, 1, 0, p;\n}"
                 : "=r"(r) : "r"((uint32_t)pred), "r"(id) : "memory");
    return r != 0u;
}
__device__ __forceinline__ float tf32_round(float x)
{
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}

#define RT_TC_LD32(v, addr)                                                                                               \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                                \
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27," \
                 "%28,%29,%30,%31}, [%32];"                                                                               \
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),        \
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),  \
                   "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]),             \
                   "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]),             \
                   "=r"(v[30]), "=r"(v[31])                                                                               \
                 : "r"(addr))
#define RT_TC_ST32(addr, v)                                                                                               \
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "                                                          \
                 "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28," \
                 "%29,%30,%31,%32};"                                                                                      \
                 ::"r"(addr), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]),     \
                   "f"(v[8]), "f"(v[9]), "f"(v[10]), "f"(v[11]), "f"(v[12]), "f"(v[13]), "f"(v[14]), "f"(v[15]),          \
                   "f"(v[16]), "f"(v[17]), "f"(v[18]), "f"(v[19]), "f"(v[20]), "f"(v[21]), "f"(v[22]), "f"(v[23]),        \
                   "f"(v[24]), "f"(v[25]), "f"(v[26]), "f"(v[27]), "f"(v[28]), "f"(v[29]), "f"(v[30]), "f"(v[31])         \
                 : "memory")

// Sign masks of 32 accumulator words -> one 32-bit candidate mask whose __clz order is the sphere order.
// I2IP with 2-bit saturation (cvt.pack.sat.u2.s32.b32): a positive s32 (= a positive float) packs to 0b11, anything else to
// 0b00, and the instruction shifts its third operand left by 4 -- a chain of 8 holds 16 values as bit pairs; two chains merge
// with ONE LOP3 (odd bits of the first, even bits of the second): 17 instructions per word.  bit 31 - 2k <=> word k,
// bit 30 - 2k <=> word 16 + k; the host places sphere 32 b + s in column 32 b + (s even ? s/2 : 16 + (s-1)/2) (rt_api.cu).
__device__ __forceinline__ uint32_t sign_mask(const uint32_t (&v)[32])
{
    uint32_t c0 = 0u, c1 = 0u;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        asm("cvt.pack.sat.u2.s32.b32 %0, %1, %2, %3;" : "=r"(c0) : "r"(v[2 * i]), "r"(v[2 * i + 1]), "r"(c0));
        asm("cvt.pack.sat.u2.s32.b32 %0, %1, %2, %3;" : "=r"(c1) : "r"(v[16 + 2 * i]), "r"(v[17 + 2 * i]), "r"(c1));
    }
    return (c0 & 0xaaaaaaaau) | (c1 & 0x55555555u);
}

// Ray -> the 32 A-operand slots (slot order must match the sphere rows built on the host, rt_api.cu)
__device__ __forceinline__ void ray_row(const RtDeviceScene &sc, bool live, float ox, float oy, float oz,
                                        float dx, float dy, float dz, float (&row)[32])
{
    ox -= sc.tc_gx; oy -= sc.tc_gy; oz -= sc.tc_gz;
    float len2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
    float oo = fmaf(oz, oz, fmaf(oy, oy, ox * ox));
    bool ok = live && len2 > 0.f && len2 < 3.0e38f && oo < 1.0e30f;
    float inv = rsqrtf(len2);
    float ux = dx * inv, uy = dy * inv, uz = dz * inv;
    float uo = fmaf(uz, oz, fmaf(uy, oy, ux * ox));
    float qx = fmaf(-uo, ux, ox), qy = fmaf(-uo, uy, oy), qz = fmaf(-uo, uz, oz);
    float kk = fmaf(sc.tc_eps, oo, -fmaf(qz, qz, fmaf(qy, qy, qx * qx)));
    float co[9] = {ux * ux, uy * uy, uz * uz, 2.f * ux * uy, 2.f * ux * uz, 2.f * uy * uz, 2.f * qx, 2.f * qy, 2.f * qz};
    if (!ok) {                              // no candidates: d = 0 / non-finite ray (the reference test rejects it), idle slot
#pragma unroll
        for (int t = 0; t < 9; ++t) co[t] = 0.f;
        kk = -3.0e38f;
    }
#pragma unroll
    for (int t = 0; t < 9; ++t) {
        float h = tf32_round(co[t]);
        row[3 * t + 0] = h; row[3 * t + 1] = h; row[3 * t + 2] = co[t] - h;
    }
    row[27] = 1.f; row[28] = 1.f; row[29] = 1.f;
    float h = tf32_round(kk);
    row[30] = h; row[31] = kk - h;
}

// ---------------------------------------------------------------- CTA layout shared by the render and the hits kernel
// shared memory: [0,512) barriers + TMEM base + done flags | B matrix | sphere records | candidate mask words | (materials)
//   bars[0] staging; per warpgroup w, at bars + 1 + w * RT_TC_WGBARS: a_ready (4 arrivals), full[NBUF] (1), empty[NBUF] (4)
struct Cta {
    uint64_t *bars;
    volatile uint32_t *s_done;
    float4 *s_b, *s_sph;
    uint32_t *s_mask;
    uint32_t ns, npad, n, nch, tmem_base;
};

// barrier init, TMA staging of the sphere matrix and records, TMEM allocation; every thread of the CTA calls it
__device__ __forceinline__ Cta cta_setup(const RtDeviceScene &sc, unsigned char *smem_raw)
{
    Cta c;
    c.bars = reinterpret_cast<uint64_t *>(smem_raw);
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(smem_raw + 448);
    c.s_done = reinterpret_cast<volatile uint32_t *>(smem_raw + 452);
    c.ns = sc.tc_ns; c.npad = sc.npad; c.n = sc.n;
    c.nch = c.ns / RT_TC_CHUNK;
    c.s_b = reinterpret_cast<float4 *>(smem_raw + 512);
    c.s_sph = c.s_b + (size_t)c.ns * RT_TC_KQ;
    c.s_mask = reinterpret_cast<uint32_t *>(c.s_sph + c.npad);
    RT_DASSERT(c.ns % (RT_TC_CHUNK * RT_TC_NBUF) == 0u && c.ns >= c.n && c.ns <= RT_TC_MAX_SPHERES);
    if (threadIdx.x == 0) {
        for (int w = 0; w < RT_TC_WGS; ++w) {
            uint64_t *bw = c.bars + 1 + w * RT_TC_WGBARS;
            bar_init(bw, 4);
            for (int b = 0; b < RT_TC_NBUF; ++b) { bar_init(bw + 1 + b, 1); bar_init(bw + 1 + RT_TC_NBUF + b, 4); }
            c.s_done[w] = 0u;
        }
    }
    stage_spheres(c.s_sph, sc.sph, c.s_b, sc.tc_b, c.npad, c.ns * RT_TC_KQ, c.bars);      // inits bars[0], fences, TMA, wait
    if ((threadIdx.x >> 5) == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(s_tmem)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    fence_before();
    __syncthreads();
    fence_after();
    c.tmem_base = *s_tmem;
    RT_DASSERT((c.tmem_base & 0xffffu) == 0u);          // the whole 512-column allocation starts at column 0
    return c;
}
__device__ __forceinline__ void cta_teardown(const Cta &c)
{
    fence_before();
    __syncthreads();
    if ((threadIdx.x >> 5) == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(c.tmem_base));
}

// MMA issuer warp of ray warpgroup w: per pass, wait for the ray rows, then per tile wait for the accumulator buffer,
// issue its 4 MMAs and commit.  Ends when the warpgroup has flagged s_done and arrived once more on a_ready.
__device__ __forceinline__ void issuer_loop(const Cta &c, uint32_t w)
{
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(RT_TC_CHUNK >> 3) << 17) | ((128u >> 4) << 24);
    const uint64_t bdesc0 = smem_desc(smem_u32(c.s_b), c.ns * 16u, 128u);
    uint64_t *bar = c.bars + 1 + w * RT_TC_WGBARS;
    const uint32_t tmem_a = c.tmem_base + w * RT_TC_WGCOLS, tmem_d = tmem_a + 32;
    RT_DASSERT((tmem_d & 0xffffu) + RT_TC_NBUF * RT_TC_CHUNK <= 512u);
    uint32_t uses = 0;                                                          // tiles issued so far
    for (uint32_t pass = 0;; ++pass) {
        bar_wait_sleep(bar + 0, pass & 1u);                                     // ray rows are in TMEM
        if (c.s_done[w]) break;
        for (uint32_t t = 0; t < c.nch; ++t) {
            const uint32_t b = t % RT_TC_NBUF;                                  // nch is a multiple of NBUF: b = uses % NBUF too
            RT_DASSERT(b == uses % RT_TC_NBUF);
            if (uses >= RT_TC_NBUF)                                             // previous tile in this buffer consumed
                bar_wait_sleep(bar + 1 + RT_TC_NBUF + b, (uses / RT_TC_NBUF - 1u) & 1u);
            fence_after();
#pragma unroll
            for (uint32_t ks = 0; ks < RT_TC_KQ / 2; ++ks) {
                // K step = 2 k-chunks (2 * ns * 16 bytes apart in s_b); tile t starts at row RT_TC_CHUNK * t
                uint64_t bd = bdesc0 + (uint64_t)(((ks * 2u * c.ns + t * RT_TC_CHUNK) * 16u) >> 4);
                mma_tf32_ts(tmem_d + b * RT_TC_CHUNK, tmem_a + ks * 8u, bd, idesc, ks > 0u ? 1u : 0u);
            }
            mma_commit(bar + 1 + b);
            ++uses;
        }
    }
    __syncwarp();
}

// per-thread view of a ray warp
struct RayWarp {
    uint32_t wg, tid, lane;
    uint32_t tmem_a, tmem_d;
    uint64_t *bar;
    uint32_t full_parity;                   // bit b: parity of the next completion of full[b]
};
__device__ __forceinline__ RayWarp ray_warp_setup(const Cta &c)
{
    RayWarp r;
    const uint32_t warp = threadIdx.x >> 5;
    r.tid = threadIdx.x; r.lane = threadIdx.x & 31u; r.wg = warp >> 2;
    r.tmem_a = c.tmem_base + r.wg * RT_TC_WGCOLS + (((warp & 3u) * 32u) << 16);
    r.tmem_d = r.tmem_a + 32;
    r.bar = c.bars + 1 + r.wg * RT_TC_WGBARS;
    r.full_parity = 0;
    return r;
}
// the warpgroup has no ray left in any slot: tell the issuer and leave (called by all 128 threads after the vote)
__device__ __forceinline__ void ray_warp_finish(const Cta &c, const RayWarp &r)
{
    if ((r.tid & 127u) == 0u) c.s_done[r.wg] = 1u;
    __syncwarp();
    if (r.lane == 0) bar_arrive(r.bar + 0);
}

// THE closest hit of the tensor path: this thread's ray against all spheres.  Must be called by all four warps of the
// warpgroup in every pass (idle slots pass live = false).  Returns the oracle's hit: smallest t in (kMIN, FLT_MAX) of the
// IEEE reference test, ties to the lowest sphere index (src/raytracer.hxx:52-118).
__device__ __forceinline__ void closest_hit(const Cta &c, RayWarp &r, const RtDeviceScene &sc, bool live,
                                            float ox, float oy, float oz, float dx, float dy, float dz,
                                            float &best_t, int &best_i, bool check, unsigned long long *stats)
{
    {
        float row[32];
        ray_row(sc, live, ox, oy, oz, dx, dy, dz, row);
        RT_TC_ST32(r.tmem_a, row);
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        fence_before();
        __syncwarp();
        if (r.lane == 0) bar_arrive(r.bar + 0);
    }
    uint32_t nz = 0u;                       // bit (31 - w): mask word w is non-empty
    best_t = FLT_MAX; best_i = -1;
#ifndef RT_TC_RA_LATE
#define RT_TC_RA_LATE 0
#endif
    const float a = ray_a(dx, dy, dz);                      // src/raytracer.hxx:56, the reference's operation order
#if !RT_TC_RA_LATE
    const float ra = ray_ra(a);
#endif
#if RT_TC_UNROLL1
#pragma unroll 1
#endif
    for (uint32_t t = 0; t < c.nch; ++t) {
        const uint32_t b = t % RT_TC_NBUF;
        bar_wait(r.bar + 1 + b, (r.full_parity >> b) & 1u);
        r.full_parity ^= 1u << b;
        fence_after();
#pragma unroll
        for (uint32_t i = 0; i < RT_TC_CHUNK / 32; i += 2) {                   // 64 columns at a time
            const bool last = i + 2 >= RT_TC_CHUNK / 32;
            uint32_t v0[32], v1[32];
            RT_DASSERT(((r.tmem_d + b * RT_TC_CHUNK + i * 32u) & 0xffffu) + 64u <= 512u);
            RT_TC_LD32(v0, r.tmem_d + b * RT_TC_CHUNK + i * 32u);
            RT_TC_LD32(v1, r.tmem_d + b * RT_TC_CHUNK + i * 32u + 32u);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (last) {                      // the whole tile is in registers: hand the buffer back
                fence_before();
                __syncwarp();
                if (r.lane == 0) bar_arrive(r.bar + 1 + RT_TC_NBUF + b);
            }
            const uint32_t w0 = t * (RT_TC_CHUNK / 32) + i;
            const uint32_t m0 = sign_mask(v0);
            if (m0) { c.s_mask[w0 * RT_TC_RAY_THREADS + r.tid] = m0; nz |= 0x80000000u >> w0; }
            const uint32_t m1 = sign_mask(v1);
            if (m1) { c.s_mask[(w0 + 1u) * RT_TC_RAY_THREADS + r.tid] = m1; nz |= 0x80000000u >> (w0 + 1u); }
#if RT_TC_SKIPSTAT
            // measurement build (two-level filter question): how often does a WHOLE WARP find no candidate in a 64-sphere tile,
            // i.e. how often could a warp-collective tcgen05.ld + sign-mask pass be skipped?  counts land in the otherwise unused
            // node_visits (warp-tiles seen) and filter_misses (warp-tiles with no candidate in any lane) statistics
            {
                const bool none = !__any_sync(RT_FULL, live && (m0 | m1) != 0u);
                if (r.lane == 0) { atomicAdd(stats + 5, 1ull); if (none) atomicAdd(stats + 7, 1ull); }
            }
#endif
        }
    }
    // pass 2: the reference test on this lane's candidates, increasing sphere index, strict <
#if RT_TC_RA_LATE
    const float ra = ray_ra(a);
#endif
    uint32_t words = live ? nz : 0u, cand = 0u, w = 0u;
    for (;;) {
        if (cand == 0u) {
            if (words == 0u) break;
            w = (uint32_t)__clz(words);
            words &= ~(0x80000000u >> w);
            cand = c.s_mask[w * RT_TC_RAY_THREADS + r.tid];
        }
        const int lz = __clz(cand);
        cand &= ~(0x80000000u >> lz);
        const uint32_t j = (w << 5) + (uint32_t)lz;
        if (j < c.n) sphere_candidate(c.s_sph[j], (int)j, ox, oy, oz, dx, dy, dz, a, ra, best_t, best_i);
    }
    RT_DASSERT(best_i < (int)c.n);
    // self-check (RT_FLAG_FILTER_CHECK): the plain test over all spheres must agree.  A RUNTIME flag on purpose: with the
    // check compiled out of the timed instantiation (a template parameter) the same kernel ran at 1.16 instead of 1.71
    // Gsamples/s (run 5 of round 2, same box, same sources otherwise: profiles/r02_ab5.log) -- a code generation effect
    // of ptxas on this 4 000-instruction kernel, not of the check, which never executes in a timed run.
    if (check && live) {
        float ct = FLT_MAX; int ci = -1;
        for (uint32_t j = 0; j < c.n; ++j)
            sphere_candidate_ref(c.s_sph[j], (int)j, ox, oy, oz, dx, dy, dz, a, ct, ci);
        if (ci != best_i || ct != best_t) atomicAdd(stats + 7, 1ull);
    }
}

#if RT_TC_REGROUP
// ---------------------------------------------------------------- material regrouping (north_star (3); prototype, default off)
// After the closest hit the 128 paths of a warpgroup are SORTED by what their shading will do -- lambert / metal /
// dielectric hit, sky or depth cap, regeneration -- so that each warp shades (mostly) one kind.  With one path per lane a
// permutation inside a warp cannot reduce divergence, so the exchange goes through shared memory across the four warps:
// per-warp ballots give the key counts, one named barrier publishes them, every thread computes its destination from
// the 4 x 5 counts, writes its 15 words of state (13 of the path, t, hit index) there, a second barrier, reads its new
// state back.  Measured against the unsorted kernel in DESIGN.md section 4 (tools/build_variant.sh regroup -DRT_TC_REGROUP=1).
#define RT_TC_XCHG_WORDS 15
__device__ __forceinline__ void regroup(const RtRenderArgs &A, const RayWarp &r, uint32_t *s_cnt, uint32_t *s_x,
                                        const uint32_t *s_meta, PathSlot &S, float &best_t, int &best_i)
{
    const uint32_t w = (r.tid >> 5) & 3u, lane = r.lane, lt = (1u << lane) - 1u;
    uint32_t key;
    if (S.st != ST_ACTIVE) key = 4u;                                         // will regenerate
    else if (best_i < 0 || S.bounce + 1u >= A.depth) key = 3u;               // sky / depth cap: the short branches
    else key = (s_meta ? s_meta[best_i] : __ldg(A.sc.sph_meta + best_i)) >> 30;   // material kind of the hit
    uint32_t mine = 0u, cnt[5];
#pragma unroll
    for (uint32_t k = 0; k < 5u; ++k) {
        const uint32_t b = __ballot_sync(RT_FULL, key == k);
        cnt[k] = __popc(b);
        if (key == k) mine = __popc(b & lt);
    }
    uint32_t *wc = s_cnt + r.wg * 20u;                                       // [warp][key]
    if (lane < 5u) wc[w * 5u + lane] = cnt[lane];
    asm volatile("bar.sync %0, 128;" ::"r"(8u + r.wg) : "memory");
    uint32_t pos = mine;
#pragma unroll
    for (uint32_t k = 0; k < 5u; ++k) {
#pragma unroll
        for (uint32_t ww = 0; ww < 4u; ++ww) {
            const uint32_t v = wc[ww * 5u + k];
            if (k < key || (k == key && ww < w)) pos += v;
        }
    }
    uint32_t *x = s_x + r.wg * 128u;                                         // [word][RT_TC_RAY_THREADS]
#define RT_X(i) x[(i) * RT_TC_RAY_THREADS + pos]
    RT_X(0) = __float_as_uint(S.ox); RT_X(1) = __float_as_uint(S.oy); RT_X(2) = __float_as_uint(S.oz);
    RT_X(3) = __float_as_uint(S.dx); RT_X(4) = __float_as_uint(S.dy); RT_X(5) = __float_as_uint(S.dz);
    RT_X(6) = __float_as_uint(S.ar); RT_X(7) = __float_as_uint(S.ag); RT_X(8) = __float_as_uint(S.ab);
    RT_X(9) = S.pix; RT_X(10) = S.smp; RT_X(11) = S.bounce | ((uint32_t)S.st << 30);
    RT_X(12) = __float_as_uint(best_t); RT_X(13) = (uint32_t)best_i;
#undef RT_X
    asm volatile("bar.sync %0, 128;" ::"r"(8u + r.wg) : "memory");
    const uint32_t me = r.tid & 127u;
#define RT_Y(i) x[(i) * RT_TC_RAY_THREADS + me]
    S.ox = __uint_as_float(RT_Y(0)); S.oy = __uint_as_float(RT_Y(1)); S.oz = __uint_as_float(RT_Y(2));
    S.dx = __uint_as_float(RT_Y(3)); S.dy = __uint_as_float(RT_Y(4)); S.dz = __uint_as_float(RT_Y(5));
    S.ar = __uint_as_float(RT_Y(6)); S.ag = __uint_as_float(RT_Y(7)); S.ab = __uint_as_float(RT_Y(8));
    S.pix = RT_Y(9); S.smp = RT_Y(10);
    const uint32_t bs = RT_Y(11);
    S.bounce = bs & 0x3fffffffu; S.st = (int)(bs >> 30);
    best_t = __uint_as_float(RT_Y(12)); best_i = (int)RT_Y(13);
#undef RT_Y
}
#endif

}  // namespace tc

// ---------------------------------------------------------------- the persistent bounce loop on the tensor path
// Shading = that of the translation unit: rtf -- throughput shading (the timed default); rtx -- parity shading (RT_FLAG_EXACT).
// QUEUE: work from the device-side tile queue instead of the launch's tile list.
template <bool QUEUE>
__global__ void __launch_bounds__(RT_TC_THREADS, 1) k_render_tc(const __grid_constant__ RtRenderArgs A)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const tc::Cta c = tc::cta_setup(A.sc, smem_raw);
    // radii, material indices and materials next to the spheres when they fit (rt_render_tc_smem_bytes)
    const bool mats_in_smem = A.smem_mats != 0u;
    float4 *s_mat = reinterpret_cast<float4 *>(c.s_mask + (size_t)(c.ns / 32u) * RT_TC_RAY_THREADS);
    float *s_rad = reinterpret_cast<float *>(s_mat + A.sc.m);
    uint32_t *s_meta = reinterpret_cast<uint32_t *>(s_rad + c.npad);
    if (mats_in_smem) {
        for (uint32_t i = threadIdx.x; i < A.sc.m; i += RT_TC_THREADS) s_mat[i] = __ldg(A.sc.mat + i);
        for (uint32_t i = threadIdx.x; i < c.n; i += RT_TC_THREADS) { s_rad[i] = __ldg(A.sc.sph_rad + i); s_meta[i] = __ldg(A.sc.sph_meta + i); }
        __syncthreads();
    }
    const uint32_t warp = threadIdx.x >> 5;
    if (warp >= RT_TC_WGS * 4) {
        // ============================================================ MMA issuers: one warp per ray warpgroup
#ifdef RT_TC_REG_RAY
        asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");                        // hand registers to the ray warpgroups
#endif
        if (warp - RT_TC_WGS * 4 < RT_TC_WGS) tc::issuer_loop(c, warp - RT_TC_WGS * 4);
    } else {
        // ============================================================ ray warps: one path per thread
#ifdef RT_TC_REG_RAY
        asm volatile("setmaxnreg.inc.sync.aligned.u32 " RT_TC_REG_RAY ";");
#endif
        tc::RayWarp rw = tc::ray_warp_setup(c);
        PathSlot S;
        S.st = ST_EMPTY; S.pix = 0; S.smp = 0; S.bounce = 0;
        S.ox = S.oy = S.oz = 0.f; S.dx = S.dy = S.dz = 0.f;
        S.ar = S.ag = S.ab = 1.f;
        WarpPool pool; pool.next = 0; pool.end = 0; pool.exhausted = false;
        ThreadStats ts = {0, 0, 0, 0, 0, {0, 0}};
        float best_t = FLT_MAX; int best_i = -1;
#if RT_TC_REGROUP
        // exchange buffers of the regrouping prototype, behind everything else in the dynamic shared memory
        uint32_t *s_cnt = reinterpret_cast<uint32_t *>(mats_in_smem ? reinterpret_cast<unsigned char *>(s_meta + c.npad)
                                                                     : reinterpret_cast<unsigned char *>(s_mat));
        uint32_t *s_x = s_cnt + RT_TC_WGS * 20u;
#endif
        for (;;) {
#if RT_TC_REGROUP
            tc::regroup(A, rw, s_cnt, s_x, mats_in_smem ? s_meta : nullptr, S, best_t, best_i);
#endif
            shade_slot<false, QUEUE>(A, c.s_sph, S, best_t, best_i, pool, ts, false, mats_in_smem ? s_rad : nullptr,
                                   mats_in_smem ? s_meta : nullptr, mats_in_smem ? s_mat : nullptr);
            if (!tc::wg_any(1u + rw.wg, S.st != ST_EMPTY)) {                         // whole warpgroup out of work
                tc::ray_warp_finish(c, rw);
                break;
            }
            tc::closest_hit(c, rw, A.sc, S.st == ST_ACTIVE, S.ox, S.oy, S.oz, S.dx, S.dy, S.dz, best_t, best_i,
                            A.check != 0u, A.stats);
        }
        uint32_t v[5] = {ts.samples, ts.segments, ts.sky, ts.absorbed, ts.capped};
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            uint32_t x = v[k];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(RT_FULL, x, o);
            if (rw.lane == 0 && x) {
                atomicAdd(A.stats + k, (unsigned long long)x);
                if (k == 1) atomicAdd(A.stats + 6, (unsigned long long)x * c.n);
            }
        }
    }
    tc::cta_teardown(c);
}

// ---------------------------------------------------------------- parity pass through the SAME closest hit
#if !RT_EXACT
// rt_primary_hits / rt_closest_hits of the throughput build on a tensor-eligible scene: one ray per ray thread and pass,
// grid-stride over the rays; tc::closest_hit is the function k_render_tc calls.
__global__ void __launch_bounds__(RT_TC_THREADS, 1) k_hits_tc(const __grid_constant__ RtHitArgs A)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const tc::Cta c = tc::cta_setup(A.sc, smem_raw);
    const uint32_t warp = threadIdx.x >> 5;
    if (warp >= RT_TC_WGS * 4) {
#ifdef RT_TC_REG_RAY
        asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
#endif
        if (warp - RT_TC_WGS * 4 < RT_TC_WGS) tc::issuer_loop(c, warp - RT_TC_WGS * 4);
    } else {
#ifdef RT_TC_REG_RAY
        asm volatile("setmaxnreg.inc.sync.aligned.u32 " RT_TC_REG_RAY ";");
#endif
        tc::RayWarp rw = tc::ray_warp_setup(c);
        const unsigned long long total = A.rays ? A.nrays : (unsigned long long)A.w * A.h;
        const unsigned long long stride = (unsigned long long)gridDim.x * RT_TC_RAY_THREADS;
        for (unsigned long long i = (unsigned long long)blockIdx.x * RT_TC_RAY_THREADS + threadIdx.x;; i += stride) {
            const bool live = i < total;
            if (!tc::wg_any(1u + rw.wg, live)) {
                tc::ray_warp_finish(c, rw);
                break;
            }
            float ox = 0.f, oy = 0.f, oz = 0.f, dx = 0.f, dy = 0.f, dz = 0.f;
            if (live) hits_ray(A, i, ox, oy, oz, dx, dy, dz);
            float bt; int bi;
            tc::closest_hit(c, rw, A.sc, live, ox, oy, oz, dx, dy, dz, bt, bi, false, A.stats);
            if (live) hits_store(A, i, bi, bt, bi >= 0 ? c.s_sph[bi] : make_float4(0.f, 0.f, 0.f, 0.f), ox, oy, oz, dx, dy, dz);
        }
    }
    tc::cta_teardown(c);
}
#endif  // !RT_EXACT
