// rt_tensor.cuh -- brute-force closest hit with the candidate filter on the 5th-generation tensor cores.
// Included by rt_device.cuh in the throughput build only (namespace rtf).
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
// The product is only a CONSERVATIVE FILTER, exactly like the FMA-pipe filter it replaces: the sphere is
// inflated by eps (|c|^2 + r^2 + |o|^2), eps = 2^-16 (measured error of q: < 2^-20.4 of that scale,
// tools/proto_tc.cu), so no sphere the reference test accepts is dropped; the hit that is returned is decided
// by the reference test (sphere_candidate_fast) on the few candidates, in increasing sphere index.
//
// Per SM: ONE CTA = RT_TC_WGS ray warpgroups (128 rays each, one ray per thread = one TMEM lane) + one warpgroup of
// MMA-issuer warps (one per ray warpgroup; with 4 ray warpgroups they give their registers away: setmaxnreg).
// TMEM (512 columns): per warpgroup 32 columns of ray rows (A) + RT_TC_NBUF accumulator tiles (D) of RT_TC_CHUNK columns.
// Handshakes are mbarriers only: a_ready (ray rows stored) -> issuer; full[b] (tcgen05.commit) -> ray warps;
// empty[b] (accumulators copied to registers) -> issuer.  A tile is 64 spheres = 4 MMAs (M128 N64 K8).
// Measured shapes (huge scene, Gsamples/s): 3 warpgroups x 128 columns 1.33, 3 x (2 x 64) 1.35, 4 x 96 1.43, 4 x 64 1.46,
// 4 x (3 x 32) 1.18; candidate tests inlined per tile (RT_TC_INLINE_PASS2) 1.42; with the inversion sampler 4 x 64 1.64,
// 5 x 64 (shipped) 1.82, + I2IP.U2 masks 1.85-1.89.
// Epilogue: tcgen05.ld 2 x 32 columns -> 64 registers -> two sign-mask words (16 I2IP.U2 + 1 LOP3 each, RT_TC_MASKMODE)
// -> the non-empty mask words go to shared memory -> pass 2 as in the FMA-pipe build.
#pragma once

// RT_TC_WGS / RT_TC_CHUNK / RT_TC_NBUF (rt_types.h): ray warpgroups per CTA, spheres per accumulator tile, tiles per warpgroup
#define RT_TC_RAY_THREADS (RT_TC_WGS * 128)
#if RT_TC_SELFISSUE
#define RT_TC_ISSUER_THREADS 0          // the last ray warp to arrive issues the MMAs itself: no issuer warps
#else
#define RT_TC_ISSUER_THREADS (((RT_TC_WGS + 3) / 4) * 128)   // whole warpgroups (setmaxnreg is per warpgroup), one issuer warp per ray warpgroup
#endif
#define RT_TC_THREADS (RT_TC_RAY_THREADS + RT_TC_ISSUER_THREADS)
// register hand-over (the pool is the CTA's own allocation: what the issuers free must cover what the ray warps take)
//   4 ray warpgroups: 640 threads x 96 at launch; issuers 24, ray warps 112  (128 x 72 freed >= 512 x 16 taken)
//   5 ray warpgroups: 896 threads x 72 at launch; issuers 24, ray warps 88   (256 x 48 freed >= 640 x 16 taken)
#if RT_TC_SELFISSUE
// no hand-over: 640 threads get up to 96 registers each from the launch bound alone
#elif RT_TC_WGS == 4
#define RT_TC_REG_RAY "112"
#elif RT_TC_WGS == 5
#define RT_TC_REG_RAY "88"
#endif
// with 4 ray warpgroups: 640 threads x 96 registers at launch; the issuer warpgroup drops to 24, the ray warpgroups take 112
// (the pool is the CTA's own allocation: 128 x 72 freed >= 4 x 128 x 16 taken)
#define RT_TC_KQ 8                        // K = 32 tf32 slots = 8 chunks of 16 bytes
#define RT_TC_WGCOLS (32 + RT_TC_NBUF * RT_TC_CHUNK)   // TMEM columns per warpgroup
#define RT_TC_WGBARS (1 + 2 * RT_TC_NBUF)              // a_ready, full[NBUF], empty[NBUF]
static_assert(RT_TC_WGS * RT_TC_WGCOLS <= 512, "TMEM has 512 columns");
static_assert(RT_TC_CHUNK % 32 == 0 && RT_TC_CHUNK >= 32 && RT_TC_CHUNK <= 256, "tile width: whole mask words, one MMA N");
static_assert((1 + RT_TC_WGS * RT_TC_WGBARS) * 8 <= 448 && 452 + 12 * RT_TC_WGS <= 512, "barriers, TMEM base, done flags and arrival counters live in the first 512 bytes of shared memory");
#if RT_TC_SELFISSUE
static_assert(RT_TC_NBUF == 1 && RT_TC_MASK_POSITIVE && !RT_TC_INLINE_PASS2, "the self-issuing handshake: one accumulator tile per warpgroup, I2IP masks");
#endif

#ifdef RT_TC_PROF
#define TC_T(k) do { long long t_ = clock64(); prof[k] += t_ - tprev; tprev = t_; } while (0)
#else
#define TC_T(k) do { } while (0)
#endif

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
__device__ __forceinline__ void bar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n.reg .pred P1;\nTC_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra TC_DONE;\nbra TC_WAIT;\nTC_DONE:\n}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// issuer-side wait: may sleep long (suspend-time hint), so the idle issuer warps do not take issue slots
__device__ __forceinline__ void bar_wait_sleep(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n.reg .pred P1;\nTC_WAITS:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, 0x989680;\n"
        "@P1 bra TC_DONES;\nbra TC_WAITS;\nTC_DONES:\n}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// issuer-side wait for the ray rows (a long wait: the ray warps are in pass 2 / shading): back off with nanosleep
__device__ __forceinline__ void bar_wait_nanosleep(uint64_t *bar, uint32_t parity, uint32_t ns)
{
    for (;;) {
        uint32_t ok;
        asm volatile("{\n.reg .pred P1;\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\nselp.u32 %0, 1, 0, P1;\n}"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
        if (ok) break;
        __nanosleep(ns);
    }
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
// Arrival counter of the self-issuing handshake (RT_TC_SELFISSUE): returns the count before this arrival; the
// arrival whose old value is 3 mod 4 is the last of its warpgroup and issues the next tile's MMAs.  acq_rel at CTA
// scope orders the other warps' completed tcgen05.st / tcgen05.ld before the issue.
__device__ __forceinline__ uint32_t cnt_arrive(uint32_t *cnt)
{
    uint32_t old;
#if RT_TC_SELFISSUE == 2
    asm volatile("atom.relaxed.cta.shared::cta.add.u32 %0, [%1], 1;" : "=r"(old) : "r"(smem_u32(cnt)) : "memory");
#else
    asm volatile("atom.acq_rel.cta.shared::cta.add.u32 %0, [%1], 1;" : "=r"(old) : "r"(smem_u32(cnt)) : "memory");
#endif
    return old;
}
// The same on an mbarrier with 4 expected arrivals (RT_TC_SELFISSUE == 3): arrive (release) returns the state before the
// arrival; a pending count of 1 means this was the last one, and a test_wait (acquire) on that phase -- complete by
// construction -- orders the other warps' work before the issue.  Returns 3 for the last arriver, 0 otherwise.
__device__ __forceinline__ uint32_t bar_arrive_last(uint64_t *bar)
{
    uint32_t last;
    asm volatile("{\n.reg .b64 st;\n.reg .u32 pc;\n.reg .pred p, q;\n"
                 "mbarrier.arrive.shared::cta.b64 st, [%1];\n"
                 "mbarrier.pending_count.b64 pc, st;\n"
                 "setp.eq.u32 p, pc, 1;\n"
                 "@p mbarrier.test_wait.shared::cta.b64 q, [%1], st;\n"
                 "selp.u32 %0, 3, 0, p;\n}" : "=r"(last) : "r"(smem_u32(bar)) : "memory");
    return last;
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
// RT_TC_MASKMODE (rt_types.h) selects the reduction (default 4); the host lays the sphere matrix out accordingly:
//   0  I2IP: 16 two-value saturating s32->u8 packs + 8 LOP3.  bit (8p + j) <=> word 4j + 3 - p is a positive float;
//      the host places sphere 32 b + k in column 32 b + v with k = 7 + 8 (v & 3) - (v >> 2).  24 instructions per word.
//   1-3  the ray rows are stored NEGATED, a candidate is a NEGATIVE accumulator, word i -> bit 31 - i, identity layout:
//      1  FMA pipe: t = sat(-2^100 v) in {0,1}, FP32 Horner chains c = 2c + t (exact: chains stay below 2^16), the
//         two 16-bit halves read out of the mantissa with the 2^23 trick + one PRMT;
//      2  ALU pipe: funnel shifts, 4 chains of 8 + 3 combines;
//      3  FMA pipe on even words, ALU pipe on odd words (both pipes at once).
//      (An accumulator of magnitude below 2^-100 would make t fractional; the terms are O(1)..O(|c|^2) with 24-bit
//       significands, so a nonzero sum that small does not occur -- and a true candidate is >= eps * scale away from 0.)
__device__ __forceinline__ uint32_t sign_mask_i2ip(const uint32_t (&v)[32])
{
    uint32_t m = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        uint32_t t, w;
        asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(t) : "r"(v[4 * j + 0]), "r"(v[4 * j + 1]), "r"(0u));
        asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(w) : "r"(v[4 * j + 2]), "r"(v[4 * j + 3]), "r"(t));
        m |= w & (0x01010101u << j);
    }
    return m;
}
//   4  I2IP with 2-bit saturation: a positive s32 packs to 0b11, anything else to 0b00, and the instruction shifts its third
//      operand left by 4 -- a chain of 8 holds 16 values as bit pairs; two chains merge with ONE LOP3 (odd bits of the
//      first, even bits of the second): 17 instructions per word instead of 24.  bit 31 - 2k <=> word k, bit 30 - 2k <=>
//      word 16 + k; the host places sphere 32 b + s in column 32 b + (s even ? s/2 : 16 + (s-1)/2).
__device__ __forceinline__ uint32_t sign_mask_i2ip2(const uint32_t (&v)[32])
{
    uint32_t c0 = 0u, c1 = 0u;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        asm("cvt.pack.sat.u2.s32.b32 %0, %1, %2, %3;" : "=r"(c0) : "r"(v[2 * i]), "r"(v[2 * i + 1]), "r"(c0));
        asm("cvt.pack.sat.u2.s32.b32 %0, %1, %2, %3;" : "=r"(c1) : "r"(v[16 + 2 * i]), "r"(v[17 + 2 * i]), "r"(c1));
    }
    return (c0 & 0xaaaaaaaau) | (c1 & 0x55555555u);
}
__device__ __forceinline__ uint32_t sign_mask_alu(const uint32_t (&v)[32])
{
    uint32_t c[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        c[k] = 0u;
#pragma unroll
        for (int i = 0; i < 8; ++i) c[k] = __funnelshift_l(v[8 * k + i], c[k], 1);
    }
    return (c[0] << 24) | (c[1] << 16) | (c[2] << 8) | c[3];
}
__device__ __forceinline__ uint32_t sign_mask_fma(const uint32_t (&v)[32])
{
    float c[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        c[k] = __saturatef(__uint_as_float(v[8 * k]) * -1.2676506002282294e30f);
#pragma unroll
        for (int i = 1; i < 8; ++i) c[k] = fmaf(c[k], 2.f, __saturatef(__uint_as_float(v[8 * k + i]) * -1.2676506002282294e30f));
    }
    const float hi = fmaf(c[0], 256.f, c[1]) + 8388608.f, lo = fmaf(c[2], 256.f, c[3]) + 8388608.f;
    return __byte_perm(__float_as_uint(lo), __float_as_uint(hi), 0x5410);
}
template <int WORD>
__device__ __forceinline__ uint32_t sign_mask(const uint32_t (&v)[32])
{
#if RT_TC_MASKMODE == 0
    return sign_mask_i2ip(v);
#elif RT_TC_MASKMODE == 4
    return sign_mask_i2ip2(v);
#elif RT_TC_MASKMODE == 1
    return sign_mask_fma(v);
#elif RT_TC_MASKMODE == 2
    return sign_mask_alu(v);
#else
    return (WORD & 1) ? sign_mask_alu(v) : sign_mask_fma(v);
#endif
}

// Ray -> the 32 A-operand slots (slot order must match tc_sphere_row on the host, rt_api.cu)
__device__ __forceinline__ void ray_row(const RtDeviceScene &sc, bool live, float ox, float oy, float oz,
                                        float dx, float dy, float dz, float (&row)[32])
{
    ox -= sc.tc_gx; oy -= sc.tc_gy; oz -= sc.tc_gz;
    float len2 = dx * dx + dy * dy + dz * dz;
    float oo = ox * ox + oy * oy + oz * oz;
    bool ok = live && len2 > 0.f && len2 < 3.0e38f && oo < 1.0e30f;
    float inv = rsqrtf(len2);
    float ux = dx * inv, uy = dy * inv, uz = dz * inv;
    float uo = ux * ox + uy * oy + uz * oz;
    float qx = ox - uo * ux, qy = oy - uo * uy, qz = oz - uo * uz;
    float kk = fmaf(sc.tc_eps, oo, -(qx * qx + qy * qy + qz * qz));
    // RT_TC_NEG = -1: the row holds -q's coefficients, candidates come out as negative accumulators
    float co[9] = {ux * ux, uy * uy, uz * uz, 2.f * ux * uy, 2.f * ux * uz, 2.f * uy * uz, 2.f * qx, 2.f * qy, 2.f * qz};
#pragma unroll
    for (int t = 0; t < 9; ++t) co[t] *= RT_TC_NEG;
    kk *= RT_TC_NEG;
    if (!ok) {                              // no candidates: d = 0 / non-finite ray (the reference test rejects it), idle slot
#pragma unroll
        for (int t = 0; t < 9; ++t) co[t] = 0.f;
        kk = -3.0e38f * RT_TC_NEG;
    }
#pragma unroll
    for (int t = 0; t < 9; ++t) {
        float h = tf32_round(co[t]);
        row[3 * t + 0] = h; row[3 * t + 1] = h; row[3 * t + 2] = co[t] - h;
    }
    row[27] = RT_TC_NEG; row[28] = RT_TC_NEG; row[29] = RT_TC_NEG;
    float h = tf32_round(kk);
    row[30] = h; row[31] = kk - h;
}

}  // namespace tc

// shared memory: [0,512) barriers + TMEM base + done flags | B matrix | sphere records | candidate mask words
//   bars[0] staging; per warpgroup w, at bars + 1 + w * RT_TC_WGBARS: a_ready (4 arrivals), full[NBUF] (1), empty[NBUF] (4)
__global__ void __launch_bounds__(RT_TC_THREADS, 1) k_render_tc(const __grid_constant__ RtRenderArgs A)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem_raw);
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(smem_raw + 448);
    volatile uint32_t *s_done = reinterpret_cast<volatile uint32_t *>(smem_raw + 452);
#if RT_TC_SELFISSUE || RT_TC_NOVOTE
    uint32_t *s_cnt = reinterpret_cast<uint32_t *>(smem_raw + 452 + 4 * RT_TC_WGS);                  // arrivals (a_ready / tile consumed)
    uint32_t *s_fin = reinterpret_cast<uint32_t *>(smem_raw + 452 + 8 * RT_TC_WGS);                  // warps that ran out of work for good
#endif
    const uint32_t ns = A.sc.tc_ns, npad = A.sc.npad, n = A.sc.n;
    float4 *s_b = reinterpret_cast<float4 *>(smem_raw + 512);
    float4 *s_sph = s_b + (size_t)ns * RT_TC_KQ;
    uint32_t *s_mask = reinterpret_cast<uint32_t *>(s_sph + npad);
    // radii, material indices and materials next to the spheres when they fit (rt_render_tc_smem_bytes)
    const bool mats_in_smem = A.smem_mats != 0u;
    float4 *s_mat = reinterpret_cast<float4 *>(s_mask + (size_t)(ns / 32u) * RT_TC_RAY_THREADS);
    float *s_rad = reinterpret_cast<float *>(s_mat + A.sc.m);
    uint32_t *s_meta = reinterpret_cast<uint32_t *>(s_rad + npad);
    if (mats_in_smem) {
        for (uint32_t i = threadIdx.x; i < A.sc.m; i += RT_TC_THREADS) s_mat[i] = __ldg(A.sc.mat + i);
        for (uint32_t i = threadIdx.x; i < n; i += RT_TC_THREADS) { s_rad[i] = __ldg(A.sc.sph_rad + i); s_meta[i] = __ldg(A.sc.sph_meta + i); }
    }
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u;
    const uint32_t nch = ns / RT_TC_CHUNK;
#if RT_TC_STASH
    // path state that the tile loop does not touch is parked here while the 64 accumulators of a tile are in registers
    // (so that both halves of a tile can be loaded at once); [word][thread], at the end of the dynamic shared memory
    uint32_t *s_stash = reinterpret_cast<uint32_t *>(mats_in_smem ? reinterpret_cast<unsigned char *>(s_meta + npad)
                                                                   : reinterpret_cast<unsigned char *>(s_mat));
#endif

    if (tid == 0) {
        for (int w = 0; w < RT_TC_WGS; ++w) {
            uint64_t *bw = bars + 1 + w * RT_TC_WGBARS;
            tc::bar_init(bw, 4);
            for (int b = 0; b < RT_TC_NBUF; ++b) { tc::bar_init(bw + 1 + b, 1); tc::bar_init(bw + 1 + RT_TC_NBUF + b, 4); }
            s_done[w] = 0u;
#if RT_TC_SELFISSUE || RT_TC_NOVOTE
            s_cnt[w] = 0u; s_fin[w] = 0u;
#endif
        }
    }
    stage_spheres(s_sph, A.sc.sph, s_b, A.sc.tc_b, npad, ns * RT_TC_KQ, bars);      // inits bars[0], fences, TMA, wait
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(s_tmem)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc::fence_before();
    __syncthreads();
    tc::fence_after();
    const uint32_t tmem_base = *s_tmem;

#if !RT_TC_SELFISSUE
    if (warp >= RT_TC_WGS * 4) {
        // ============================================================ MMA issuer: one thread per warpgroup
        const uint32_t w = warp - RT_TC_WGS * 4;
#ifdef RT_TC_REG_RAY
        asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");                        // hand registers to the ray warpgroups
#endif
        if (w < RT_TC_WGS) {
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(RT_TC_CHUNK >> 3) << 17) | ((128u >> 4) << 24);
        const uint64_t bdesc0 = tc::smem_desc(smem_u32(s_b), ns * 16u, 128u);
        uint64_t *bar = bars + 1 + w * RT_TC_WGBARS;
        const uint32_t tmem_a = tmem_base + w * RT_TC_WGCOLS, tmem_d = tmem_a + 32;
        uint32_t uses = 0;                                                          // tiles issued so far
        for (uint32_t pass = 0;; ++pass) {
#if RT_TC_AREADY_SLEEP
            tc::bar_wait_nanosleep(bar + 0, pass & 1u, RT_TC_AREADY_SLEEP);
#else
            tc::bar_wait_sleep(bar + 0, pass & 1u);                                 // ray rows are in TMEM
#endif
#if RT_TC_NOVOTE
            if (*reinterpret_cast<volatile uint32_t *>(s_fin + w) >= 4u) {          // all four warps are out of work for good:
                if (lane == 0) { s_done[w] = 1u; tc::bar_arrive(bar + 1); }         // wake them (release) without an MMA
                break;
            }
#else
            if (s_done[w]) break;
#endif
            for (uint32_t c = 0; c < nch; ++c) {
                const uint32_t b = c % RT_TC_NBUF;                                  // nch is a multiple of NBUF: b = uses % NBUF too
                if (uses >= RT_TC_NBUF)                                             // previous tile in this buffer consumed
                    tc::bar_wait_sleep(bar + 1 + RT_TC_NBUF + b, (uses / RT_TC_NBUF - 1u) & 1u);
                tc::fence_after();
#pragma unroll
                for (uint32_t ks = 0; ks < RT_TC_KQ / 2; ++ks) {
                    // K step = 2 k-chunks (2 * ns * 16 bytes apart in s_b); tile c starts at row RT_TC_CHUNK * c
                    uint64_t bd = bdesc0 + (uint64_t)(((ks * 2u * ns + c * RT_TC_CHUNK) * 16u) >> 4);
                    tc::mma_tf32_ts(tmem_d + b * RT_TC_CHUNK, tmem_a + ks * 8u, bd, idesc, ks > 0u ? 1u : 0u);
                }
                tc::mma_commit(bar + 1 + b);
                ++uses;
            }
        }
        }
        __syncwarp();
    } else
#endif
    {
        // ============================================================ ray warps: one path per thread
#ifdef RT_TC_REG_RAY
        asm volatile("setmaxnreg.inc.sync.aligned.u32 " RT_TC_REG_RAY ";");
#endif
        const uint32_t wg = warp >> 2;
        const uint32_t lane_off = ((warp & 3u) * 32u) << 16;
        const uint32_t tmem_a = tmem_base + wg * RT_TC_WGCOLS + lane_off, tmem_d = tmem_a + 32;
        uint64_t *bar = bars + 1 + wg * RT_TC_WGBARS;
        uint32_t full_parity = 0;                   // bit b: parity of the next completion of full[b]
#if RT_TC_SELFISSUE || RT_TC_NOVOTE
        bool fin_counted = false;
#endif
#if RT_TC_SELFISSUE
        // the MMA operands of this warpgroup (used by whichever warp arrives last)
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(RT_TC_CHUNK >> 3) << 17) | ((128u >> 4) << 24);
        const uint64_t bdesc0 = tc::smem_desc(smem_u32(s_b), ns * 16u, 128u);
        const uint32_t tmem_a0 = tmem_base + wg * RT_TC_WGCOLS, tmem_d0 = tmem_a0 + 32;
        auto issue_tile = [&](uint32_t c) {          // converged warp; one elected lane issues
            tc::fence_after();
#pragma unroll
            for (uint32_t ks = 0; ks < RT_TC_KQ / 2; ++ks) {
                uint64_t bd = bdesc0 + (uint64_t)(((ks * 2u * ns + c * RT_TC_CHUNK) * 16u) >> 4);
                tc::mma_tf32_ts(tmem_d0, tmem_a0 + ks * 8u, bd, idesc, ks > 0u ? 1u : 0u);
            }
            tc::mma_commit(bar + 1);
        };
#endif

        PathSlot S;
        S.st = ST_EMPTY; S.pix = 0; S.smp = 0; S.bounce = 0;
        S.ox = S.oy = S.oz = 0.f; S.dx = S.dy = S.dz = 0.f;
        S.ar = S.ag = S.ab = 1.f;
        WarpPool pool; pool.next = 0; pool.end = 0; pool.exhausted = false;
        ThreadStats ts = {0, 0, 0, 0, 0, {0, 0}};
        float best_t = FLT_MAX; int best_i = -1;

#ifdef RT_TC_PROF
        long long prof[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tprev = clock64();
        uint32_t npass = 0;
        unsigned long long cand_sum = 0, cand_max = 0, live_sum = 0;
#endif
        for (;;) {
            shade_slot<false>(A, s_sph, S, best_t, best_i, pool, ts, false, mats_in_smem ? s_rad : nullptr,
                              mats_in_smem ? s_meta : nullptr, mats_in_smem ? s_mat : nullptr);
            TC_T(0);
            const bool live = (S.st == ST_ACTIVE);
#if RT_TC_SELFISSUE
            // no warpgroup vote: a warp whose 32 slots are empty has an exhausted pool (final), says so once, and keeps
            // taking part with idle rows until the last arriver of a pass finds all four warps finished
            if (!fin_counted && !__any_sync(RT_FULL, S.st != ST_EMPTY)) {
                fin_counted = true;
                if (lane == 0) atomicAdd(s_fin + wg, 1u);
            }
            TC_T(1);
            {
                float row[32];
                tc::ray_row(A.sc, live, S.ox, S.oy, S.oz, S.dx, S.dy, S.dz, row);
                RT_TC_ST32(tmem_a, row);
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                tc::fence_before();
                __syncwarp();
                uint32_t old = 0;
#if RT_TC_SELFISSUE == 3
                if (lane == 0) old = tc::bar_arrive_last(bar + 0);
#else
                if (lane == 0) old = tc::cnt_arrive(s_cnt + wg);
#endif
                old = __shfl_sync(RT_FULL, old, 0);
                if ((old & 3u) == 3u) {                                             // all four warps' rows are in TMEM
                    if (*reinterpret_cast<volatile uint32_t *>(s_fin + wg) >= 4u) { // ... and all four are out of work
                        if (lane == 0) { s_done[wg] = 1u; tc::bar_arrive(bar + 1); }   // wake the others (release) without an MMA
                    } else issue_tile(0u);
                }
            }
            TC_T(2);
#elif RT_TC_NOVOTE
            // no warpgroup vote: a warp whose 32 slots are empty has an exhausted pool (final) and says so once; the issuer
            // reads the count when all four have arrived and ends the warpgroup through full[0]
            if (!fin_counted && !__any_sync(RT_FULL, S.st != ST_EMPTY)) {
                fin_counted = true;
                if (lane == 0) atomicAdd(s_fin + wg, 1u);
            }
            TC_T(1);
            {
                float row[32];
                tc::ray_row(A.sc, live, S.ox, S.oy, S.oz, S.dx, S.dy, S.dz, row);
                RT_TC_ST32(tmem_a, row);
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                tc::fence_before();
                __syncwarp();
                if (lane == 0) tc::bar_arrive(bar + 0);
            }
            TC_T(2);
#else
            if (!tc::wg_any(1u + wg, S.st != ST_EMPTY)) {                           // whole warpgroup out of work
                if ((tid & 127u) == 0u) s_done[wg] = 1u;
                __syncwarp();
                if (lane == 0) tc::bar_arrive(bar + 0);
                break;
            }
            TC_T(1);
            {
                float row[32];
                tc::ray_row(A.sc, live, S.ox, S.oy, S.oz, S.dx, S.dy, S.dz, row);
                RT_TC_ST32(tmem_a, row);
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                tc::fence_before();
                __syncwarp();
                if (lane == 0) tc::bar_arrive(bar + 0);
            }
            TC_T(2);
#endif
            uint32_t nz = 0u;                       // bit (31 - w): mask word w is non-empty
            best_t = FLT_MAX; best_i = -1;
#if RT_TC_STASH
            {
                uint32_t *q = s_stash + tid;
                q[0 * RT_TC_RAY_THREADS] = __float_as_uint(S.ar); q[1 * RT_TC_RAY_THREADS] = __float_as_uint(S.ag);
                q[2 * RT_TC_RAY_THREADS] = __float_as_uint(S.ab); q[3 * RT_TC_RAY_THREADS] = S.pix;
                q[4 * RT_TC_RAY_THREADS] = S.smp; q[5 * RT_TC_RAY_THREADS] = S.bounce;
                q[6 * RT_TC_RAY_THREADS] = ts.samples; q[7 * RT_TC_RAY_THREADS] = ts.segments;
                q[8 * RT_TC_RAY_THREADS] = ts.sky; q[9 * RT_TC_RAY_THREADS] = ts.absorbed;
                q[10 * RT_TC_RAY_THREADS] = ts.capped; q[11 * RT_TC_RAY_THREADS] = pool.next;
                q[12 * RT_TC_RAY_THREADS] = pool.end;
            }
#else
            const float ray_a = S.dx * S.dx + S.dy * S.dy + S.dz * S.dz;
            const float ray_inv_a = __frcp_rn(ray_a);
#endif
#if RT_TC_STASH
#pragma unroll 1                                    // one loop body: both halves of a tile in disjoint registers, loaded back to back
#endif
            for (uint32_t c = 0; c < nch; ++c) {
                const uint32_t b = c % RT_TC_NBUF;
                tc::bar_wait(bar + 1 + b, (full_parity >> b) & 1u);
                full_parity ^= 1u << b;
#if RT_TC_SELFISSUE || RT_TC_NOVOTE
                if (c == 0u && s_done[wg]) goto tc_finished;                        // woken by the exit arrival, not by an MMA
#endif
                tc::fence_after();
                TC_T(3);
#if RT_TC_MASK_POSITIVE
#pragma unroll
                for (uint32_t i = 0; i < RT_TC_CHUNK / 32; i += 2) {           // 64 columns (one odd 32 at the end) at a time
                    const bool two = i + 1 < RT_TC_CHUNK / 32, last = i + 2 >= RT_TC_CHUNK / 32;
                    uint32_t v0[32], v1[32];
                    RT_TC_LD32(v0, tmem_d + b * RT_TC_CHUNK + i * 32u);
                    if (two) RT_TC_LD32(v1, tmem_d + b * RT_TC_CHUNK + i * 32u + 32u);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    TC_T(4);
#if RT_TC_SELFISSUE
                    uint32_t old = 0;
                    if (last && c + 1u < nch) {      // the whole tile is in registers: count this warp in ...
                        tc::fence_before();
                        __syncwarp();
#if RT_TC_SELFISSUE == 3
                        if (lane == 0) old = tc::bar_arrive_last(bar + 0);
#else
                        if (lane == 0) old = tc::cnt_arrive(s_cnt + wg);
#endif
                        old = __shfl_sync(RT_FULL, old, 0);
                        if ((old & 3u) == 3u) issue_tile(c + 1u);    // ... and the last of the four issues the next tile
                    }
                    const uint32_t w0 = c * (RT_TC_CHUNK / 32) + i;
                    const uint32_t m0 = tc::sign_mask<0>(v0);
#else
                    if (last) {                      // the whole tile is in registers: hand the buffer back
                        tc::fence_before();
                        __syncwarp();
                        if (lane == 0) tc::bar_arrive(bar + 1 + RT_TC_NBUF + b);
                    }
                    const uint32_t w0 = c * (RT_TC_CHUNK / 32) + i;
                    const uint32_t m0 = tc::sign_mask<0>(v0);
#endif
#if RT_TC_INLINE_PASS2
                    // pass 2 right here, while the MMAs of the next tile run: this lane's candidates of these 64 columns
                    unsigned long long mm = (unsigned long long)m0 << 32;
                    if (two) mm |= tc::sign_mask<1>(v1);
                    TC_T(5);
                    while (mm) {
                        const int lz = __clzll((long long)mm);
                        mm &= ~(0x8000000000000000ull >> lz);
                        const uint32_t j = (w0 << 5) + (uint32_t)lz;
                        if (j < n) sphere_candidate_fast(s_sph[j], (int)j, S.ox, S.oy, S.oz, S.dx, S.dy, S.dz, ray_a, ray_inv_a, best_t, best_i);
                    }
                    TC_T(6);
#else
                    if (m0) { s_mask[w0 * RT_TC_RAY_THREADS + tid] = m0; nz |= 0x80000000u >> w0; }
                    if (two) {
                        const uint32_t m1 = tc::sign_mask<1>(v1);
                        if (m1) { s_mask[(w0 + 1u) * RT_TC_RAY_THREADS + tid] = m1; nz |= 0x80000000u >> (w0 + 1u); }
                    }
                    TC_T(5);
#endif
                }
#else
#pragma unroll
                for (uint32_t i = 0; i < RT_TC_CHUNK / 32; ++i) {              // 32 columns at a time (register budget)
                    uint32_t v0[32];
                    RT_TC_LD32(v0, tmem_d + b * RT_TC_CHUNK + i * 32u);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    TC_T(4);
                    if (i + 1 == RT_TC_CHUNK / 32) {  // the whole tile is in registers: hand the buffer back
                        tc::fence_before();
                        __syncwarp();
                        if (lane == 0) tc::bar_arrive(bar + 1 + RT_TC_NBUF + b);
                    }
                    const uint32_t w0 = c * (RT_TC_CHUNK / 32) + i;
                    const uint32_t m0 = (i & 1u) ? tc::sign_mask<1>(v0) : tc::sign_mask<0>(v0);
                    if (m0) { s_mask[w0 * RT_TC_RAY_THREADS + tid] = m0; nz |= 0x80000000u >> w0; }
                    TC_T(5);
                }
#endif
            }
            // pass 2: the reference test on this lane's candidates, increasing sphere index, strict <
#if RT_TC_STASH
            {
                const uint32_t *q = s_stash + tid;
                S.ar = __uint_as_float(q[0 * RT_TC_RAY_THREADS]); S.ag = __uint_as_float(q[1 * RT_TC_RAY_THREADS]);
                S.ab = __uint_as_float(q[2 * RT_TC_RAY_THREADS]); S.pix = q[3 * RT_TC_RAY_THREADS];
                S.smp = q[4 * RT_TC_RAY_THREADS]; S.bounce = q[5 * RT_TC_RAY_THREADS];
                ts.samples = q[6 * RT_TC_RAY_THREADS]; ts.segments = q[7 * RT_TC_RAY_THREADS];
                ts.sky = q[8 * RT_TC_RAY_THREADS]; ts.absorbed = q[9 * RT_TC_RAY_THREADS];
                ts.capped = q[10 * RT_TC_RAY_THREADS]; pool.next = q[11 * RT_TC_RAY_THREADS];
                pool.end = q[12 * RT_TC_RAY_THREADS];
            }
            const float ray_a = S.dx * S.dx + S.dy * S.dy + S.dz * S.dz;
            const float ray_inv_a = __frcp_rn(ray_a);
#endif
            {
                const float a = ray_a, inv_a = ray_inv_a;
#if !(RT_TC_INLINE_PASS2 && RT_TC_MASK_POSITIVE)
                uint32_t words = live ? nz : 0u, cand = 0u, w = 0u;
#ifdef RT_TC_PROF
                uint32_t my_cands = 0;
#endif
#if RT_TC_PASS2_PAIR
                // two candidates per iteration: the two tests are independent until the merge, so their shared-memory,
                // FMA and MUFU latencies overlap (2.8 candidates per ray on average, 5.4 for the slowest lane of a warp)
                auto next = [&]() -> int {
                    if (cand == 0u) {
                        if (words == 0u) return -1;
                        w = (uint32_t)__clz(words);
                        words &= ~(0x80000000u >> w);
                        cand = s_mask[w * RT_TC_RAY_THREADS + tid];
                    }
                    const int lz = __clz(cand);
                    cand &= ~(0x80000000u >> lz);
                    const uint32_t j = (w << 5) + (uint32_t)lz;
                    return j < n ? (int)j : -1;          // padding columns never test positive; guard anyway
                };
                for (;;) {
                    const int j0 = next();
                    if (j0 < 0) break;
                    const int j1 = next();
                    const float4 s0 = s_sph[j0], s1 = s_sph[j1 < 0 ? j0 : j1];
                    const float t0 = sphere_root_fast(s0, S.ox, S.oy, S.oz, S.dx, S.dy, S.dz, a, inv_a);
                    const float t1 = sphere_root_fast(s1, S.ox, S.oy, S.oz, S.dx, S.dy, S.dz, a, inv_a);
                    if (t0 < best_t) { best_t = t0; best_i = j0; }
                    if (j1 >= 0 && t1 < best_t) { best_t = t1; best_i = j1; }
#ifdef RT_TC_PROF
                    my_cands += (j1 >= 0) ? 2u : 1u;
#endif
                }
#else
                for (;;) {
                    if (cand == 0u) {
                        if (words == 0u) break;
                        w = (uint32_t)__clz(words);
                        words &= ~(0x80000000u >> w);
                        cand = s_mask[w * RT_TC_RAY_THREADS + tid];
                    }
                    const int lz = __clz(cand);
                    cand &= ~(0x80000000u >> lz);
                    const uint32_t j = (w << 5) + (uint32_t)lz;
                    if (j < n) sphere_candidate_fast(s_sph[j], (int)j, S.ox, S.oy, S.oz, S.dx, S.dy, S.dz, a, inv_a, best_t, best_i);
#ifdef RT_TC_PROF
                    ++my_cands;
#endif
                }
#endif
#ifdef RT_TC_PROF
                cand_sum += __reduce_add_sync(RT_FULL, my_cands);
                cand_max += __reduce_max_sync(RT_FULL, my_cands);
                live_sum += __popc(__ballot_sync(RT_FULL, live));
#endif
#endif
                if (A.check && live) {              // self-check: the same test over all spheres must agree
                    float ct = FLT_MAX; int ci = -1;
                    for (uint32_t j = 0; j < n; ++j)
                        sphere_candidate_fast(s_sph[j], (int)j, S.ox, S.oy, S.oz, S.dx, S.dy, S.dz, a, inv_a, ct, ci);
                    if (ci != best_i || ct != best_t) atomicAdd(A.stats + 7, 1ull);
                }
            }
            TC_T(6);
#ifdef RT_TC_PROF
            ++npass;
#endif
        }

#if RT_TC_SELFISSUE || RT_TC_NOVOTE
    tc_finished:;
#endif
#ifdef RT_TC_PROF
        if (blockIdx.x == 3 && lane == 0)
            printf("warp %2u passes %u | per pass: shade %6.0f  wg-barrier %5.0f  row+st %5.0f  wait-mma %6.0f  ldtm %5.0f  masks %6.0f  pass2 %6.0f\n", warp, npass,
                   (double)prof[0] / npass, (double)prof[1] / npass, (double)prof[2] / npass, (double)prof[3] / npass, (double)prof[4] / npass,
                   (double)prof[5] / npass, (double)prof[6] / npass);
        if (blockIdx.x == 3 && lane == 0 && (warp & 3u) == 0u)
            printf("warp %2u candidates per live ray %.2f, warp max per pass %.2f, live lanes %.1f\n", warp,
                   (double)cand_sum / (double)live_sum, (double)cand_max / npass, (double)live_sum / npass);
#endif
        uint32_t v[5] = {ts.samples, ts.segments, ts.sky, ts.absorbed, ts.capped};
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            uint32_t x = v[k];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(RT_FULL, x, o);
            if (lane == 0 && x) {
                atomicAdd(A.stats + k, (unsigned long long)x);
                if (k == 1) atomicAdd(A.stats + 6, (unsigned long long)x * n);
            }
        }
    }
    tc::fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base));
}
