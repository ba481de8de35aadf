// rt_device.cuh -- device code of the path-tracing hot path, compiled TWICE:
//   rt_kernels_exact.cu : -DRT_EXACT=1, nvcc --fmad=false -> namespace rtx.  The parity kernels: every float operation is an
//                         IEEE mul/add/div/sqrt in the reference's operation order (rejection sampler, IEEE normalise /
//                         divide, Schlick in double); renders are bit-identical to the CPU oracle (oracle/rt_oracle.c) path
//                         by path.  Includes the tensor-core kernel (rt_tensor.cuh) with this shading: RT_FLAG_EXACT on a
//                         48..1024-sphere scene runs the timed kernel's structure and closest hit.
//   rt_kernels_fast.cu  : -DRT_EXACT=0, FMA contraction on -> namespace rtf.  The throughput kernels.  The CLOSEST HIT is
//                         written with explicit round-to-nearest intrinsics (__fmul_rn, __fadd_rn, __fsqrt_rn, __fdiv_rn:
//                         never contracted), so it is the same IEEE arithmetic in both units: a conservative filter --
//                         tensor cores or explicit FMA chains -- only selects the candidates, the reference test decides,
//                         and hit index, t and normal are bit-identical to the oracle for the same ray.  The shading of this
//                         unit samples the unit ball by inversion, uses rsqrt / reciprocal in the normalisations and lets
//                         the compiler contract (the timed default; same law, checked statistically).
//
// Reference semantics restated here (paths relative to the reference repo):
//   intersect            src/raytracer.hxx:52-92      hit_world (first minimum)   src/raytracer.hxx:94-118
//   apply_material       src/raytracer.hxx:120-199    schlick (double)            src/raytracer.hxx:45-50
//   random_in_unit_sphere src/raytracer.hxx:32-43     camera::ray                 src/camera.hxx:46-57
//   app::color           src/main.cxx:52-75           background                  src/main.cxx:47-50,71
//   reflect/refract/normalize/mix                     src/math.hxx:219-227,294-335
#pragma once
#include <cfloat>
#if defined(RT_TC_PROF) || RT_BVH_PATHSTAT
#include <cstdio>
#endif
#include "rt_types.h"

#ifndef RT_EXACT
#error "define RT_EXACT to 0 or 1"
#endif

#if defined(RT_NS_OVERRIDE)
#define RT_NS RT_NS_OVERRIDE       // rt_kernels_bvh.cu: namespace rtb
#elif RT_EXACT
#define RT_NS rtx
#else
#define RT_NS rtf
#endif

namespace RT_NS {

#define RT_FULL 0xffffffffu

// -DRT_DEBUG: device-side asserts on the TMEM column ranges, barrier bookkeeping, work-item / tile / feed indices and hit
// indices (compute-sanitizer cannot attach on the GPU pool; the GPU test-suite is run once against this build instead:
// tools/build_variant.sh debug -DRT_DEBUG=1; RTIOW_LIB=build/variants/debug/librtiow_b200.so pytest -m gpu)
#ifdef RT_DEBUG
#include <cassert>
#define RT_DASSERT(x) assert(x)
#else
#define RT_DASSERT(x) do { } while (0)
#endif

struct f3 { float x, y, z; };
__device__ __forceinline__ f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }

// ------------------------------------------------------------------ PTX helpers (TMA bulk copy)
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t phase)
{
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "RT_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra RT_DONE;\n"
        "bra RT_WAIT;\n"
        "RT_DONE:\n"
        "}" ::"r"(smem_u32(bar)), "r"(phase) : "memory");
}

// Stage the SoA sphere records {cx,cy,cz,r*r} (and, in the throughput build, the filter records) into
// shared memory with TMA bulk copies (<= 64 KB pieces) completing on one mbarrier; every thread waits on it.
// Called once per CTA (the render kernel is persistent).
__device__ __forceinline__ void stage_spheres(float4 *s_a, const float4 *g_a, float4 *s_b, const float4 *g_b,
                                              uint32_t na, uint32_t nb, uint64_t *bar)
{
    if (threadIdx.x == 0) mbar_init(bar, 1);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t bytes_a = na * 16u, bytes_b = nb * 16u;
        mbar_expect_tx(bar, bytes_a + bytes_b);
        for (uint32_t off = 0; off < bytes_a; off += 65536u)
            tma_bulk_g2s((char *)s_a + off, (const char *)g_a + off, min(65536u, bytes_a - off), bar);
        for (uint32_t off = 0; off < bytes_b; off += 65536u)
            tma_bulk_g2s((char *)s_b + off, (const char *)g_b + off, min(65536u, bytes_b - off), bar);
    }
    mbar_wait(bar, 0);
}

// ------------------------------------------------------------------ RNG: Philox4x32-10, counter based
// counter = (pixel, sample, event, block), key = seed.  event 0 = camera ray (block 0 words 0,1 = pixel
// jitter; lens-ball attempts start at block 1), event b+1 = the scatter after segment b (ball attempts
// start at block 0, attempt k uses words 0..2 of block k; the dielectric uses word 0 of block 0).
__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1)
{
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}
__device__ __forceinline__ float u01(uint32_t w) { return (float)(w >> 8) * 5.9604644775390625e-8f; }
__device__ __forceinline__ float s11(uint32_t w) { return (float)(w >> 8) * 5.9604644775390625e-8f * 2.f + -1.f; }

#if RT_EXACT
#define RT_PAR_DEFAULT true
#else
#define RT_PAR_DEFAULT false
#endif
// Throughput shading: random_in_unit_sphere (src/raytracer.hxx:32-43) by inversion instead of rejection -- the same law
// (uniform in the unit ball) from three words: z/r uniform in [-1,1), azimuth uniform, radius U^(1/3).
__device__ __forceinline__ void ball_by_inversion(uint32_t wz, uint32_t wphi, uint32_t wr, float &x, float &y, float &z)
{
    float sn, cs;
    const float zz = s11(wz);
    __sincosf(u01(wphi) * 6.28318548f, &sn, &cs);
    const float rr = exp2f(__log2f(u01(wr)) * 0.333333343f);              // U^(1/3)
    const float sxy = rr * sqrtf(fmaxf(1.f - zz * zz, 0.f));
    x = sxy * cs; y = sxy * sn; z = rr * zz;
}
// x and y of such a point (all the lens offset uses, src/camera.hxx:51-53): rho = sqrt(1 - U^(2/3)), azimuth uniform.
__device__ __forceinline__ void ball_xy_by_inversion(uint32_t wrho, uint32_t wphi, float &x, float &y)
{
    float sn, cs;
    __sincosf(u01(wphi) * 6.28318548f, &sn, &cs);
    const float q = exp2f(__log2f(u01(wrho)) * 0.666666687f);             // U^(2/3)
    const float rho = sqrtf(fmaxf(1.f - q, 0.f));
    x = rho * cs; y = rho * sn;
}
// RT_FLAG_BOOK_OPTICS lens: uniform in the unit disk.
__device__ __forceinline__ void disk_by_inversion(uint32_t wrho, uint32_t wphi, float &x, float &y)
{
    float sn, cs;
    __sincosf(u01(wphi) * 6.28318548f, &sn, &cs);
    const float rho = sqrtf(u01(wrho));
    x = rho * cs; y = rho * sn;
}

// ------------------------------------------------------------------ vector math in the reference's order
__device__ __forceinline__ float dot3(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// src/math.hxx:219-227
template <bool PAR = RT_PAR_DEFAULT>
__device__ __forceinline__ f3 normalize3(f3 a)
{
    if constexpr (PAR) {
        float len = sqrtf(a.x * a.x + a.y * a.y + a.z * a.z);
        if (fabsf(len) > FLT_MIN) { a.x = a.x / len; a.y = a.y / len; a.z = a.z / len; }
    } else {
        float len2 = fmaf(a.z, a.z, fmaf(a.y, a.y, a.x * a.x));
        if (len2 > 1.4e-38f) {                             // length > FLT_MIN: any normal len2
            float inv = rsqrtf(len2);
            a.x *= inv; a.y *= inv; a.z *= inv;
        }
    }
    return a;
}
// src/math.hxx:294-298
__device__ __forceinline__ f3 reflect3(f3 I, f3 N)
{
    float dn = dot3(N, I);
    return mk3(I.x - (N.x * dn) * 2.f, I.y - (N.y * dn) * 2.f, I.z - (N.z * dn) * 2.f);
}
// src/math.hxx:300-309 with its quirk: dot*eta is added to every component; k < 0 gives NaN.
__device__ __forceinline__ f3 refract3(f3 I, f3 N, float eta)
{
    float dv = dot3(N, I);
    float k = 1.f - eta * eta * (1.f - dv * dv);
    float sk = sqrtf(k);
    float s = dv * eta;
    float z = (k >= 0.f) ? 1.f : 0.f;
    return mk3((I.x * eta - (N.x * sk + s)) * z, (I.y * eta - (N.y * sk + s)) * z, (I.z * eta - (N.z * sk + s)) * z);
}
// RT_FLAG_BOOK_OPTICS: Snell refraction (GLSL refract); k < 0 (total internal reflection) gives the zero vector, which
// the caller (src/raytracer.hxx:175-191) turns into a certain reflection.
__device__ __forceinline__ f3 refract3_book(f3 I, f3 N, float eta)
{
    float dv = dot3(N, I);
    float k = 1.f - eta * eta * (1.f - dv * dv);
    if (!(k >= 0.f)) return mk3(0.f, 0.f, 0.f);
    float s = eta * dv + sqrtf(k);
    return mk3(I.x * eta - N.x * s, I.y * eta - N.y * s, I.z * eta - N.z * s);
}
// src/raytracer.hxx:45-50
template <bool PAR = RT_PAR_DEFAULT>
__device__ __forceinline__ float schlick(float ri, float cosine)
{
    if constexpr (PAR) {
        double q = (double)((1.f - ri) / (1.f + ri));
        double r0 = q * q;
        double x = (double)(1.f - cosine);
        double x2 = x * x;
        double x5 = x2 * x2 * x;
        return (float)(r0 + (1. - r0) * x5);
    } else {
        float q = (1.f - ri) / (1.f + ri);
        float r0 = q * q;
        float x = 1.f - cosine;
        float x2 = x * x;
        return r0 + (1.f - r0) * (x2 * x2 * x);
    }
}

// ------------------------------------------------------------------ closest hit, brute force
// The reference test, src/raytracer.hxx:55-89, in explicit round-to-nearest intrinsics: every operation is a separate IEEE
// multiply / add / sqrt / divide in the reference's order, whatever the translation unit's contraction setting.
// a = dot(d, d), src/raytracer.hxx:56
__device__ __forceinline__ float ray_a(float dx, float dy, float dz)
{
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}
// oc, b, c and the discriminant b*b - a*c (src/raytracer.hxx:55-61); s.w = radius * radius
__device__ __forceinline__ float sphere_disc(float4 s, float ox, float oy, float oz, float dx, float dy, float dz,
                                             float a, float &b, float &c)
{
    const float ocx = __fsub_rn(ox, s.x), ocy = __fsub_rn(oy, s.y), ocz = __fsub_rn(oz, s.z);
    b = __fadd_rn(__fadd_rn(__fmul_rn(ocx, dx), __fmul_rn(ocy, dy)), __fmul_rn(ocz, dz));
    c = __fsub_rn(__fadd_rn(__fadd_rn(__fmul_rn(ocx, ocx), __fmul_rn(ocy, ocy)), __fmul_rn(ocz, ocz)), s.w);
    return __fsub_rn(__fmul_rn(b, b), __fmul_rn(a, c));
}
__device__ __forceinline__ float sphere_disc(float4 s, float ox, float oy, float oz, float dx, float dy, float dz,
                                             float a, float &b)
{
    float c;
    return sphere_disc(s, ox, oy, oz, dx, dy, dz, a, b, c);
}

// Early-outs of the throughput kernels' candidate test.  Each one only SKIPS work whose outcome is already decided: when
// it fires, the plain reference test below would leave (best_t, best_i) untouched, so the result stays the oracle's bit for
// bit (RT_FLAG_FILTER_CHECK compares against the plain test over all spheres for every segment).
//   bit 0  b > 0 and c > 0: the origin is outside the sphere and the sphere lies behind the ray.  a c >= 0 makes
//          disc = RN(RN(b b) - RN(a c)) <= RN(b b), so sq = RN(sqrt(disc)) <= RN(sqrt(RN(b b))) = b and both roots
//          (-b -+ sq) / a are <= 0 < kMIN in the reference's own float arithmetic.  (The tensor / FMA filters test the
//          LINE, so spheres behind a bounce ray's origin are candidates: this saves their sqrt and two divisions.)
//   bit 1  approximate roots with an error bound.  t1', t2' from sqrt.approx / rcp.approx differ from the IEEE roots by
//          less than m = 2^-19 (|b| + sq) / a (derivation: DESIGN.md section 4; the individual errors add up to < 2^-20).
//          Skip when the smallest root that can still be valid is beyond the running best (t' - m > best_t: a loser) or
//          when even t2' + m < kMIN (both roots invalid: the sphere the path just left).  NaN / inf in the approximation
//          make every comparison false, i.e. fall through to the IEEE evaluation.
#ifndef RT_CAND_EARLYOUT
#define RT_CAND_EARLYOUT 1
#endif
__device__ __forceinline__ bool candidate_decided(float disc, float b, float c, float a, float best_t)
{
#if !RT_EXACT
    if ((RT_CAND_EARLYOUT & 1) && b > 0.f && c > 0.f) return true;
    if (RT_CAND_EARLYOUT & 2) {
        float sq, ra;
        asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(disc));
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(ra) : "f"(a));
        const float m = (fabsf(b) + sq) * ra * 1.9073486328125e-6f;                 // 2^-19
        const float t1 = (-b - sq) * ra, t2 = (sq - b) * ra;
        if (t2 + m < RT_KMIN) return true;                                          // both roots below kMIN
        const float tl = (t1 + m < RT_KMIN) ? t2 : t1;                             // smallest root that can still be valid
        if (tl - m > best_t) return true;                                           // cannot beat the running closest hit
    }
#endif
    return false;
}

// n / a, correctly rounded, from the correctly rounded reciprocal ra = RN(1 / a) of the ray's a, computed ONCE per ray
// (ray_ra): two residual corrections with exact FMA residuals (Markstein: with ra = RN(1/a) and q1 faithful,
// RN(q1 + (n - a q1) ra) = RN(n / a)).  5 dependent FMA-pipe instructions instead of the ~10 + MUFU + range check of
// div.rn.  Verified against IEEE division on 4 x 10^8 random pairs with stressed significands (tools/check_mdiv.c: no
// mismatch where the quotient is a normal number; a subnormal or overflowing quotient fails t > kMIN / t < best_t either
// way).  ray_ra returns 0 outside 2^-60 < a < 2^60 -- never reached by a ray direction -- and div_by_a then divides.
#ifndef RT_CAND_MDIV
#define RT_CAND_MDIV 1
#endif
__device__ __forceinline__ float ray_ra(float a)
{
#if RT_CAND_MDIV && !RT_EXACT
    return (a > 8.7e-19f && a < 1.15e18f) ? __frcp_rn(a) : 0.f;
#else
    return 0.f;
#endif
}
__device__ __forceinline__ float div_by_a(float n, float a, float ra)
{
#if RT_CAND_MDIV && !RT_EXACT
    if (ra != 0.f) {
        const float q0 = __fmul_rn(n, ra);
        const float q1 = __fmaf_rn(__fmaf_rn(-a, q0, n), ra, q0);
        return __fmaf_rn(__fmaf_rn(-a, q1, n), ra, q1);
    }
#endif
    return __fdiv_rn(n, a);
}

// Per-candidate reference test (src/raytracer.hxx:63-89) against the running closest hit.  Candidates are
// visited in increasing sphere index and the comparison is strict, so ties keep the LOWEST index --
// hit_world's stable_partition + first min_element (src/raytracer.hxx:107-117).
// sphere_candidate_ref is the plain test; sphere_candidate adds the early-outs above in the throughput unit.
__device__ __forceinline__ void sphere_candidate_ref(float4 s, int j, float ox, float oy, float oz,
                                                     float dx, float dy, float dz, float a, float &best_t, int &best_i)
{
    float b;
    float disc = sphere_disc(s, ox, oy, oz, dx, dy, dz, a, b);
    if (disc > 0.f) {
        float sq = __fsqrt_rn(disc);
        float t = __fdiv_rn(__fsub_rn(-b, sq), a);
        if (t < best_t && t > RT_KMIN) { best_t = t; best_i = j; }
        else {
            t = __fdiv_rn(__fadd_rn(-b, sq), a);
            if (t < best_t && t > RT_KMIN) { best_t = t; best_i = j; }
        }
    }
}
__device__ __forceinline__ void sphere_candidate(float4 s, int j, float ox, float oy, float oz, float dx, float dy, float dz,
                                                 float a, float ra, float &best_t, int &best_i)
{
#ifndef RT_CAND_SINGLE_EXIT
#define RT_CAND_SINGLE_EXIT 1
#endif
    float b, c;
    const float disc = sphere_disc(s, ox, oy, oz, dx, dy, dz, a, b, c);
#if RT_CAND_SINGLE_EXIT
    // one exit, one compare-and-select at the end: t = the root the reference would look at last, FLT_MAX if none
    float t = FLT_MAX;
    if (disc > 0.f && !candidate_decided(disc, b, c, a, best_t)) {
        const float sq = __fsqrt_rn(disc);
        t = div_by_a(__fsub_rn(-b, sq), a, ra);
        if (!(t < best_t && t > RT_KMIN)) t = div_by_a(__fadd_rn(-b, sq), a, ra);
    }
    if (t < best_t && t > RT_KMIN) { best_t = t; best_i = j; }
#else
    if (disc > 0.f && !candidate_decided(disc, b, c, a, best_t)) {
        float sq = __fsqrt_rn(disc);
        float t = div_by_a(__fsub_rn(-b, sq), a, ra);
        if (t < best_t && t > RT_KMIN) { best_t = t; best_i = j; }
        else {
            t = div_by_a(__fadd_rn(-b, sq), a, ra);
            if (t < best_t && t > RT_KMIN) { best_t = t; best_i = j; }
        }
    }
#endif
}

// Brute-force closest hit for R rays per thread over the shared-memory sphere buffer.
// Pass 1 (hot): per sphere 1 LDS.128 (warp-broadcast) + per ray the discriminant and one funnel shift
//   that pushes its sign bit into a 32-sphere mask word -- no branch, no sqrt, no divide.
// Pass 2: each lane walks its own candidate bits (typically 3-10 of 488 spheres) and runs the exact
//   reference test.  Lanes skip empty mask words independently, so the warp iterates max-over-lanes
//   of the candidate COUNT, not the number of mask words.
// s_mask layout: [word][ray][thread] (conflict-free).
template <int R>
__device__ __forceinline__ void closest_hit_brute(const float4 *__restrict__ s_sph, uint32_t n, uint32_t npad,
                                                  uint32_t *__restrict__ s_mask, const bool (&live)[R],
                                                  const float (&ox)[R], const float (&oy)[R], const float (&oz)[R],
                                                  const float (&dx)[R], const float (&dy)[R], const float (&dz)[R],
                                                  float (&best_t)[R], int (&best_i)[R])
{
    float a[R];
    uint32_t nz[R];                                  // bit (31 - w): mask word w of the chunk is non-empty
#pragma unroll
    for (int r = 0; r < R; ++r) {
        a[r] = ray_a(dx[r], dy[r], dz[r]);
        best_t[r] = FLT_MAX;
        best_i[r] = -1;
        nz[r] = 0u;
    }
    const uint32_t nthreads = blockDim.x;
    for (uint32_t chunk = 0; chunk < npad; chunk += 32u * RT_CHUNK_WORDS) {
        const uint32_t nwords = min((uint32_t)RT_CHUNK_WORDS, (npad - chunk) >> 5);
        for (uint32_t w = 0; w < nwords; ++w) {
            uint32_t m[R];
#pragma unroll
            for (int r = 0; r < R; ++r) m[r] = 0u;
            const float4 *sp = s_sph + chunk + (w << 5);
#pragma unroll 16
            for (int k = 0; k < 32; ++k) {
                float4 s = sp[k];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    float b;
                    float disc = sphere_disc(s, ox[r], oy[r], oz[r], dx[r], dy[r], dz[r], a[r], b);
                    // sign bit clear (disc >= +0 or NaN) -> candidate; bit (31-k) of the finished word
                    m[r] = __funnelshift_l(__float_as_uint(disc), m[r], 1);
                }
            }
#pragma unroll
            for (int r = 0; r < R; ++r) {
                s_mask[(w * R + r) * nthreads + threadIdx.x] = ~m[r];
                if (~m[r] != 0u) nz[r] |= 0x80000000u >> w;            // word w holds candidates
            }
        }
        // one candidate per lane per iteration: the warp iterates max-over-lanes of the candidate COUNT
#pragma unroll
        for (int r = 0; r < R; ++r) {
            uint32_t words = live[r] ? nz[r] : 0u, cand = 0, w = 0;
            nz[r] = 0u;
            for (;;) {
                if (cand == 0u) {
                    if (words == 0u) break;
                    w = (uint32_t)__clz(words);
                    words &= ~(0x80000000u >> w);
                    cand = s_mask[(w * R + r) * nthreads + threadIdx.x];
                }
                int bit = __clz(cand);
                cand &= ~(0x80000000u >> bit);
                uint32_t j = chunk + (w << 5) + (uint32_t)bit;
                if (j < n) sphere_candidate(s_sph[j], (int)j, ox[r], oy[r], oz[r], dx[r], dy[r], dz[r], a[r], 0.f, best_t[r], best_i[r]);
            }
        }
    }
}

#if !RT_EXACT
// ------------------------------------------------------------------ closest hit, brute force, throughput build
// Pass 1 is a CONSERVATIVE candidate filter, pass 2 the reference test (sphere_candidate: the IEEE arithmetic of the
// parity build) on the few candidates, so the hit that is returned is the oracle's bit for bit; the filter only has to
// never drop a sphere the reference test would accept.
//
// With u = d/|d|:  disc/a = (oc.u)^2 - (|oc|^2 - r^2) = b^2 + 2 o.c + (r^2 - |c|^2) - |o|^2,  b = o.u - c.u.
// Everything that depends on the ray only or on the sphere only is hoisted, which leaves per ray-sphere
//     3 FFMA (b)  +  1 FADD + 3 FFMA (ncc)  +  1 FFMA.SAT (disc)  +  1 FFMA (mask)   = 9 FMA-pipe issues
// against 11 FP32 + 1 funnel shift for the literal formula -- and on sm_100 a funnel shift (ALU pipe, 16
// lanes) costs two dispatch cycles, an FFMA one (measured, tools/ubench_mix.cu).
//   * All inputs are pre-scaled by powers of two (b' = sigma*b, disc' = S*disc with S = sigma^2 ~ 2^100), so
//     the saturating FMA returns exactly 0 for disc' <= 0 and exactly 1 for any representable positive
//     disc' (>= 1 after scaling): a step function for free.
//   * mask = 2*mask + t (one FFMA) builds an exact 24-bit integer in a float: bit (23-k) <=> sphere k of the group.
//   * Safety margin: the sphere is inflated by eps*(|c|^2 + r^2 + |o|^2), eps = 2^-18, which covers the
//     rounding of this expanded form (<= ~2^-20 (|o|^2+|c|^2)) and of the reference evaluation.
//   * Rays with d = 0 / non-finite d get no candidates (the reference test rejects them: disc = 0 or NaN);
//     rays whose |o| would overflow the scaling take every sphere as a candidate.
struct FastRay { float nux, nuy, nuz, ou, px, py, pz, noo; };

__device__ __forceinline__ FastRay fast_ray_setup(const RtDeviceScene &sc, bool live, float ox, float oy, float oz,
                                                  float dx, float dy, float dz)
{
    FastRay f;
    float len2 = dx * dx + dy * dy + dz * dz;
    float oo = ox * ox + oy * oy + oz * oz;
    bool ok = live && len2 > 0.f && len2 < 3.0e38f && oo < 3.0e38f;
    float inv = rsqrtf(len2);
    float ux = dx * inv, uy = dy * inv, uz = dz * inv;
    float ns = -sc.sigma;
    f.nux = ns * ux; f.nuy = ns * uy; f.nuz = ns * uz;
    f.ou = sc.sigma * (ox * ux + oy * uy + oz * uz);
    float s2 = 2.f * sc.S;
    f.px = s2 * ox; f.py = s2 * oy; f.pz = s2 * oz;
    f.noo = -(sc.S * (1.f - (float)RT_FILTER_EPS)) * oo;
    if (!ok) {                       // no candidates at all
        f.nux = f.nuy = f.nuz = 0.f; f.ou = 0.f; f.px = f.py = f.pz = 0.f;
        f.noo = -__int_as_float(0x7f800000);
    } else if (oo > sc.o_limit2) {   // origin outside the range the scaling was chosen for: everything is a candidate
        f.nux = f.nuy = f.nuz = 0.f; f.px = f.py = f.pz = 0.f; f.noo = 0.f;
        f.ou = __int_as_float(0x7f800000);
    }
    return f;
}

// s_mask layout: [word][ray][thread]
template <int R>
__device__ __forceinline__ void closest_hit_brute_fast(const RtDeviceScene &sc, const float4 *__restrict__ s_sph,
                                                       const float4 *__restrict__ s_filt, uint32_t *__restrict__ s_mask,
                                                       const bool (&live)[R],
                                                       const float (&ox)[R], const float (&oy)[R], const float (&oz)[R],
                                                       const float (&dx)[R], const float (&dy)[R], const float (&dz)[R],
                                                       float (&best_t)[R], int (&best_i)[R])
{
    FastRay f[R];
    uint32_t nz[R];                                  // bit (31 - w): mask word w of the chunk is non-empty
#pragma unroll
    for (int r = 0; r < R; ++r) {
        f[r] = fast_ray_setup(sc, live[r], ox[r], oy[r], oz[r], dx[r], dy[r], dz[r]);
        best_t[r] = FLT_MAX;
        best_i[r] = -1;
        nz[r] = 0u;
    }
    const uint32_t nthreads = blockDim.x;
    const uint32_t n = sc.n, nfilt = sc.nfilt;
    if (n <= RT_SMALL_SCENE) {                       // a handful of spheres: the candidate test on each one costs less than one filter group
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const float a = ray_a(dx[r], dy[r], dz[r]), ra = ray_ra(a);
            if (live[r] && a > 0.f && a < 3.0e38f)
                for (uint32_t j = 0; j < n; ++j)
                    sphere_candidate(s_sph[j], (int)j, ox[r], oy[r], oz[r], dx[r], dy[r], dz[r], a, ra, best_t[r], best_i[r]);
        }
        return;
    }
    for (uint32_t chunk = 0; chunk < nfilt; chunk += RT_FGROUP * RT_FCHUNK_WORDS) {
        const uint32_t nwords = min((uint32_t)RT_FCHUNK_WORDS, (nfilt - chunk) / RT_FGROUP);
        for (uint32_t w = 0; w < nwords; ++w) {
            float acc[R];
#pragma unroll
            for (int r = 0; r < R; ++r) acc[r] = 0.f;
            const float4 *sp = s_filt + chunk + w * RT_FGROUP;
#pragma unroll 8
            for (int k = 0; k < RT_FGROUP; ++k) {
                float4 s = sp[k];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    float b = fmaf(f[r].nuz, s.z, fmaf(f[r].nuy, s.y, fmaf(f[r].nux, s.x, f[r].ou)));
                    float ncc = fmaf(f[r].pz, s.z, fmaf(f[r].py, s.y, fmaf(f[r].px, s.x, s.w + f[r].noo)));
                    float t = __saturatef(fmaf(b, b, ncc));
                    acc[r] = fmaf(acc[r], 2.f, t);           // Horner: the first sphere of the group ends in bit 23
                }
            }
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if (acc[r] != 0.f) {                                     // rare: most groups hold no candidate
                    s_mask[(w * R + r) * nthreads + threadIdx.x] = __float2uint_rz(acc[r]);
                    nz[r] |= 0x80000000u >> w;
                }
            }
        }
        // one candidate per lane per iteration: the warp iterates max-over-lanes of the candidate COUNT
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const float a = ray_a(dx[r], dy[r], dz[r]), ra = ray_ra(a);
            uint32_t words = nz[r], cand = 0, w = 0;
            nz[r] = 0u;
            for (;;) {
                if (cand == 0u) {
                    if (words == 0u) break;
                    w = (uint32_t)__clz(words);
                    words &= ~(0x80000000u >> w);
                    cand = s_mask[(w * R + r) * nthreads + threadIdx.x];
                }
                int lz = __clz(cand);                        // bit 23 (lz = 8) is the first sphere of the group
                cand &= ~(0x80000000u >> lz);
                uint32_t j = chunk + w * RT_FGROUP + (uint32_t)(lz - 8);
                if (j < n) sphere_candidate(s_sph[j], (int)j, ox[r], oy[r], oz[r], dx[r], dy[r], dz[r], a, ra, best_t[r], best_i[r]);
            }
        }
    }
}
#endif  // !RT_EXACT

// ------------------------------------------------------------------ closest hit, LBVH traversal
// Must return what the brute-force loop returns: the closest t in (kMIN, inf) over ALL spheres, ties to the
// lowest sphere index.  Two things make that hold in floating point:
//   * leaves run the reference test per sphere (root selection against FLT_MAX, src/raytracer.hxx:63-89) and
//     merge with the lexicographic rule (t, index), because leaves arrive in tree order, not index order;
//   * boxes are tested CONSERVATIVELY.  The float test accepts a ray when dist^2(line, c) < r^2 + E with
//     E <= ~8 * 2^-24 * |oc|^2, i.e. at a distance up to r + 2^-10.5 |oc|: far spheres are "fatter" for the reference
//     arithmetic than geometrically.  Each child box is therefore inflated per ray by eta * D + pad0, where
//     D is the L1 distance from the ray origin to the box's farthest corner (>= |oc| of every sphere in it),
//     eta = 2^-10, and pad0 covers the rounding of o - c at the scene's scale; the slab interval gets a 1e-5
//     relative slack.  Culling uses tnear <= best_t (not <), so equal-t candidates with a lower index survive.
struct BvhCounters { uint32_t nodes, leaves; };

// Leaf test on a sphere record s = {centre, r*r} at sorted position pos: the reference test with the root chosen against
// tmax = FLT_MAX (exactly src/raytracer.hxx:63-89 as hit_world calls it), merged in (t, index) order because leaves arrive
// in tree order.  The original sphere index is only fetched for a root that can still win.
__device__ __forceinline__ void bvh_leaf_sphere(const RtDeviceScene &sc, uint32_t pos, float4 s, float ox, float oy, float oz,
                                                float dx, float dy, float dz, float a, float &best_t, int &best_i, BvhCounters &cnt)
{
    ++cnt.leaves;
    float b, c;
    float disc = sphere_disc(s, ox, oy, oz, dx, dy, dz, a, b, c);
    if (disc > 0.f && !candidate_decided(disc, b, c, a, best_t)) {
        const float ra = ray_ra(a);
        float sq = __fsqrt_rn(disc);
        float t = div_by_a(__fsub_rn(-b, sq), a, ra);
        bool ok = t < FLT_MAX && t > RT_KMIN;
        if (!ok) { t = div_by_a(__fadd_rn(-b, sq), a, ra); ok = t < FLT_MAX && t > RT_KMIN; }
        if (ok && t <= best_t) {
            const int j = (int)__ldg(sc.bvh_prims + pos);
            if (t < best_t || j < best_i) { best_t = t; best_i = j; }
        }
    }
}
// leaf child `ref` of a node: the sphere comes from the sorted leaf array, or -- RT_BVH_LEAFINNODE -- from the node itself
// (the child slot of a leaf holds the sphere's centre and |r| in place of a box: no second memory round trip)
__device__ __forceinline__ void bvh_leaf(const RtDeviceScene &sc, uint32_t ref, float ncx, float ncy, float ncz, float nr,
                                         float ox, float oy, float oz, float dx, float dy, float dz, float a,
                                         float &best_t, int &best_i, BvhCounters &cnt)
{
    const uint32_t pos = ref & 0x7fffffffu;
#if RT_BVH_LEAFINNODE
    bvh_leaf_sphere(sc, pos, make_float4(ncx, ncy, ncz, nr * nr), ox, oy, oz, dx, dy, dz, a, best_t, best_i, cnt);
#else
    bvh_leaf_sphere(sc, pos, __ldg(sc.bvh_leaf_sph + pos), ox, oy, oz, dx, dy, dz, a, best_t, best_i, cnt);
#endif
}
#define RT_BVH_LEAF_L(n0, n1, n2, ref, ...) bvh_leaf(sc, ref, n0.x, n0.y, n0.z, n0.w, __VA_ARGS__)
#define RT_BVH_LEAF_R(n0, n1, n2, ref, ...) bvh_leaf(sc, ref, n1.z, n1.w, n2.x, n2.y, __VA_ARGS__)

__device__ __forceinline__ bool bvh_box(float lox, float loy, float loz, float hix, float hiy, float hiz,
                                        float ox, float oy, float oz, float ix, float iy, float iz,
                                        float pad0, float best_t, float &tnear)
{
    // offsets of the six planes from the origin, shared by the pad bound and the slab test (their rounding, at most
    // 2^-24 of `far`, disappears in the 2^-10 pad)
    const float dlx = lox - ox, dhx = hix - ox, dly = loy - oy, dhy = hiy - oy, dlz = loz - oz, dhz = hiz - oz;
    float far = fmaxf(fabsf(dlx), fabsf(dhx)) + fmaxf(fabsf(dly), fabsf(dhy)) + fmaxf(fabsf(dlz), fabsf(dhz));
    float pad = fmaf(far, 9.765625e-4f, pad0);
    float t0 = (dlx - pad) * ix, t1 = (dhx + pad) * ix;
    float tn = fminf(t0, t1), tf = fmaxf(t0, t1);
    t0 = (dly - pad) * iy; t1 = (dhy + pad) * iy;
    tn = fmaxf(tn, fminf(t0, t1)); tf = fminf(tf, fmaxf(t0, t1));
    t0 = (dlz - pad) * iz; t1 = (dhz + pad) * iz;
    tn = fmaxf(tn, fminf(t0, t1)); tf = fminf(tf, fmaxf(t0, t1));
    tnear = tn;
    // relative slack on both ends; a negative tnear (origin inside) passes trivially
    const float tn_lo = fmaf(fabsf(tn), -1e-5f, tn);
    return tn_lo <= fmaf(fabsf(tf), 1e-5f, tf) && tf * 1.00001f >= RT_KMIN && tn_lo <= best_t;
}

// The same conservative test on a box stored as centre c, half extent h and hsum = hx + hy + hz (RT_BVH_CENTER):
// the farthest corner is at L1 distance |c - o|_1 + hsum, and the slab interval of an axis is d i -+ |(h + pad) i| with
// d = c - o, i = 1 / direction.  A ray parallel to an axis (i = +-inf) makes that axis' bounds NaN or +-inf; fminf / fmaxf
// drop a NaN, i.e. the axis is ignored -- a superset of the exact answer, which is all a conservative test needs.
__device__ __forceinline__ bool bvh_box_c(float cx, float cy, float cz, float hx, float hy, float hz, float hsum,
                                          float ox, float oy, float oz, float ix, float iy, float iz,
                                          float pad0, float best_t, float &tnear)
{
    const float dx = cx - ox, dy = cy - oy, dz = cz - oz;
    const float far = (fabsf(dx) + fabsf(dy)) + (fabsf(dz) + hsum);
    const float pad = fmaf(far, 9.765625e-4f, pad0);
    const float px = fabsf((hx + pad) * ix), py = fabsf((hy + pad) * iy), pz = fabsf((hz + pad) * iz);
    float tn = fmaf(dx, ix, -px), tf = fmaf(dx, ix, px);
    tn = fmaxf(tn, fmaf(dy, iy, -py)); tf = fminf(tf, fmaf(dy, iy, py));
    tn = fmaxf(tn, fmaf(dz, iz, -pz)); tf = fminf(tf, fmaf(dz, iz, pz));
    tnear = tn;
    const float tn_lo = fmaf(fabsf(tn), -1e-5f, tn);
    return tn_lo <= fmaf(fabsf(tf), 1e-5f, tf) && tf * 1.00001f >= RT_KMIN && tn_lo <= best_t;
}
// both child boxes of a node (n0..n2 = the first three float4 of the node, hs = the two L1 half extents in centre form)
#if RT_BVH_CENTER
#define RT_BVH_BOX_L(n0, n1, n2, n3, ...) bvh_box_c(n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, n3.z, __VA_ARGS__)
#define RT_BVH_BOX_R(n0, n1, n2, n3, ...) bvh_box_c(n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, n3.w, __VA_ARGS__)
#else
#define RT_BVH_BOX_L(n0, n1, n2, n3, ...) bvh_box(n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, __VA_ARGS__)
#define RT_BVH_BOX_R(n0, n1, n2, n3, ...) bvh_box(n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, __VA_ARGS__)
#endif

// keep_best: start from the caller's (best_t, best_i) -- a hit that was already found is a valid bound, the result is the same minimum
__device__ __forceinline__ void closest_hit_lbvh(const RtDeviceScene &sc, bool live, float ox, float oy, float oz,
                                                 float dx, float dy, float dz, float &best_t, int &best_i, BvhCounters &cnt,
                                                 bool keep_best = false)
{
    if (!keep_best) {
        best_t = FLT_MAX;
        best_i = -1;
    }
    const float a = ray_a(dx, dy, dz);
    if (!live || !(a > 0.f) || !(a < 3.0e38f)) return;      // d = 0 or non-finite: the reference test finds nothing
    const float ix = 1.f / dx, iy = 1.f / dy, iz = 1.f / dz;
    uint32_t stack[96];
    int sp = 0;
    uint32_t node = sc.bvh_root;
    for (;;) {
        const float4 *np = sc.bvh_nodes + 4 * (size_t)node;
        float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3);
        ++cnt.nodes;
        uint32_t left = __float_as_uint(n3.x), right = __float_as_uint(n3.y);
        float tl, tr;
        bool hl = RT_BVH_BOX_L(n0, n1, n2, n3, ox, oy, oz, ix, iy, iz, sc.bvh_pad0, best_t, tl);
        bool hr = RT_BVH_BOX_R(n0, n1, n2, n3, ox, oy, oz, ix, iy, iz, sc.bvh_pad0, best_t, tr);
        if (hl && (left & 0x80000000u)) { RT_BVH_LEAF_L(n0, n1, n2, left, ox, oy, oz, dx, dy, dz, a, best_t, best_i, cnt); hl = false; }
        if (hr && (right & 0x80000000u)) { RT_BVH_LEAF_R(n0, n1, n2, right, ox, oy, oz, dx, dy, dz, a, best_t, best_i, cnt); hr = false; }
        if (hl && hr) {                                       // both inner: near child first, far child on the stack
            bool left_first = tl <= tr;
            stack[sp++] = left_first ? right : left;
            node = left_first ? left : right;
        } else if (hl) node = left;
        else if (hr) node = right;
        else {
            if (sp == 0) break;
            node = stack[--sp];
        }
    }
}

// ------------------------------------------------------------------ warp work pool
// A warp takes RT_BATCH consecutive work items (pixel, sample) with one atomic and deals them to the
// lanes whose path ended, by ballot + popc prefix -- terminated paths are replaced immediately, so every
// lane carries a live ray into the next sphere loop (compaction by regeneration).
// next/end: the warp's current batch of work items.  Queue mode packs `end` as tile << 9 | items left in the batch (a batch
// is RT_BATCH = 256 items of ONE tile), so that the device-side queue costs the render loops no extra register.
struct WarpPool { uint32_t next, end; bool exhausted; };

// ---- device-side tile queue (multi-GPU): acquire the next tile for this GPU -- own range first, then steal.
// Range r = tiles [r T / N, (r+1) T / N) of the frame's (interleaved) tile list; cursors live in rank 0's memory.
#ifndef RT_QUEUE_FETCH_NOINLINE
#define RT_QUEUE_FETCH_NOINLINE 0    // measured (1M-sphere LBVH through the queue, 256 spp): inline 853, out of line 806 Msamples/s (tile-list order: 890)
#endif
#if RT_QUEUE_FETCH_NOINLINE
__device__ __noinline__             // once per tile: kept out of the render loops
#else
__device__ __forceinline__
#endif
uint32_t queue_fetch(const RtRenderArgs &A)
{
    if (*reinterpret_cast<volatile uint32_t *>(&A.feed->drained)) return RT_FEED_NONE;
    for (uint32_t d = 0; d < A.q_nranks; ++d) {
        const uint32_t v = (A.q_rank + d) % A.q_nranks;
        const uint32_t hi = ((v + 1u) * A.q_ntiles) / A.q_nranks;          // ntiles <= 2^16, nranks <= 64: 32 bits are enough
        uint32_t *cur = A.q_cursors + v * RT_QUEUE_STRIDE;
        if (d > 0u && *reinterpret_cast<volatile uint32_t *>(cur) >= hi) continue;     // a dry victim: do not bump it again
        const uint32_t t = atomicAdd_system(cur, 1u);
        if (t < hi) return t;
    }
    A.feed->drained = 1u;
    return RT_FEED_NONE;
}
// tile of feed slot k (called by ONE lane): the first warp to reach a slot talks to the queue, the others wait for it
__device__ __forceinline__ uint32_t feed_slot_tile(const RtRenderArgs &A, uint32_t k)
{
    uint32_t *slot = A.feed->slot + k;
    uint32_t v = *reinterpret_cast<volatile uint32_t *>(slot);
    if (v == RT_FEED_EMPTY) {
        v = atomicCAS(slot, RT_FEED_EMPTY, RT_FEED_FETCHING);
        if (v == RT_FEED_EMPTY) {
            v = queue_fetch(A);
            unsigned long long now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            A.feed->slot[2u * A.feed_slots + k] = (uint32_t)(now >> 10);      // acquisition time, ~microseconds (rt_queue_trace)
            __threadfence();
            *reinterpret_cast<volatile uint32_t *>(slot) = v;
            return v;
        }
    }
    while (v == RT_FEED_FETCHING) {
        __nanosleep(200);
        v = *reinterpret_cast<volatile uint32_t *>(slot);
    }
    return v;
}

// Queue mode, lane 0: the next batch of work -> tile << 32 | first item within the tile (tile >= RT_FEED_NONE: no more work).
// Slots are opened in order, so a dry slot means that everything behind it is dry too.
template <uint32_t BATCH>
__device__ __forceinline__ unsigned long long queue_next_batch(const RtRenderArgs &A)
{
    uint32_t *ctr = A.feed->slot + A.feed_slots;
    for (;;) {
        const uint32_t k = *reinterpret_cast<volatile uint32_t *>(&A.feed->cur);
        const uint32_t t = (k < A.feed_slots) ? feed_slot_tile(A, k) : RT_FEED_NONE;
        if (t >= RT_FEED_NONE) return (unsigned long long)RT_FEED_NONE << 32;
        const uint32_t j = atomicAdd(ctr + k, BATCH);
        if (j < A.TPS) return ((unsigned long long)t << 32) | j;
        atomicCAS(&A.feed->cur, k, k + 1u);                 // this tile is handed out completely: open the next slot
    }
}

// Next batch of work items for the warp.  Returns false when there is no more work.
// QUEUE (template parameter of every render kernel): the work domain is the device-side tile queue instead of the
// launch's own tile list; separate instantiations, so the single-GPU kernels carry none of the queue code.
// BATCH: work items a warp takes at once (RT_BATCH = 256; a template parameter so that a kernel can choose: RT_BVH_BATCH).
template <bool QUEUE, uint32_t BATCH>
__device__ __forceinline__ bool pool_refill(WarpPool &p, const RtRenderArgs &A)
{
    const uint32_t lane = threadIdx.x & 31u;
    if constexpr (QUEUE) {
        // a batch lies inside ONE tile (TPS is a multiple of RT_BATCH)
        unsigned long long tj = 0ull;
        if (lane == 0) tj = queue_next_batch<BATCH>(A);
        const uint32_t t = __shfl_sync(RT_FULL, (uint32_t)(tj >> 32), 0);
        const uint32_t j = __shfl_sync(RT_FULL, (uint32_t)tj, 0);
        if (t >= RT_FEED_NONE) return false;
        RT_DASSERT(t < A.q_ntiles && t < (1u << 23) && j + BATCH <= A.TPS);
        p.next = j; p.end = (t << 9) | BATCH;
        return true;
    } else {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(A.work_counter, BATCH);
        base = __shfl_sync(RT_FULL, base, 0);
        if (base >= A.total_items) return false;
        p.next = base;
        p.end = min(base + BATCH, A.total_items);
        return true;
    }
}

// Work items for the lanes that want one; `tile` = the item's tile in queue mode (items are then tile-relative).
template <bool QUEUE, uint32_t BATCH>
__device__ __forceinline__ uint32_t pool_take(WarpPool &p, bool want, const RtRenderArgs &A, uint32_t &tile)
{
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t item = 0xffffffffu;
    uint32_t need = __ballot_sync(RT_FULL, want);
    while (need) {
        if (QUEUE ? (p.end & 511u) == 0u : p.next >= p.end) {
            if (p.exhausted) break;
            if (!pool_refill<QUEUE, BATCH>(p, A)) { p.exhausted = true; break; }
        }
        const uint32_t avail = QUEUE ? (p.end & 511u) : p.end - p.next;
        bool mine = want && item == 0xffffffffu;
        uint32_t rank = __popc(need & lt);
        if (mine && rank < avail) { item = p.next + rank; if (QUEUE) tile = p.end >> 9; }
        const uint32_t adv = min((uint32_t)__popc(need), avail);
        p.next += adv;
        if (QUEUE) p.end -= adv;
        need = __ballot_sync(RT_FULL, want && item == 0xffffffffu);
    }
    return item;
}

// x / d for 32-bit x with M = floor(2^32 / d) (d >= 2; host-computed): the estimate is at most one short.
__device__ __forceinline__ uint32_t div_magic(uint32_t x, uint32_t d, uint32_t M, uint32_t &rem)
{
    uint32_t q = __umulhi(x, M);
    uint32_t r = x - q * d;
    if (r >= d) { ++q; r -= d; }
    rem = r;
    return q;
}

// ------------------------------------------------------------------ the persistent render kernel
enum { ST_EMPTY = 0, ST_ACTIVE = 1, ST_DEAD = 2 };
enum { RM_NONE = 0, RM_UNIFORM = 1, RM_BALL = 2, RM_CAMERA = 3 };

struct PathSlot {
    float ox, oy, oz, dx, dy, dz;   // current ray (direction NOT normalised, as in the reference)
    float ar, ag, ab;               // attenuation
    uint32_t pix, smp, bounce;      // global pixel index (y*W+x), sample index, segments traced so far
    int st;
};

struct ThreadStats { uint32_t samples, segments, sky, absorbed, capped; BvhCounters bvh; };

__device__ __forceinline__ unsigned long long to_fixed(float c)
{
    return (c > 0.f) ? __float2ull_rn(c * RT_FIXED_ONE) : 0ull;
}

// One shading pass for one path slot: finish the segment just traced (background / material), hand a new
// camera sample to the slot if its path ended, draw the random numbers every lane needs in ONE shared
// Philox loop, then build the next ray.  Must be called by all 32 lanes (pool_take is warp-cooperative).
template <bool LBVH, bool QUEUE = false, bool PAR = RT_PAR_DEFAULT, uint32_t BATCH = RT_BATCH>
__device__ __forceinline__ void shade_slot(const RtRenderArgs &A, const float4 *__restrict__ s_sph, PathSlot &S,
                                           float best_t, int best_i, WarpPool &pool, ThreadStats &ts, bool hold = false,
                                           const float *__restrict__ s_rad = nullptr, const uint32_t *__restrict__ s_meta = nullptr,
                                           const float4 *__restrict__ s_mat = nullptr)
{
    // s_rad / s_meta / s_mat: shared-memory copies of the radii, material indices and materials when the kernel staged
    // them (two dependent L2 round trips per hit otherwise)
    // hold: this lane's closest-hit query is still running (resumable LBVH traversal); it only takes part in the
    // warp-cooperative work pool and leaves its path untouched
    const RtDeviceScene &sc = A.sc;
    bool want = !hold && (S.st != ST_ACTIVE);
    int mode = RM_NONE;
    int kind = -1;
    f3 p = mk3(0, 0, 0), n = mk3(0, 0, 0), unit = mk3(0, 0, 0), refl = mk3(0, 0, 0), refr = mk3(0, 0, 0);
    float4 m = make_float4(0, 0, 0, 0);
    float prob = 1.f;

    if (!hold && S.st == ST_ACTIVE) {
        ++ts.segments;
        unit = normalize3<PAR>(mk3(S.dx, S.dy, S.dz));                   // ray::unit_direction, src/math.hxx:351
        if (best_i < 0) {                                                // miss: src/main.cxx:71 (quirk F6: + 1.f)
            float t = .5f * unit.y + (A.book ? .5f : 1.f);              // book optics: the usual [0,1] blend
            float omt = 1.f - t;
            float cr = (omt + .5f * t) * S.ar, cg = (omt + .7f * t) * S.ag, cb = (omt + 1.f * t) * S.ab;
            unsigned long long *px = A.accum + 4ull * S.pix;
            atomicAdd(px + 0, to_fixed(cr));
            atomicAdd(px + 1, to_fixed(cg));
            atomicAdd(px + 2, to_fixed(cb));
            ++ts.sky;
            want = true;
        } else if (S.bounce + 1u >= A.depth) {                           // depth exhausted -> black, src/main.cxx:74
            ++ts.capped;
            want = true;
        } else {
            RT_DASSERT(best_i < (int)sc.n);
            float4 s = LBVH ? __ldg(sc.sph + best_i) : s_sph[best_i];
            float rad = s_rad ? s_rad[best_i] : __ldg(sc.sph_rad + best_i);
            uint32_t meta = s_meta ? s_meta[best_i] : __ldg(sc.sph_meta + best_i);
            m = s_mat ? s_mat[meta & 0x3fffffffu] : __ldg(sc.mat + (meta & 0x3fffffffu));
            kind = (int)(meta >> 30);
            p = mk3(S.ox + S.dx * best_t, S.oy + S.dy * best_t, S.oz + S.dz * best_t);   // ray::point_at
            if constexpr (PAR) {
                n = mk3((p.x - s.x) / rad, (p.y - s.y) / rad, (p.z - s.z) / rad);        // src/raytracer.hxx:70
            } else {
                float inv_rad = __frcp_rn(rad);
                n = mk3((p.x - s.x) * inv_rad, (p.y - s.y) * inv_rad, (p.z - s.z) * inv_rad);
            }
            if (kind == 0) mode = RM_BALL;
            else if (kind == 1) { refl = reflect3(unit, n); mode = RM_BALL; }
            else {                                                        // dielectric, src/raytracer.hxx:158-178
                f3 outward = mk3(-n.x, -n.y, -n.z);
                float ri = m.w;
                float cosine = dot3(unit, n);
                if (cosine <= 0.f) {
                    outward = mk3(outward.x * -1.f, outward.y * -1.f, outward.z * -1.f);
                    ri = 1.f / ri;
                    cosine = cosine * -1.f;
                }
                refr = A.book ? refract3_book(unit, outward, ri) : refract3(unit, outward, ri);
                if (refr.x * refr.x + refr.y * refr.y + refr.z * refr.z > 0.f) prob = schlick<PAR>(ri, cosine);
                refl = reflect3(unit, n);
                mode = RM_UNIFORM;
            }
        }
    }

    // ---- regeneration: give every ended path a new (pixel, sample)
    uint32_t px = 0, py = 0;
    uint32_t ev = S.bounce + 1u;
    for (bool need = want;;) {
        uint32_t qtile = 0u;
        uint32_t item = pool_take<QUEUE, BATCH>(pool, need, A, qtile);
        if (need) {
            if (item != 0xffffffffu) {
                // item order: tile-major, then sample, then pixel -- a tile's accumulators (TP * 32 B) stay
                // cache-resident while all of its samples are traced.  Queue mode: the item is already tile-relative.
                uint32_t q, r, rx;
                uint32_t tl = QUEUE ? (q = item, qtile) : (A.TPS > 1u) ? div_magic(item, A.TPS, A.mTPS, q) : (q = 0u, item);
                uint32_t s_i = (A.TP > 1u) ? div_magic(q, A.TP, A.mTP, r) : (r = 0u, q);
                uint4 T = A.tiles ? __ldg(A.tiles + A.t0 + tl) : A.tile0;
                RT_DASSERT(s_i < A.spp && r < A.TP && T.x + T.z <= A.W && T.y + T.w <= A.H);
                if (r < T.z * T.w) {                                      // edge tiles leave slots unused
                    uint32_t ty;
                    if (!A.tiles && T.z > 1u) ty = div_magic(r, T.z, A.mW0, rx);
                    else { ty = r / T.z; rx = r - ty * T.z; }
                    px = T.x + rx;
                    py = T.y + ty;
                    S.pix = py * A.W + px;
                    S.smp = A.sample_base + s_i;
                    S.bounce = 0;
                    S.ar = S.ag = S.ab = 1.f;
                    S.st = ST_ACTIVE;
                    ev = 0;
                    mode = RM_CAMERA;
                    ++ts.samples;
                    need = false;
                }
            } else {
                S.st = ST_EMPTY;
                need = false;
            }
        }
        if (!__any_sync(RT_FULL, need)) break;
    }

    float r0 = 0.f, r1 = 0.f, r2 = 0.f, ju = 0.f, jv = 0.f;
    const bool camera_lane = (mode == RM_CAMERA);
    if constexpr (PAR) {
    // ---- parity shading: one Philox loop serves camera jitter, lens / lambert / metal ball samples (the reference's
    // rejection loop, src/raytracer.hxx:32-43, one block per attempt) and the dielectric coin
    uint32_t blk = 0;
    while (mode != RM_NONE) {
        uint4 w = philox4x32_10(S.pix, S.smp, ev, blk, A.key0, A.key1);
        if (mode == RM_CAMERA) {
            ju = u01(w.x); jv = u01(w.y);
            blk = 1;
            mode = (A.cam.lens_radius != 0.f) ? RM_BALL : RM_NONE;
        } else if (mode == RM_UNIFORM) {
            r0 = u01(w.x);
            mode = RM_NONE;
        } else {                                                          // src/raytracer.hxx:32-43
            r0 = s11(w.x); r1 = s11(w.y); r2 = s11(w.z);
            if (camera_lane && A.book) r2 = 0.f;                          // book optics: the lens sample is a DISK point
            bool reject = sqrtf(r0 * r0 + r1 * r1 + r2 * r2) > 1.f;
            if (reject) ++blk; else mode = RM_NONE;
        }
    }
    } else {
    // ---- throughput shading: ONE Philox block per lane and no rejection loop.  random_in_unit_sphere
    // (src/raytracer.hxx:32-43) returns a point uniform in the unit ball; the same distribution is drawn by inversion
    // (z uniform in [-1,1), azimuth uniform, radius U^(1/3)), which costs one block instead of a loop that runs for
    // the slowest lane of the warp (mean 1.9 attempts per lane, ~5.5 for the maximum over 32 lanes).  The lens offset
    // uses only x and y of the ball point (src/camera.hxx:51-53); their joint law is rho = sqrt(1 - U^(2/3)) with a
    // uniform azimuth, so the camera ray needs the two words the pixel jitter leaves over.
    // The parity shading above keeps the reference's rejection loop word for word; the two are compared statistically.
    if (mode != RM_NONE) {
        const uint4 w = philox4x32_10(S.pix, S.smp, ev, 0u, A.key0, A.key1);
        if (camera_lane) {
            ju = u01(w.x); jv = u01(w.y);
            if (A.book) disk_by_inversion(w.z, w.w, r0, r1);
            else ball_xy_by_inversion(w.z, w.w, r0, r1);
        } else {
            ball_by_inversion(w.x, w.y, w.z, r0, r1, r2);
            if (mode == RM_UNIFORM) r0 = u01(w.x);                        // the dielectric coin is word 0, as in the parity build
        }
    }
    }

    // ---- next ray
    if (camera_lane) {                                                    // src/main.cxx:192-200, src/camera.hxx:46-57
        const RtDeviceCamera &c = A.cam;
        float u = (float)px / (float)A.W + ju / (float)A.W;
        float v = (float)py / (float)A.H + jv / (float)A.H;
        float rdx = r0 * c.lens_radius, rdy = r1 * c.lens_radius;
        float offx = u * rdx, offy = v * rdy, offz = 0.f;
        float omv = 1.f - v;
        if (A.book) {       // book optics: lens offset in the camera basis, direction measured from the eye (fixes F3, F4)
            offx = c.bux * rdx + c.bvx * rdy; offy = c.buy * rdx + c.bvy * rdy; offz = c.buz * rdx + c.bvz * rdy;
        }
        S.ox = c.ox + offx; S.oy = c.oy + offy; S.oz = c.oz + offz;
        S.dx = ((c.llx + c.hx * u) + c.vx * omv) - offx;
        S.dy = ((c.lly + c.hy * u) + c.vy * omv) - offy;
        S.dz = ((c.llz + c.hz * u) + c.vz * omv) - offz;
        if (A.book) { S.dx = S.dx - c.ox; S.dy = S.dy - c.oy; S.dz = S.dz - c.oz; }
    } else if (kind >= 0) {
        bool scattered = true;
        f3 nd;
        if (kind == 0) {                                                  // lambert, src/raytracer.hxx:128-140
            nd = mk3(((p.x + n.x) + r0) - p.x, ((p.y + n.y) + r1) - p.y, ((p.z + n.z) + r2) - p.z);
        } else if (kind == 1) {                                           // metal, src/raytracer.hxx:142-156
            nd = mk3(refl.x + r0 * m.w, refl.y + r1 * m.w, refl.z + r2 * m.w);
            scattered = dot3(nd, n) > 0.f;
        } else {                                                          // dielectric, src/raytracer.hxx:180-192
            nd = (r0 < prob) ? refl : refr;
        }
        if (scattered) {
            S.ox = p.x; S.oy = p.y; S.oz = p.z;
            S.dx = nd.x; S.dy = nd.y; S.dz = nd.z;
            S.ar *= m.x; S.ag *= m.y; S.ab *= m.z;                        // src/main.cxx:65
            ++S.bounce;
        } else {
            ++ts.absorbed;                                                // returns black, src/main.cxx:68
            S.st = ST_DEAD;                                               // regenerated in the next pass
        }
    }
}

#if !RT_EXACT
// Sampler dump for the distribution tests (tests/test_gpu_parity.py): point i of the ball sampler (lens = 0) or of the
// lens sampler (lens = 1: x, y, 0) exactly as shade_slot draws them, from Philox block (i, 0, event 1, 0).
__global__ void k_sample_ball(float *out, unsigned long long n, uint32_t key0, uint32_t key1, int lens)
{
    unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 w = philox4x32_10((uint32_t)i, (uint32_t)(i >> 32), 1u, 0u, key0, key1);
    float x = 0.f, y = 0.f, z = 0.f;
    if (lens) ball_xy_by_inversion(w.z, w.w, x, y);
    else ball_by_inversion(w.x, w.y, w.z, x, y, z);
    out[3 * i + 0] = x; out[3 * i + 1] = y; out[3 * i + 2] = z;
}
#endif

template <bool LBVH, bool CHECK = false, bool QUEUE = false>
__global__ void __launch_bounds__(RT_THREADS, 3) k_render(const __grid_constant__ RtRenderArgs A)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4 *s_sph = reinterpret_cast<float4 *>(smem_raw + 128);
    uint32_t *s_mask = nullptr;
#if !RT_EXACT
    float4 *s_filt = nullptr;
#endif
    if (!LBVH) {
        uint64_t *bar = reinterpret_cast<uint64_t *>(smem_raw);
#if RT_EXACT
        s_mask = reinterpret_cast<uint32_t *>(smem_raw + 128 + (size_t)A.sc.npad * 16u);
        stage_spheres(s_sph, A.sc.sph, nullptr, nullptr, A.sc.npad, 0, bar);
#else
        s_filt = s_sph + A.sc.npad;
        s_mask = reinterpret_cast<uint32_t *>(smem_raw + 128 + ((size_t)A.sc.npad + A.sc.nfilt) * 16u);
        stage_spheres(s_sph, A.sc.sph, s_filt, A.sc.filt, A.sc.npad, A.sc.nfilt, bar);
#endif
    }

    PathSlot S[RT_RPT];
#pragma unroll
    for (int r = 0; r < RT_RPT; ++r) {
        S[r].st = ST_EMPTY; S[r].pix = 0; S[r].smp = 0; S[r].bounce = 0;
        S[r].ox = S[r].oy = S[r].oz = 0.f; S[r].dx = S[r].dy = S[r].dz = 0.f;
        S[r].ar = S[r].ag = S[r].ab = 1.f;
    }
    WarpPool pool; pool.next = 0; pool.end = 0; pool.exhausted = false;
    ThreadStats ts = {0, 0, 0, 0, 0, {0, 0}};
    float best_t[RT_RPT]; int best_i[RT_RPT];
#pragma unroll
    for (int r = 0; r < RT_RPT; ++r) { best_t[r] = FLT_MAX; best_i[r] = -1; }

    for (;;) {
        bool any_live = false;
#pragma unroll
        for (int r = 0; r < RT_RPT; ++r) {
            shade_slot<LBVH, QUEUE>(A, s_sph, S[r], best_t[r], best_i[r], pool, ts);
            any_live |= (S[r].st != ST_EMPTY);
        }
        if (!__any_sync(RT_FULL, any_live)) break;

        if (LBVH) {
#pragma unroll
            for (int r = 0; r < RT_RPT; ++r)
                closest_hit_lbvh(A.sc, S[r].st == ST_ACTIVE, S[r].ox, S[r].oy, S[r].oz, S[r].dx, S[r].dy, S[r].dz,
                                 best_t[r], best_i[r], ts.bvh);
        } else {
            bool live[RT_RPT];
            float ox[RT_RPT], oy[RT_RPT], oz[RT_RPT], dx[RT_RPT], dy[RT_RPT], dz[RT_RPT];
#pragma unroll
            for (int r = 0; r < RT_RPT; ++r) {
                live[r] = (S[r].st == ST_ACTIVE);
                ox[r] = S[r].ox; oy[r] = S[r].oy; oz[r] = S[r].oz;
                dx[r] = S[r].dx; dy[r] = S[r].dy; dz[r] = S[r].dz;
            }
#if RT_EXACT
            closest_hit_brute<RT_RPT>(s_sph, A.sc.n, A.sc.npad, s_mask, live, ox, oy, oz, dx, dy, dz, best_t, best_i);
#else
            closest_hit_brute_fast<RT_RPT>(A.sc, s_sph, s_filt, s_mask, live, ox, oy, oz, dx, dy, dz, best_t, best_i);
            if (CHECK) {                            // RT_FLAG_FILTER_CHECK: the same test over all spheres must agree
#pragma unroll
                for (int r = 0; r < RT_RPT; ++r) {
                    if (!live[r]) continue;
                    const float a = ray_a(dx[r], dy[r], dz[r]);
                    float ct = FLT_MAX; int ci = -1;
                    for (uint32_t j = 0; j < A.sc.n; ++j)
                        sphere_candidate_ref(s_sph[j], (int)j, ox[r], oy[r], oz[r], dx[r], dy[r], dz[r], a, ct, ci);
                    if (ci != best_i[r] || ct != best_t[r]) atomicAdd(A.stats + 7, 1ull);
                }
            }
#endif
        }
    }

    // per-warp reduction of the path statistics, one atomic per counter per warp
    uint32_t v[7] = {ts.samples, ts.segments, ts.sky, ts.absorbed, ts.capped, ts.bvh.nodes, ts.bvh.leaves};
#pragma unroll
    for (int k = 0; k < 7; ++k) {
        uint32_t x = v[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(RT_FULL, x, o);
        if ((threadIdx.x & 31u) == 0 && x) {
            atomicAdd(A.stats + k, (unsigned long long)x);
            if (k == 1 && !LBVH) atomicAdd(A.stats + 6, (unsigned long long)x * A.sc.n);   // sphere tests
        }
    }
}

#if !RT_EXACT
// ------------------------------------------------------------------ small scenes (throughput build, n <= RT_SMALL_SCENE)
// With a handful of spheres the closest hit is a few dozen instructions and the kernel is all shading.  One path per
// thread and no filter code keeps the loop body small: the general kernel (two paths per thread, filter + candidate
// loops inlined) stalled 2.1 of 9.4 cycles per issue on instruction fetch (ncu: stalled_no_instruction) on the
// 5-sphere scene.  Spheres, radii, material indices and materials all live in shared memory.
#ifndef RT_SMALL_CTAS
#define RT_SMALL_CTAS 4
#endif
template <bool QUEUE>
__global__ void __launch_bounds__(RT_THREADS, RT_SMALL_CTAS) k_render_small(const __grid_constant__ RtRenderArgs A)
{
    __shared__ float4 s_sph[RT_SMALL_SCENE];
    __shared__ float4 s_mat[RT_SMALL_SCENE];
    __shared__ float s_rad[RT_SMALL_SCENE];
    __shared__ uint32_t s_meta[RT_SMALL_SCENE];
    const uint32_t n = A.sc.n;
    const bool mats = A.sc.m <= RT_SMALL_SCENE;
    if (threadIdx.x < n) {
        s_sph[threadIdx.x] = __ldg(A.sc.sph + threadIdx.x);
        s_rad[threadIdx.x] = __ldg(A.sc.sph_rad + threadIdx.x);
        s_meta[threadIdx.x] = __ldg(A.sc.sph_meta + threadIdx.x);
    }
    if (mats && threadIdx.x < A.sc.m) s_mat[threadIdx.x] = __ldg(A.sc.mat + threadIdx.x);
    __syncthreads();

    PathSlot S;
    S.st = ST_EMPTY; S.pix = 0; S.smp = 0; S.bounce = 0;
    S.ox = S.oy = S.oz = 0.f; S.dx = S.dy = S.dz = 0.f;
    S.ar = S.ag = S.ab = 1.f;
    WarpPool pool; pool.next = 0; pool.end = 0; pool.exhausted = false;
    ThreadStats ts = {0, 0, 0, 0, 0, {0, 0}};
    float best_t = FLT_MAX; int best_i = -1;
    for (;;) {
        shade_slot<false, QUEUE>(A, s_sph, S, best_t, best_i, pool, ts, false, s_rad, s_meta, mats ? s_mat : nullptr);
        if (!__any_sync(RT_FULL, S.st != ST_EMPTY)) break;
        best_t = FLT_MAX; best_i = -1;
        const float a = ray_a(S.dx, S.dy, S.dz), ra = ray_ra(a);
        if (S.st == ST_ACTIVE && a > 0.f && a < 3.0e38f) {
            for (uint32_t j = 0; j < n; ++j)
                sphere_candidate(s_sph[j], (int)j, S.ox, S.oy, S.oz, S.dx, S.dy, S.dz, a, ra, best_t, best_i);
        }
    }
    uint32_t v[5] = {ts.samples, ts.segments, ts.sky, ts.absorbed, ts.capped};
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        uint32_t x = v[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(RT_FULL, x, o);
        if ((threadIdx.x & 31u) == 0 && x) {
            atomicAdd(A.stats + k, (unsigned long long)x);
            if (k == 1) atomicAdd(A.stats + 6, (unsigned long long)x * n);   // sphere tests
        }
    }
}
#endif

// ------------------------------------------------------------------ parity pass helpers (k_hits, k_hits_tc)
// ray i of the pass: caller-supplied, or the primary ray of pixel i of the tile: u = x/W, v = y/H, lens offset 0
// (src/main.cxx:192-195 without jitter, src/camera.hxx:56)
__device__ __forceinline__ void hits_ray(const RtHitArgs &A, unsigned long long i, float &ox, float &oy, float &oz,
                                         float &dx, float &dy, float &dz)
{
    if (A.rays) {
        const float *r = A.rays + 6ull * i;
        ox = r[0]; oy = r[1]; oz = r[2]; dx = r[3]; dy = r[4]; dz = r[5];
    } else {
        uint32_t ty = (uint32_t)(i / A.w), tx = (uint32_t)(i - (unsigned long long)ty * A.w);
        float u = __fdiv_rn((float)(A.x0 + tx), (float)A.W);
        float v = __fdiv_rn((float)(A.y0 + ty), (float)A.H);
        float omv = __fsub_rn(1.f, v);
        const RtDeviceCamera &c = A.cam;
        ox = c.ox; oy = c.oy; oz = c.oz;
        dx = __fadd_rn(__fadd_rn(c.llx, __fmul_rn(c.hx, u)), __fmul_rn(c.vx, omv));
        dy = __fadd_rn(__fadd_rn(c.lly, __fmul_rn(c.hy, u)), __fmul_rn(c.vy, omv));
        dz = __fadd_rn(__fadd_rn(c.llz, __fmul_rn(c.hz, u)), __fmul_rn(c.vz, omv));
    }
}
// hit record of src/raytracer.hxx:66-72: index, t, normal = (o + d t - c) / r (s = the sphere record when bi >= 0)
__device__ __forceinline__ void hits_store(const RtHitArgs &A, unsigned long long i, int bi, float bt, float4 s,
                                           float ox, float oy, float oz, float dx, float dy, float dz)
{
    A.idx[i] = bi;
    if (bi >= 0) {
        float rad = __ldg(A.sc.sph_rad + bi);
        float px = __fadd_rn(ox, __fmul_rn(dx, bt)), py = __fadd_rn(oy, __fmul_rn(dy, bt)), pz = __fadd_rn(oz, __fmul_rn(dz, bt));
        A.t[i] = bt;
        A.normal[3 * i + 0] = __fdiv_rn(__fsub_rn(px, s.x), rad);
        A.normal[3 * i + 1] = __fdiv_rn(__fsub_rn(py, s.y), rad);
        A.normal[3 * i + 2] = __fdiv_rn(__fsub_rn(pz, s.z), rad);
    } else {
        A.t[i] = 0.f;
        A.normal[3 * i + 0] = 0.f; A.normal[3 * i + 1] = 0.f; A.normal[3 * i + 2] = 0.f;
    }
}

#include "rt_tensor.cuh"
#include "rt_bvh_render.cuh"

// ------------------------------------------------------------------ parity pass: primary / caller-supplied rays
// One thread per ray; the same two-pass brute-force closest hit (or the LBVH) as the render kernels of this build.
template <bool LBVH>
__global__ void __launch_bounds__(RT_THREADS) k_hits(const __grid_constant__ RtHitArgs A)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4 *s_sph = reinterpret_cast<float4 *>(smem_raw + 128);
    uint32_t *s_mask = nullptr;
#if !RT_EXACT
    float4 *s_filt = nullptr;
#endif
    if (!LBVH) {
        uint64_t *bar = reinterpret_cast<uint64_t *>(smem_raw);
#if RT_EXACT
        s_mask = reinterpret_cast<uint32_t *>(smem_raw + 128 + (size_t)A.sc.npad * 16u);
        stage_spheres(s_sph, A.sc.sph, nullptr, nullptr, A.sc.npad, 0, bar);
#else
        s_filt = s_sph + A.sc.npad;
        s_mask = reinterpret_cast<uint32_t *>(smem_raw + 128 + ((size_t)A.sc.npad + A.sc.nfilt) * 16u);
        stage_spheres(s_sph, A.sc.sph, s_filt, A.sc.filt, A.sc.npad, A.sc.nfilt, bar);
#endif
    }

    const unsigned long long total = A.rays ? A.nrays : (unsigned long long)A.w * A.h;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    const unsigned long long iters = (total + stride - 1) / stride;
    BvhCounters cnt = {0, 0};
    for (unsigned long long it = 0; it < iters; ++it) {
        unsigned long long i = it * stride + (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
        bool live[1] = {i < total};
        float ox[1] = {0.f}, oy[1] = {0.f}, oz[1] = {0.f}, dx[1] = {0.f}, dy[1] = {0.f}, dz[1] = {0.f};
        if (live[0]) hits_ray(A, i, ox[0], oy[0], oz[0], dx[0], dy[0], dz[0]);
        float bt[1]; int bi[1];
        if (LBVH) {
            closest_hit_lbvh(A.sc, live[0], ox[0], oy[0], oz[0], dx[0], dy[0], dz[0], bt[0], bi[0], cnt);
        } else {
#if RT_EXACT
            closest_hit_brute<1>(s_sph, A.sc.n, A.sc.npad, s_mask, live, ox, oy, oz, dx, dy, dz, bt, bi);
#else
            closest_hit_brute_fast<1>(A.sc, s_sph, s_filt, s_mask, live, ox, oy, oz, dx, dy, dz, bt, bi);
#endif
        }
        if (live[0])
            hits_store(A, i, bi[0], bt[0], bi[0] < 0 ? make_float4(0.f, 0.f, 0.f, 0.f) : LBVH ? __ldg(A.sc.sph + bi[0]) : s_sph[bi[0]],
                       ox[0], oy[0], oz[0], dx[0], dy[0], dz[0]);
    }
    if (LBVH) {
        uint32_t v[2] = {cnt.nodes, cnt.leaves};
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            uint32_t x = v[k];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(RT_FULL, x, o);
            if ((threadIdx.x & 31u) == 0 && x) atomicAdd(A.stats + 5 + k, (unsigned long long)x);
        }
    }
}

}  // namespace RT_NS
