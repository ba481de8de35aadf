/* tools/check_mdiv.c -- CPU check of the candidate test's division by a ray's a (rt_device.cuh, div_by_a): two FMA residual
 * corrections on n * RN(1/a) against IEEE division, random pairs with stressed significands.
 * gcc -O2 -mfma -ffp-contract=off tools/check_mdiv.c -o /tmp/check_mdiv -lm && /tmp/check_mdiv 400000000 */
#include <math.h>
#include <stdio.h>
#include <stdint.h>
#include <string.h>
#include <stdlib.h>
static inline float mdiv(float n, float a, float ra)
{
    float q0 = n * ra;
    float r0 = fmaf(-a, q0, n);
    float q1 = fmaf(r0, ra, q0);
    float r1 = fmaf(-a, q1, n);
    return fmaf(r1, ra, q1);
}
static uint64_t s = 88172645463325252ull;
static inline uint64_t xs(void) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; }
static inline float rndf(int emin, int emax)
{
    uint64_t r = xs();
    uint32_t man = (uint32_t)(r & 0x7fffff);
    int e = emin + (int)((r >> 23) % (uint64_t)(emax - emin + 1));
    uint32_t sign = (uint32_t)((r >> 60) & 1u) << 31;
    uint32_t u = sign | ((uint32_t)(e + 127) << 23) | man;
    if ((r >> 61) == 0) u |= 0x7fff00;       /* mantissas near all-ones */
    if ((r >> 61) == 1) u &= ~0x7fff00u;     /* mantissas near 1.0 */
    float f; memcpy(&f, &u, 4); return f;
}
int main(int argc, char **argv)
{
    long N = argc > 1 ? atol(argv[1]) : 100000000L;
    long bad = 0;
    for (long i = 0; i < N; ++i) {
        float a = fabsf(rndf(-40, 40)), n = rndf(-60, 60);
        float ra = 1.0f / a;
        float q = n / a, m = mdiv(n, a, ra);
        uint32_t uq, um; memcpy(&uq, &q, 4); memcpy(&um, &m, 4);
        if (uq != um && !(isnan(q) && isnan(m))) {
            /* ignore results that are subnormal / overflow (t < kMIN or not finite: same decision) */
            if (fabsf(q) > 1e-30f && fabsf(q) < 1e30f) { if (bad < 10) printf("n=%a a=%a q=%a m=%a\n", n, a, q, m); ++bad; }
        }
    }
    printf("tested %ld, mismatches in the normal range: %ld\n", N, bad);
    return 0;
}
