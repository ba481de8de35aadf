// tools/ubench_mix.cu -- issue-mix microbenchmark for sm_100a: cycles per "round" per SMSP for a round made of
// NF2 packed FFMA2, NF scalar FFMA, NS funnel shifts, NL LOP3 and ND LDS.128 (broadcast).  Not product code.
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

__device__ __forceinline__ unsigned long long ffma2(unsigned long long a, unsigned long long b, unsigned long long c)
{
    unsigned long long d;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ unsigned long long pack(float lo, float hi)
{
    return ((unsigned long long)__float_as_uint(hi) << 32) | __float_as_uint(lo);
}

template <int NF2, int NF, int NS, int NL, int ND>
__global__ void __launch_bounds__(256) k_mix(float *out, long long *cyc, float a, float b, int iters)
{
    __shared__ float4 sm[512];
    sm[threadIdx.x] = make_float4(a, b, a * 1.0001f, b);
    sm[threadIdx.x + 256] = make_float4(a, b * 1.01f, a, b);
    __syncthreads();
    unsigned long long x[8];
    float y[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { x[k] = pack(threadIdx.x * 1e-3f + k, threadIdx.x * 2e-3f + k); y[k] = threadIdx.x * 3e-3f + k; }
    uint32_t m[4] = {0, 0, 0, 0};
    unsigned long long aa = pack(a, a * 1.0001f), bb = pack(b, b * 1.01f);
    float sa = a, sb = b;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 8; ++r) {
#pragma unroll
            for (int d = 0; d < ND; ++d) {
                float4 s = sm[((i * 8 + r) * ND + d) & 511];
                aa = pack(s.x, s.y); bb = pack(s.z, s.w); sa = s.x; sb = s.w;
            }
#pragma unroll
            for (int k = 0; k < NF2; ++k) x[k & 7] = ffma2(x[k & 7], aa, bb);
#pragma unroll
            for (int k = 0; k < NF; ++k) y[k & 7] = fmaf(y[k & 7], sa, sb);
#pragma unroll
            for (int k = 0; k < NL; ++k) m[k & 3] = (uint32_t)x[k & 7] & (m[k & 3] | ~(uint32_t)(x[(k + 1) & 7] >> 32));
#pragma unroll
            for (int k = 0; k < NS; ++k) m[k & 3] = __funnelshift_l((uint32_t)(x[k & 7] >> (32 * (k & 1))) ^ __float_as_uint(y[k & 7]), m[k & 3], 1);
        }
    }
    long long t1 = clock64();
    unsigned long long s = m[0] ^ m[1] ^ m[2] ^ m[3];
#pragma unroll
    for (int k = 0; k < 8; ++k) s ^= x[k] ^ __float_as_uint(y[k]);
    if (s == 0x12345ull) out[0] = 1.f;
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

template <int NF2, int NF, int NS, int NL, int ND>
static void run(int warps_per_smsp)
{
    float *d; long long *c;
    cudaMalloc(&d, 256); cudaMalloc(&c, 8);
    int dev, sms;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int blocks = sms * warps_per_smsp / 2;      // 256 threads = 8 warps = 2 per SMSP
    int iters = 2048;
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    k_mix<NF2, NF, NS, NL, ND><<<blocks, 256>>>(d, c, 0.999f, 0.001f, 16);
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(a);
        k_mix<NF2, NF, NS, NL, ND><<<blocks, 256>>>(d, c, 0.999f, 0.001f, iters);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    long long cyc = 0;
    cudaMemcpy(&cyc, c, 8, cudaMemcpyDeviceToHost);
    double rounds_per_smsp = (double)iters * 8 * warps_per_smsp;
    double cyc_per_round = (double)cyc / rounds_per_smsp;
    int issue = NF2 + NF + NS + NL + ND;
    printf("F2=%2d F=%2d SHF=%d LOP=%d LDS=%d w/smsp=%2d | %7.3f ms | %6.2f cyc/round/SMSP | issue slots %2d | FMA lanes-cyc %2d | err=%s\n",
           NF2, NF, NS, NL, ND, warps_per_smsp, best, cyc_per_round, issue, 2 * NF2 + NF, cudaGetErrorString(cudaGetLastError()));
    cudaFree(d); cudaFree(c);
}

int main()
{
    run<8, 0, 0, 0, 0>(8);
    run<8, 0, 0, 0, 1>(8);
    run<8, 0, 2, 0, 0>(8);
    run<8, 0, 4, 0, 0>(8);
    run<8, 0, 4, 2, 1>(8);
    run<8, 0, 4, 2, 1>(4);
    run<0, 16, 0, 0, 0>(8);
    run<0, 16, 2, 0, 0>(8);
    run<0, 16, 4, 0, 0>(8);
    run<0, 16, 4, 0, 1>(8);
    run<0, 22, 2, 0, 1>(8);
    run<0, 22, 2, 0, 1>(4);
    run<4, 8, 2, 0, 1>(8);
    run<0, 8, 8, 0, 0>(8);
    run<0, 0, 8, 0, 0>(8);
    run<0, 0, 8, 8, 0>(8);
    return 0;
}
