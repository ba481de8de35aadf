// tools/ubench_signs.cu -- cycles per 32-value sign-mask word for the packing variants of rt_tensor.cuh
// (I2IP saturating packs, funnel-shift chains, a mix of both) at 4/5 warps per SM sub-partition.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/ubench_signs.cu -o tools/ubench_signs
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t word_i2ip(const uint32_t (&v)[32])
{
    uint32_t m = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        uint32_t t, w;
        asm volatile("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(t) : "r"(v[4 * j + 0]), "r"(v[4 * j + 1]), "r"(0u));
        asm volatile("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(w) : "r"(v[4 * j + 2]), "r"(v[4 * j + 3]), "r"(t));
        m |= w & (0x01010101u << j);
    }
    return m;
}
__device__ __forceinline__ uint32_t word_shf(const uint32_t (&v)[32])
{
    uint32_t c[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        c[k] = 0u;
#pragma unroll
        for (int i = 0; i < 8; ++i) asm volatile("shf.l.wrap.b32 %0, %1, %2, 1;" : "=r"(c[k]) : "r"(v[8 * k + i]), "r"(c[k]));
    }
    return ~((c[0] << 24) | (c[1] << 16) | (c[2] << 8) | c[3]);
}
__device__ __forceinline__ uint32_t word_prmt(const uint32_t (&v)[32])
{
    // sign bytes by PRMT (3 per 4 values), then 8 x (shift + lop)
    uint32_t m = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        uint32_t a, b, w;
        asm volatile("prmt.b32 %0, %1, %2, 0x0073;" : "=r"(a) : "r"(v[4 * j + 0]), "r"(v[4 * j + 1]));
        asm volatile("prmt.b32 %0, %1, %2, 0x0073;" : "=r"(b) : "r"(v[4 * j + 2]), "r"(v[4 * j + 3]));
        asm volatile("prmt.b32 %0, %1, %2, 0x5410;" : "=r"(w) : "r"(a), "r"(b));
        m |= (w & 0x80808080u) >> j;
    }
    return m;
}

template <int MODE>
__global__ void k(uint32_t *out, int iters, uint32_t seed)
{
    uint32_t v[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = (threadIdx.x * 2654435761u + i * 40503u + seed) ^ (i << 27);
    uint32_t acc = 0;
    long long t0 = clock64();
#define OPAQUE() _Pragma("unroll") for (int i = 0; i < 32; ++i) asm volatile("" : "+r"(v[i]))
    for (int it = 0; it < iters; ++it) {
        OPAQUE();
        if (MODE == 0) { acc += word_i2ip(v); OPAQUE(); acc += word_i2ip(v); }
        if (MODE == 1) { acc += word_shf(v); OPAQUE(); acc += word_shf(v); }
        if (MODE == 2) { acc += word_i2ip(v); OPAQUE(); acc += word_shf(v); }
        if (MODE == 3) { acc += word_prmt(v); OPAQUE(); acc += word_prmt(v); }
        if (MODE == 4) { acc += word_i2ip(v); OPAQUE(); acc += word_prmt(v); }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) out[1] = (uint32_t)((t1 - t0) >> 0);
    if (acc == 0x12345u) out[0] = acc;
}

template <int MODE>
void run(const char *name, int threads, uint32_t *d)
{
    const int iters = 1 << 16;
    k<MODE><<<148, threads>>>(d, 256, 1u);
    cudaDeviceSynchronize();
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    k<MODE><<<148, threads>>>(d, iters, 2u);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0.f; cudaEventElapsedTime(&ms, a, b);
    uint32_t h[2];
    cudaMemcpy(h, d, 8, cudaMemcpyDeviceToHost);
    const double warps_per_smsp = threads / 128.0;
    const double cyc = ms * 1e-3 * 1.965e9;           // SM clock under load on this pool (bench.py clocks line)
    printf("%-24s threads %4d: %7.1f cycles per word per warp, %6.1f per sub-partition   (clock64: %.1f per warp)\n", name, threads,
           cyc / (2.0 * iters), cyc / (2.0 * iters * warps_per_smsp), h[1] / (2.0 * iters));
}

int main()
{
    uint32_t *d; cudaMalloc(&d, 64);
    for (int threads : {128, 512, 640}) {
        run<0>("i2ip", threads, d);
        run<1>("shf chains", threads, d);
        run<2>("i2ip + shf alternating", threads, d);
        run<3>("prmt", threads, d);
        run<4>("i2ip + prmt alternating", threads, d);
    }
    return 0;
}
