// tools/ubench_i2ip.cu -- latency of a dependent I2IP.U2 chain (cvt.pack.sat.u2.s32.b32, the sign-mask instruction of
// rt_tensor.cuh) and throughput with 1/2/4/8 independent chains, for 1..5 warps per SM sub-partition.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/ubench_i2ip.cu -o tools/ubench_i2ip
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int CH>
__global__ void k(uint32_t *out, long long *cyc, int iters, uint32_t seed)
{
    uint32_t a = threadIdx.x * 2654435761u + seed, b = a ^ 0x9e3779b9u;
    uint32_t c[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) c[i] = i;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int i = 0; i < CH; ++i)
                asm volatile("cvt.pack.sat.u2.s32.b32 %0, %1, %2, %0;" : "+r"(c[i]) : "r"(a), "r"(b));
    }
    long long t1 = clock64();
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) s += c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int CH>
void run(int warps_per_sched, uint32_t *out, long long *cyc)
{
    const int iters = 2000;
    k<CH><<<1, 128 * warps_per_sched>>>(out, cyc, iters, 1u);
    cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("chains %d warps/sched %d: %.2f cycles per I2IP per warp (%.2f per scheduler issue)\n", CH, warps_per_sched,
           (double)h / (iters * 8.0 * CH), (double)h / (iters * 8.0 * CH * warps_per_sched));
}
int main()
{
    uint32_t *out; long long *cyc;
    cudaMalloc(&out, 1024 * 4); cudaMalloc(&cyc, 8);
    for (int w = 1; w <= 5; w += 2) { run<1>(w, out, cyc); run<2>(w, out, cyc); run<4>(w, out, cyc); run<8>(w, out, cyc); }
    return 0;
}
