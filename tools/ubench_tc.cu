// tools/ubench_tc.cu -- tensor-path feasibility probes for the candidate filter (not product code):
//   (1) legacy mma.sync m16n8k8 TF32 rate per SM            (results land in registers)
//   (2) tcgen05.ld (TMEM -> registers) bandwidth per SM     (results of tcgen05.mma land in TMEM)
//   (3) 3-input max.f32 rate (sign reduction of discriminants)
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

__global__ void __launch_bounds__(1024) k_mma_sync(float *out, int iters)
{
    uint32_t a[4] = {0x3f800000u + threadIdx.x, 0x3f000000u, 0x3e800000u, 0x3f800000u};
    uint32_t b[2] = {0x3f800000u, 0x3f000000u + threadIdx.x};
    float c[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
            asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                         : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NREG>
__global__ void __launch_bounds__(1024) k_ldtm(uint32_t *out, int iters)
{
    __shared__ uint32_t s_tmem;
    const uint32_t warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_tmem)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t base = s_tmem + (((warp & 3u) * 32u) << 16);
    uint32_t acc = 0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int c = 0; c < 512; c += NREG) {
            uint32_t col = (uint32_t)((c + warp * 32) & 511);
            if (NREG == 32) {
                uint32_t v[32];
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                             "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                             : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                               "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                               "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                               "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                             : "r"(base + col));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int i = 0; i < 32; i += 8) acc ^= v[i];
            } else {
                uint32_t v[8];
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                             : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                             : "r"(base + col));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                acc ^= v[0] ^ v[4];
            }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(s_tmem));
}

// same, but 4 loads in flight before one wait
__global__ void __launch_bounds__(1024) k_ldtm_pipe(uint32_t *out, int iters)
{
    __shared__ uint32_t s_tmem;
    const uint32_t warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_tmem)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t base = s_tmem + (((warp & 3u) * 32u) << 16);
    uint32_t acc = 0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int c = 0; c < 512; c += 64) {
            uint32_t v[4][16];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                uint32_t col = (uint32_t)((c + q * 16 + warp * 32) & 511);
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                             : "=r"(v[q][0]), "=r"(v[q][1]), "=r"(v[q][2]), "=r"(v[q][3]), "=r"(v[q][4]), "=r"(v[q][5]), "=r"(v[q][6]), "=r"(v[q][7]),
                               "=r"(v[q][8]), "=r"(v[q][9]), "=r"(v[q][10]), "=r"(v[q][11]), "=r"(v[q][12]), "=r"(v[q][13]), "=r"(v[q][14]), "=r"(v[q][15])
                             : "r"(base + col));
            }
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int q = 0; q < 4; ++q) acc ^= v[q][0] ^ v[q][8];
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(s_tmem));
}

__global__ void __launch_bounds__(1024) k_max3(float *out, const float *in, int iters)
{
    float a[8], x = in[threadIdx.x & 31], y = in[(threadIdx.x + 1) & 31];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = in[i];
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(x), "f"(y));
        x += 1.f;
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(1024) k_max2(float *out, const float *in, int iters)
{
    float a[8], x = in[threadIdx.x & 31];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = in[i];
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) asm volatile("max.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(x));
        x += 1.f;
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static float time_ms(F f)
{
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    f();
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int i = 0; i < 3; ++i) {
        cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    return best;
}

int main()
{
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const double clk = 1.965e9;
    float *out; cudaMalloc(&out, sms * 1024 * 4 * 2);
    float *in; cudaMalloc(&in, 64 * 4); cudaMemset(in, 0, 64 * 4);
    for (int warps : {4, 8, 16, 32}) {
        int iters = 20000;
        float ms = time_ms([&] { k_mma_sync<<<sms, warps * 32>>>(out, iters); });
        double macs = (double)sms * warps * iters * 8 * (16.0 * 8 * 8);
        double cyc = ms * 1e-3 * clk;
        printf("mma.sync m16n8k8 tf32: %2d warps/SM  %8.3f ms  %7.1f MAC/cyc/SM  %6.2f cyc per MMA per SMSP  (%.1f TFLOP/s)  %s\n", warps, ms,
               macs / cyc / sms, cyc / ((double)warps / 4 * iters * 8), 2 * macs / (ms * 1e-3) / 1e12, cudaGetErrorString(cudaGetLastError()));
    }
    for (int warps : {4, 8, 16, 32}) {
        int iters = 2000;
        float ms = time_ms([&] { k_ldtm<32><<<sms, warps * 32>>>((uint32_t *)out, iters); });
        double bytes = (double)sms * warps * iters * 512.0 * 32 * 4;
        double cyc = ms * 1e-3 * clk;
        printf("tcgen05.ld 32x32b.x32 (wait each): %2d warps/SM  %8.3f ms  %7.1f B/cyc/SM  %s\n", warps, ms, bytes / cyc / sms, cudaGetErrorString(cudaGetLastError()));
        ms = time_ms([&] { k_ldtm<8><<<sms, warps * 32>>>((uint32_t *)out, iters); });
        cyc = ms * 1e-3 * clk;
        printf("tcgen05.ld 32x32b.x8  (wait each): %2d warps/SM  %8.3f ms  %7.1f B/cyc/SM  %s\n", warps, ms, bytes / cyc / sms, cudaGetErrorString(cudaGetLastError()));
        ms = time_ms([&] { k_ldtm_pipe<<<sms, warps * 32>>>((uint32_t *)out, iters); });
        cyc = ms * 1e-3 * clk;
        printf("tcgen05.ld 32x32b.x16 x4 pipelined: %2d warps/SM  %8.3f ms  %7.1f B/cyc/SM  %s\n", warps, ms, bytes / cyc / sms, cudaGetErrorString(cudaGetLastError()));
    }
    for (int warps : {16, 32}) {
        int iters = 20000;
        float ms = time_ms([&] { k_max3<<<sms, warps * 32>>>(out, in, iters); });
        double cyc = ms * 1e-3 * clk;
        printf("max.f32 3-input: %2d warps/SM  %6.2f cyc per warp-inst per SMSP  %s\n", warps, cyc / ((double)warps / 4 * iters * 8), cudaGetErrorString(cudaGetLastError()));
        ms = time_ms([&] { k_max2<<<sms, warps * 32>>>(out, in, iters); });
        cyc = ms * 1e-3 * clk;
        printf("max.f32 2-input: %2d warps/SM  %6.2f cyc per warp-inst per SMSP  %s\n", warps, cyc / ((double)warps / 4 * iters * 8), cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
