// tools/ubench_umma.cu -- tcgen05.mma kind::tf32 instruction throughput for small shapes (not product code):
// cycles per MMA for N = 64/128/256, A from smem (SS) or TMEM (TS), accumulating into one D tile or rotating over D tiles.
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo)
{
    return (uint64_t)((saddr >> 4) & 0x3fffu) | ((uint64_t)((lbo >> 4) & 0x3fffu) << 16) | ((uint64_t)((sbo >> 4) & 0x3fffu) << 32) | (1ull << 46);
}
extern __shared__ __align__(1024) unsigned char smem[];
template <int N, bool TS, int ROT, int KSTEPS>
__global__ void __launch_bounds__(128, 1) k_bench(long long *out, int iters)
{
    float *s = reinterpret_cast<float *>(smem);
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem + 96 * 1024);
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(bar + 1);
    for (int i = threadIdx.x; i < 24 * 1024; i += 128) s[i] = 0.f;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(s_tmem)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tm = *s_tmem;
    if (threadIdx.x == 0) {
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
        const uint64_t adesc = make_desc(smem_u32(s), 128 * 16, 128);                 // 128 rows x K=32: 16 KB at 0
        const uint64_t bdesc = make_desc(smem_u32(s) + 16384, 256 * 16, 128);         // 256 rows x K=32: 32 KB at 16 KB
        long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int ks = 0; ks < KSTEPS; ++ks) {
                uint32_t d = tm + 64 + ((it % ROT) * N) % (512 - 64 - N + 1);
                uint64_t ad = adesc + (uint64_t)(((ks & 3) * 2 * 128 * 16) >> 4);
                uint64_t bd = bdesc + (uint64_t)(((ks & 3) * 2 * 256 * 16) >> 4);
                uint32_t acc = ks > 0 ? 1u : 0u;
                if (TS)
                    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n}"
                                 ::"r"(d), "r"(tm + (ks & 3) * 8), "l"(bd), "r"(idesc), "r"(acc) : "memory");
                else
                    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}"
                                 ::"r"(d), "l"(ad), "l"(bd), "r"(idesc), "r"(acc) : "memory");
            }
        }
        long long t1 = clock64();
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
        asm volatile("{\n.reg .pred P1;\nUW:\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], 0;\n@P1 bra UD;\nbra UW;\nUD:\n}" ::"r"(smem_u32(bar)) : "memory");
        long long t2 = clock64();
        if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm));
}
template <int N, bool TS, int ROT, int KSTEPS>
static void run(const char *name)
{
    long long *d; cudaMalloc(&d, 16);
    size_t sm = 96 * 1024 + 64;
    cudaFuncSetAttribute(k_bench<N, TS, ROT, KSTEPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    int iters = 2000;
    k_bench<N, TS, ROT, KSTEPS><<<148, 128, sm>>>(d, iters);
    k_bench<N, TS, ROT, KSTEPS><<<148, 128, sm>>>(d, iters);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    double n = (double)iters * KSTEPS;
    printf("%-34s N=%3d %s: issue %.1f cyc/MMA, complete %.1f cyc/MMA -> %.0f MAC/cyc/SM  (%s)\n", name, N, TS ? "TS" : "SS", h[0] / n, h[1] / n,
           128.0 * N * 8 * n / h[1], cudaGetErrorString(e));
    for (int it : {1, 2, 4}) {
        k_bench<N, TS, ROT, KSTEPS><<<148, 128, sm>>>(d, it);
        cudaDeviceSynchronize();
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        printf("      %d x %d MMAs + commit + wait: issue %lld cycles, until mbarrier completes %lld cycles\n", it, KSTEPS, h[0], h[1]);
    }
    cudaFree(d);
}
int main()
{
    run<64, false, 1, 4>("one D, 4 k-steps accumulate");
    run<64, true, 1, 4>("one D, 4 k-steps accumulate");
    run<128, false, 1, 4>("one D, 4 k-steps accumulate");
    run<128, true, 1, 4>("one D, 4 k-steps accumulate");
    run<256, false, 1, 4>("one D, 4 k-steps accumulate");
    run<256, true, 1, 4>("one D, 4 k-steps accumulate");
    run<64, true, 4, 4>("4 D tiles rotating");
    run<128, true, 3, 4>("3 D tiles rotating");
    run<64, true, 4, 1>("4 D tiles rotating, 1 k-step");
    run<256, true, 1, 1>("one D, 1 k-step, overwrite");
    run<128, true, 1, 1>("one D, 1 k-step, overwrite");
    run<64, true, 1, 1>("one D, 1 k-step, overwrite");
    return 0;
}
