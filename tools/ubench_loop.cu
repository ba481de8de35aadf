// tools/ubench_loop.cu -- candidate-filter inner-loop variants for sm_100a (not product code):
// cycles per ray-sphere for the expanded 9-FMA-pipe-op filter with sphere data from constant memory
// (uniform registers) vs shared memory (LDS.128), R rays per thread.
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
__constant__ float4 c_sph[4096];

template <int R, bool SMEM>
__global__ void __launch_bounds__(256) k_loop(const float *__restrict__ rays, const float4 *__restrict__ gsph, uint32_t *out, int n, int reps)
{
    extern __shared__ float4 s_sph[];
    if (SMEM) { for (int i = threadIdx.x; i < n; i += blockDim.x) s_sph[i] = gsph[i]; __syncthreads(); }
    float nux[R], nuy[R], nuz[R], ou[R], px[R], py[R], pz[R], noo[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const float *q = rays + ((threadIdx.x * R + r) & 1023) * 8;
        nux[r] = q[0]; nuy[r] = q[1]; nuz[r] = q[2]; ou[r] = q[3]; px[r] = q[4]; py[r] = q[5]; pz[r] = q[6]; noo[r] = q[7];
    }
    uint32_t res = 0;
    for (int rep = 0; rep < reps; ++rep) {
        for (int g = 0; g < n; g += 24) {
            float acc[R];
#pragma unroll
            for (int r = 0; r < R; ++r) acc[r] = 0.f;
#pragma unroll
            for (int k = 0; k < 24; ++k) {
                float4 s = SMEM ? s_sph[g + k] : c_sph[g + k];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    float b = fmaf(nuz[r], s.z, fmaf(nuy[r], s.y, fmaf(nux[r], s.x, ou[r])));
                    float ncc = fmaf(pz[r], s.z, fmaf(py[r], s.y, fmaf(px[r], s.x, s.w + noo[r])));
                    float t = __saturatef(fmaf(b, b, ncc));
                    acc[r] = fmaf(t, (float)(1u << k), acc[r]);
                }
            }
#pragma unroll
            for (int r = 0; r < R; ++r) { res ^= (uint32_t)acc[r] * (g + 1 + r); ou[r] += 1e-7f; }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = res;
}

template <int R, bool SMEM>
static void run(int ctas_per_sm)
{
    const int n = 504, reps = 200;
    float *rays; float4 *gs; uint32_t *out;
    cudaMalloc(&rays, 1024 * 8 * 4); cudaMalloc(&gs, 4096 * 16); cudaMalloc(&out, 148 * 8 * 256 * 4);
    float hr[1024 * 8]; for (int i = 0; i < 1024 * 8; ++i) hr[i] = (float)((i * 2654435761u) % 1000) * 1e-3f - .5f;
    float4 hs[4096]; for (int i = 0; i < 4096; ++i) hs[i] = make_float4((i % 23) - 11.f, .2f, (i / 23) - 11.f, -5.f);
    cudaMemcpy(rays, hr, sizeof hr, cudaMemcpyHostToDevice);
    cudaMemcpy(gs, hs, sizeof hs, cudaMemcpyHostToDevice);
    cudaMemcpyToSymbol(c_sph, hs, sizeof hs);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int blocks = sms * ctas_per_sm;
    size_t smem = SMEM ? n * 16 : 0;
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    k_loop<R, SMEM><<<blocks, 256, smem>>>(rays, gs, out, n, 2);
    float best = 1e30f;
    for (int i = 0; i < 3; ++i) {
        cudaEventRecord(a);
        k_loop<R, SMEM><<<blocks, 256, smem>>>(rays, gs, out, n, reps);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    // per SMSP: warps = ctas*8/4; ray-spheres per warp = R * n * reps
    double warps_per_smsp = ctas_per_sm * 2.0;
    double cyc = best * 1e-3 * 1.965e9;
    printf("R=%d %s ctas/SM=%d: %7.3f ms  %6.2f cycles per ray-sphere per SMSP  (%s)\n", R, SMEM ? "smem " : "const", ctas_per_sm, best,
           cyc / (warps_per_smsp * R * n * reps), cudaGetErrorString(cudaGetLastError()));
    cudaFree(rays); cudaFree(gs); cudaFree(out);
}


// (d) fully unrolled over NG groups: sphere fields become constant-bank operands with immediate offsets
template <int R, int NG>
__global__ void __launch_bounds__(256) k_unrolled(const float *__restrict__ rays, uint32_t *out, int reps)
{
    float nux[R], nuy[R], nuz[R], ou[R], px[R], py[R], pz[R], noo[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const float *q = rays + ((threadIdx.x * R + r) & 1023) * 8;
        nux[r] = q[0]; nuy[r] = q[1]; nuz[r] = q[2]; ou[r] = q[3]; px[r] = q[4]; py[r] = q[5]; pz[r] = q[6]; noo[r] = q[7];
    }
    uint32_t res = 0;
    for (int rep = 0; rep < reps; ++rep) {
#pragma unroll
        for (int g = 0; g < NG; ++g) {
            float acc[R];
#pragma unroll
            for (int r = 0; r < R; ++r) acc[r] = 0.f;
#pragma unroll
            for (int k = 0; k < 24; ++k) {
                float4 s = c_sph[g * 24 + k];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    float b = fmaf(nuz[r], s.z, fmaf(nuy[r], s.y, fmaf(nux[r], s.x, ou[r])));
                    float ncc = fmaf(pz[r], s.z, fmaf(py[r], s.y, fmaf(px[r], s.x, s.w + noo[r])));
                    float t = __saturatef(fmaf(b, b, ncc));
                    acc[r] = fmaf(acc[r], 2.f, t);
                }
            }
#pragma unroll
            for (int r = 0; r < R; ++r) { res ^= (uint32_t)acc[r] * (g + 1 + r); ou[r] += 1e-7f; }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = res;
}

template <int R, int NG>
static void run_unrolled(int ctas_per_sm)
{
    const int reps = 200 * 21 / NG;
    float *rays; uint32_t *out;
    cudaMalloc(&rays, 1024 * 8 * 4); cudaMalloc(&out, 148 * 8 * 256 * 4);
    float hr[1024 * 8]; for (int i = 0; i < 1024 * 8; ++i) hr[i] = (float)((i * 2654435761u) % 1000) * 1e-3f - .5f;
    float4 hs[4096]; for (int i = 0; i < 4096; ++i) hs[i] = make_float4((i % 23) - 11.f, .2f, (i / 23) - 11.f, -5.f);
    cudaMemcpy(rays, hr, sizeof hr, cudaMemcpyHostToDevice);
    cudaMemcpyToSymbol(c_sph, hs, sizeof hs);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int blocks = sms * ctas_per_sm;
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    k_unrolled<R, NG><<<blocks, 256>>>(rays, out, 2);
    float best = 1e30f;
    for (int i = 0; i < 3; ++i) {
        cudaEventRecord(a);
        k_unrolled<R, NG><<<blocks, 256>>>(rays, out, reps);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    double warps_per_smsp = ctas_per_sm * 2.0;
    double cyc = best * 1e-3 * 1.965e9;
    printf("R=%d unrolled NG=%2d (%3d spheres, %4d KB code est) ctas/SM=%d: %7.3f ms  %6.2f cycles per ray-sphere per SMSP  (%s)\n", R, NG, NG * 24,
           NG * 24 * 9 * R * 16 / 1024, ctas_per_sm, best, cyc / (warps_per_smsp * R * NG * 24.0 * reps), cudaGetErrorString(cudaGetLastError()));
    cudaFree(rays); cudaFree(out);
}

int main()
{
    run_unrolled<2, 1>(3); run_unrolled<2, 4>(3); run_unrolled<2, 21>(3); run_unrolled<1, 21>(6); run_unrolled<3, 21>(2);

    run<2, false>(2); run<2, false>(3); run<2, true>(2); run<2, true>(3);
    run<4, false>(1); run<4, false>(2); run<4, true>(1); run<4, true>(2);
    run<1, false>(4); run<1, true>(4); run<3, false>(2); run<3, true>(2);
    return 0;
}
