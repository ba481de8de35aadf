// rt_common.cu -- mode-independent kernels: sample-count update, resolve (mean -> gamma -> 8 bit),
// accumulator clear, and the FP32 FMA-chain probe that measures the roofline denominator.
#include "rt_common.h"

// shared memory of the brute-force kernels: mbarrier + sphere records (+ filter records) + candidate masks
size_t rt_render_smem_bytes(uint32_t npad, uint32_t nfilt, bool exact, bool lbvh)
{
    if (lbvh) return (size_t)RT_BVH_SMEM_DEPTH * RT_THREADS * 4u;      // traversal stacks of k_render_bvh: [RT_BVH_SMEM_DEPTH][RT_THREADS]
    if (exact) return 128 + (size_t)npad * 16u + (size_t)RT_CHUNK_WORDS * RT_RPT * RT_THREADS * 4u;
    return 128 + ((size_t)npad + nfilt) * 16u + (size_t)RT_FCHUNK_WORDS * RT_RPT * RT_THREADS * 4u;
}
// tensor-filter kernel: barriers + B matrix (8 x 16 B per column) + sphere records + 32-sphere mask words (one per ray thread)
// (+ radii, material indices and materials when they still fit in the 227 KB of one CTA)
size_t rt_render_tc_smem_bytes(uint32_t npad, uint32_t tc_ns, uint32_t m, bool *mats_in_smem)
{
    size_t b = 512 + (size_t)tc_ns * 8u * 16u + (size_t)npad * 16u + (size_t)(tc_ns / 32u) * (RT_TC_WGS * 128u) * 4u;
#if RT_TC_REGROUP
    b += (size_t)RT_TC_WGS * 20u * 4u + (size_t)15 * (RT_TC_WGS * 128u) * 4u;   // regrouping prototype: key counts + exchange buffers
#endif
    const size_t extra = (size_t)m * 16u + (size_t)npad * 8u;
    const bool fits = b + extra <= RT_TC_SMEM_LIMIT;
    if (mats_in_smem) *mats_in_smem = fits;
    return fits ? b + extra : b;
}
size_t rt_hits_smem_bytes(uint32_t npad, uint32_t nfilt, bool exact, bool lbvh)
{
    if (lbvh) return 128;
    if (exact) return 128 + (size_t)npad * 16u + (size_t)RT_CHUNK_WORDS * 1 * RT_THREADS * 4u;
    return 128 + ((size_t)npad + nfilt) * 16u + (size_t)RT_FCHUNK_WORDS * 1 * RT_THREADS * 4u;
}

// count += spp for every pixel of the listed tiles (the 4th accumulator channel).
__global__ void k_add_count(unsigned long long *accum, uint32_t W, uint4 tile0, const uint4 *tiles, uint32_t nt,
                            uint32_t TP, unsigned long long spp)
{
    unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (unsigned long long)nt * TP) return;
    uint32_t tl = (uint32_t)(i / TP), r = (uint32_t)(i - (unsigned long long)tl * TP);
    uint4 T = tiles ? tiles[tl] : tile0;
    if (r >= T.z * T.w) return;
    uint32_t ty = r / T.z, tx = r - ty * T.z;
    accum[4ull * ((unsigned long long)(T.y + ty) * W + T.x + tx) + 3] += spp;
}

// mean = sum / (count * 2^32) in double, rounded to float (src/main.cxx:205-207: sum / spp);
// gamma = pow(mean, 1/2.2f) (src/main.cxx:39-45) -- evaluated in double and rounded to float so that it
// agrees with glibc's (correctly rounded in practice) powf; 8 bit = (uint8)(255.f * c), truncation
// (src/main.cxx:77-85).  One thread per pixel quad: 4 x 32 B in, 12 B (3 x u32) out, coalesced.
__device__ __forceinline__ void resolve_pixel(const unsigned long long *px, float *mean3, unsigned char *u8)
{
    ulonglong2 a = *reinterpret_cast<const ulonglong2 *>(px);
    ulonglong2 b = *reinterpret_cast<const ulonglong2 *>(px + 2);
    unsigned long long v[3] = {a.x, a.y, b.x};
    double denom = (double)b.y * 4294967296.0;
    const double g = (double)(1.f / 2.2f);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float mean = b.y ? (float)((double)v[k] / denom) : 0.f;
        mean3[k] = mean;
        float c = (float)pow((double)mean, g);
        u8[k] = (unsigned char)(unsigned int)(int)(255.f * c);
    }
}

__global__ void k_resolve(const unsigned long long *__restrict__ accum, unsigned long long npix,
                          unsigned char *__restrict__ rgb8, float *__restrict__ meanf)
{
    unsigned long long q = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;   // pixel quad
    unsigned long long i0 = q * 4ull;
    if (i0 >= npix) return;
    __align__(16) unsigned char u8[12];
    float mean[12];
    int cnt = (int)min(4ull, npix - i0);
    for (int k = 0; k < 4; ++k) {
        if (k < cnt) resolve_pixel(accum + 4ull * (i0 + k), mean + 3 * k, u8 + 3 * k);
    }
    if (cnt == 4) {
        if (rgb8) {
            uint32_t *dst = reinterpret_cast<uint32_t *>(rgb8 + 3ull * i0);   // 12-byte aligned: i0 % 4 == 0
            const uint32_t *src = reinterpret_cast<const uint32_t *>(u8);
            dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2];
        }
        if (meanf) {
            float4 *dst = reinterpret_cast<float4 *>(meanf + 3ull * i0);       // 48-byte aligned
            dst[0] = make_float4(mean[0], mean[1], mean[2], mean[3]);
            dst[1] = make_float4(mean[4], mean[5], mean[6], mean[7]);
            dst[2] = make_float4(mean[8], mean[9], mean[10], mean[11]);
        }
    } else {
        for (int k = 0; k < 3 * cnt; ++k) {
            if (rgb8) rgb8[3ull * i0 + k] = u8[k];
            if (meanf) meanf[3ull * i0 + k] = mean[k];
        }
    }
}

// ---------------------------------------------------------------- device-side tile queue (multi-GPU), see rt_types.h
// cursor of range r = first tile of the range, r * ntiles / nranks
__global__ void k_queue_reset(uint32_t *cursors, uint32_t ntiles, uint32_t nranks)
{
    const uint32_t r = threadIdx.x;
    if (r < nranks) cursors[r * RT_QUEUE_STRIDE] = (r * ntiles) / nranks;      // ntiles <= 2^16, nranks <= 64
}
// count += spp on every pixel of the tiles this GPU acquired (the feed's slot table, read after the render kernel)
__global__ void k_add_count_feed(unsigned long long *accum, uint32_t W, const uint4 *tiles, const uint32_t *slots,
                                 uint32_t TP, unsigned long long spp)
{
    const uint32_t t = slots[blockIdx.y];
    if (t >= RT_FEED_NONE) return;
    const uint4 T = tiles[t];
    const uint32_t npx = T.z * T.w;
    for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < npx; r += gridDim.x * blockDim.x) {
        const uint32_t ty = r / T.z, tx = r - ty * T.z;
        accum[4ull * ((unsigned long long)(T.y + ty) * W + T.x + tx) + 3] += spp;
    }
}
// Resolve (mean -> gamma -> 8 bit) the tiles this GPU rendered and store the texels into `dst`: this GPU's own RGB8 buffer,
// or RANK 0's frame mapped over NVLink (peer access / CUDA IPC) -- the gather of the multi-GPU frame is this store.
// One thread per 4 consecutive pixels of a tile row when the tile is 4-aligned (12-byte stores, 384 contiguous bytes per
// warp), single pixels otherwise.
__global__ void k_resolve_tiles(const unsigned long long *__restrict__ accum, uint32_t W, const uint4 *__restrict__ tiles,
                                const uint32_t *__restrict__ slots, unsigned char *__restrict__ dst)
{
    const uint32_t t = slots[blockIdx.y];
    if (t >= RT_FEED_NONE) return;
    const uint4 T = tiles[t];
    float mean[3];
    if (((T.x | T.z | W) & 3u) == 0u) {
        const uint32_t qw = T.z >> 2, nq = qw * T.w;
        for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < nq; q += gridDim.x * blockDim.x) {
            const uint32_t ty = q / qw, tx = (q - ty * qw) << 2;
            const unsigned long long p = (unsigned long long)(T.y + ty) * W + T.x + tx;
            __align__(16) unsigned char u8[12];
#pragma unroll
            for (int k = 0; k < 4; ++k) resolve_pixel(accum + 4ull * (p + k), mean, u8 + 3 * k);
            uint32_t *d = reinterpret_cast<uint32_t *>(dst + 3ull * p);
            const uint32_t *s = reinterpret_cast<const uint32_t *>(u8);
            d[0] = s[0]; d[1] = s[1]; d[2] = s[2];
        }
    } else {
        const uint32_t npx = T.z * T.w;
        for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < npx; r += gridDim.x * blockDim.x) {
            const uint32_t ty = r / T.z, tx = r - ty * T.z;
            const unsigned long long p = (unsigned long long)(T.y + ty) * W + T.x + tx;
            unsigned char u8[3];
            resolve_pixel(accum + 4ull * p, mean, u8);
            dst[3ull * p] = u8[0]; dst[3ull * p + 1] = u8[1]; dst[3ull * p + 2] = u8[2];
        }
    }
}
cudaError_t rt_launch_queue_reset(uint32_t *cursors, uint32_t ntiles, uint32_t nranks, cudaStream_t s)
{
    k_queue_reset<<<1, 64, 0, s>>>(cursors, ntiles, nranks);
    return cudaGetLastError();
}
cudaError_t rt_launch_add_count_feed(unsigned long long *accum, uint32_t W, const uint4 *tiles, const uint32_t *slots,
                                     uint32_t nslots, uint32_t TP, unsigned long long spp, cudaStream_t s)
{
    if (!nslots) return cudaSuccess;
    k_add_count_feed<<<dim3((TP + 1023) / 1024, nslots), 256, 0, s>>>(accum, W, tiles, slots, TP, spp);
    return cudaGetLastError();
}
cudaError_t rt_launch_resolve_tiles(const unsigned long long *accum, uint32_t W, const uint4 *tiles, const uint32_t *slots,
                                    uint32_t nslots, uint32_t TP, unsigned char *dst, cudaStream_t s)
{
    if (!nslots) return cudaSuccess;
    k_resolve_tiles<<<dim3((TP / 4 + 255) / 256, nslots), 256, 0, s>>>(accum, W, tiles, slots, dst);
    return cudaGetLastError();
}

// FP32 peak probe: 8 independent FMA chains per thread, 4096 x 8 FMAs each, full occupancy.
__global__ void __launch_bounds__(256) k_fp32_probe(float *out, float a, float b, int iters)
{
    float x0 = threadIdx.x * 1e-3f, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f;
    float x4 = x0 + 4.f, x5 = x0 + 5.f, x6 = x0 + 6.f, x7 = x0 + 7.f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
            x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
        }
    }
    float s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (s == 123.456f) out[0] = s;   // keeps the chains alive without a store on the timed path
}

// TF32 tensor peak probe: one CTA per SM, one thread streams tcgen05.mma kind::tf32 (M128 N256 K8, A from TMEM, B from
// shared memory) into one accumulator tile; measures the denominator of the tensor roofline in the same run.
__device__ __forceinline__ uint32_t probe_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void __launch_bounds__(128, 1) k_tf32_probe(int iters)
{
    extern __shared__ __align__(1024) unsigned char psm[];
    float *s = reinterpret_cast<float *>(psm);                                 // B: 256 rows x K=32 tf32 = 32 KB of zeros
    uint64_t *bar = reinterpret_cast<uint64_t *>(psm + 32768);
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(psm + 32768 + 8);
    for (int i = threadIdx.x; i < 8192; i += 128) s[i] = 0.f;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(probe_smem_u32(bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(probe_smem_u32(s_tmem)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = *s_tmem;
    if (threadIdx.x == 0) {
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((256u >> 3) << 17) | ((128u >> 4) << 24);
        const uint32_t sa = probe_smem_u32(s);
        const uint64_t bdesc = (uint64_t)((sa >> 4) & 0x3fffu) | ((uint64_t)((256u * 16u) >> 4) << 16) | ((uint64_t)(128u >> 4) << 32) | (1ull << 46);
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (uint32_t ks = 0; ks < 4; ++ks) {
                uint64_t bd = bdesc + (uint64_t)((ks * 2u * 256u * 16u) >> 4);
                uint32_t acc = ks > 0 ? 1u : 0u;
                asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n}"
                             ::"r"(tm + 64u), "r"(tm + ks * 8u), "l"(bd), "r"(idesc), "r"(acc) : "memory");
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(probe_smem_u32(bar)) : "memory");
        asm volatile("{\n.reg .pred P1;\nPRB_W:\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], 0;\n@P1 bra PRB_D;\nbra PRB_W;\nPRB_D:\n}"
                     ::"r"(probe_smem_u32(bar)) : "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm));
}

cudaError_t rt_launch_tf32_probe(int blocks, int iters, cudaStream_t s)
{
    const size_t smem = 32768 + 64;
    k_tf32_probe<<<blocks, 128, smem, s>>>(iters);
    return cudaGetLastError();
}

cudaError_t rt_launch_add_count(unsigned long long *accum, uint32_t W, uint4 tile0, const uint4 *tiles, uint32_t nt,
                                uint32_t TP, unsigned long long spp, cudaStream_t s)
{
    unsigned long long n = (unsigned long long)nt * TP;
    if (!n) return cudaSuccess;
    k_add_count<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(accum, W, tile0, tiles, nt, TP, spp);
    return cudaGetLastError();
}

cudaError_t rt_launch_resolve(const unsigned long long *accum, unsigned long long npix, unsigned char *rgb8, float *meanf,
                              cudaStream_t s)
{
    unsigned long long quads = (npix + 3) / 4;
    if (!quads) return cudaSuccess;
    k_resolve<<<(unsigned)((quads + 255) / 256), 256, 0, s>>>(accum, npix, rgb8, meanf);
    return cudaGetLastError();
}

cudaError_t rt_launch_fp32_probe(float *scratch, int blocks, int iters, cudaStream_t s)
{
    k_fp32_probe<<<blocks, 256, 0, s>>>(scratch, 0.999f, 0.001f, iters);
    return cudaGetLastError();
}
