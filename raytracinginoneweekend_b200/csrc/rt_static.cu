// rt_static.cu -- rt_closest_hits_static (include/rt_api.h): the closest-hit pass for callers that may not allocate.
// No counterpart in the reference.
//
// rt_closest_hits (rt_api.cu) stages rays, outputs and the sphere records in cudaMalloc'ed memory that belongs to a
// context.  This variant needs no context and allocates nothing: all of its scratch lives in ONE file-scope device
// array of a fixed, modest size (kScratchBytes, ~2.9 MB; every GPU that loads the library holds it until exit):
//
//     [ sphere records {cx, cy, cz, r*r}, padded to 32 with w = -inf | radii | rays n x 6 | index | t | normal n x 3 ]
//
// and the kernel is the SAME one rt_closest_hits runs with RT_FLAG_EXACT (rtx::k_hits<false>: the reference test of
// src/raytracer.hxx:55-72 over all spheres in IEEE arithmetic), pointed at that array -- so index, t and normal are
// bit-identical to the existing entry point.
//
// Rules this file follows:
//  * In HOST code the array's name is a host-side stand-in, not a device address.  The address every launch and copy
//    uses comes from cudaGetSymbolAddress, taken once per device with that device current (scratch_for_current_device).
//  * No cudaMalloc / cudaFree / cudaMallocAsync / cudaHostAlloc and no allocating library call on this path.
//  * All calls on one GPU share the array and it keeps the last call's contents: a call rewrites every byte it later
//    reads (records up to npad, radii, rays, all three outputs) before reading it.  Calls that overlap on one GPU
//    (two streams, two host threads) race on it -- serialise them.
#include <cstdint>
#include <cstdio>
#include <mutex>
#include <string>

#include <cuda_runtime.h>

#include "../../include/rt_api.h"
#include "rt_common.h"

extern "C" int rt_internal_set_error(int code, const char *msg);   // rt_api.cu: sets the calling thread's rt_last_error

namespace {

constexpr size_t kMaxSpheres = RT_STATIC_MAX_SPHERES;
constexpr size_t kMaxRays = RT_STATIC_MAX_RAYS;
// byte offsets inside the scratch array (all multiples of 256)
constexpr size_t kOffSph = 0;                                   // float4[kMaxSpheres]
constexpr size_t kOffRad = kOffSph + kMaxSpheres * 16;          // float[kMaxSpheres]
constexpr size_t kOffRays = kOffRad + kMaxSpheres * 4;          // float[kMaxRays * 6]
constexpr size_t kOffIdx = kOffRays + kMaxRays * 24;            // int32[kMaxRays]
constexpr size_t kOffT = kOffIdx + kMaxRays * 4;                // float[kMaxRays]
constexpr size_t kOffNrm = kOffT + kMaxRays * 4;                // float[kMaxRays * 3]
constexpr size_t kScratchBytes = kOffNrm + kMaxRays * 12;
static_assert(kMaxSpheres % 64 == 0 && kMaxRays % 64 == 0, "offsets stay 256-byte aligned");
static_assert(kScratchBytes <= (4u << 20), "the resident scratch is meant to stay modest");

}  // namespace

static __device__ __align__(256) unsigned char g_static_scratch[kScratchBytes];

namespace {

// The caller's spheres arrive as {cx, cy, cz, r} in the record slots; turn them into the kernel's records in place:
// w = r*r (src/raytracer.hxx:59; one IEEE multiply, as rt_scene_upload does on the host), the radius on the side for the
// normal, and padding that can never be a candidate.  Device code may name the array directly.
__global__ void k_static_prep(uint32_t n, uint32_t npad)
{
    float4 *sph = reinterpret_cast<float4 *>(g_static_scratch + kOffSph);
    float *rad = reinterpret_cast<float *>(g_static_scratch + kOffRad);
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npad) return;
    if (i < n) {
        float4 s = sph[i];
        rad[i] = s.w;
        s.w = __fmul_rn(s.w, s.w);
        sph[i] = s;
    } else {
        sph[i] = make_float4(0.f, 0.f, 0.f, -__int_as_float(0x7f800000));
        rad[i] = 1.f;
    }
}

int sfail(int code, const std::string &msg) { return rt_internal_set_error(code, msg.c_str()); }
int cfail(cudaError_t e, const char *what)
{
    return sfail(RT_ERR_CUDA, std::string("rt_closest_hits_static: ") + what + ": " + cudaGetErrorString(e));
}

constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned char *g_dev_ptr[kMaxDevices];     // device address of g_static_scratch on device d (nullptr: not asked yet)
int g_dev_smem[kMaxDevices];               // opt-in shared memory per block of device d

// Set-up, once per device, with that device current: the DEVICE address of the scratch array.
int scratch_for_current_device(unsigned char **out, int *max_smem)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cfail(e, "cudaGetDevice");
    if (dev < 0 || dev >= kMaxDevices) return sfail(RT_ERR_INVALID, "rt_closest_hits_static: device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_dev_ptr[dev]) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return cfail(e, "cudaGetSymbolAddress(scratch)");
        if (!p) return sfail(RT_ERR_CUDA, "rt_closest_hits_static: cudaGetSymbolAddress returned a null address");
        int smem = 0;
        e = cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        if (e != cudaSuccess) return cfail(e, "cudaDeviceGetAttribute");
        g_dev_smem[dev] = smem;
        g_dev_ptr[dev] = static_cast<unsigned char *>(p);
    }
    *out = g_dev_ptr[dev];
    *max_smem = g_dev_smem[dev];
    return RT_OK;
}

}  // namespace

extern "C" int rt_closest_hits_static(const float *spheres_xyzr, uint32_t n_spheres, const float *rays6, uint32_t n_rays,
                                      int32_t *sphere_index, float *t, float *normal_xyz, void *cuda_stream, uint32_t flags)
{
    if (flags & ~(uint32_t)RT_STATIC_NO_SYNC) return sfail(RT_ERR_INVALID, "rt_closest_hits_static: unknown flag");
    if (!spheres_xyzr || n_spheres == 0) return sfail(RT_ERR_INVALID, "rt_closest_hits_static: empty scene or NULL spheres");
    if (n_spheres > kMaxSpheres)
        return sfail(RT_ERR_INVALID, "rt_closest_hits_static: " + std::to_string(n_spheres) + " spheres exceed RT_STATIC_MAX_SPHERES = " +
                                         std::to_string(kMaxSpheres) + " (the scratch is a fixed device array; use rt_closest_hits)");
    if (n_rays > kMaxRays)
        return sfail(RT_ERR_INVALID, "rt_closest_hits_static: " + std::to_string(n_rays) + " rays exceed RT_STATIC_MAX_RAYS = " +
                                         std::to_string(kMaxRays) + " per call (the scratch is a fixed device array; split the batch or use rt_closest_hits)");
    if (n_rays == 0) return RT_OK;
    if (!rays6 || !sphere_index || !t || !normal_xyz) return sfail(RT_ERR_INVALID, "rt_closest_hits_static: NULL argument");

    unsigned char *base = nullptr;
    int max_smem = 0;
    int r = scratch_for_current_device(&base, &max_smem);
    if (r) return r;
    const uint32_t npad = (n_spheres + 31u) & ~31u;
    const size_t smem = rt_hits_smem_bytes(npad, 0, true, false);
    if (smem > (size_t)max_smem) return sfail(RT_ERR_INVALID, "rt_closest_hits_static: the sphere records do not fit this device's shared memory");

    cudaStream_t st = (cudaStream_t)cuda_stream;
    float4 *d_sph = reinterpret_cast<float4 *>(base + kOffSph);
    float *d_rad = reinterpret_cast<float *>(base + kOffRad);
    float *d_rays = reinterpret_cast<float *>(base + kOffRays);
    int32_t *d_idx = reinterpret_cast<int32_t *>(base + kOffIdx);
    float *d_t = reinterpret_cast<float *>(base + kOffT);
    float *d_n = reinterpret_cast<float *>(base + kOffNrm);

    // inputs: host or device pointers alike (unified addressing)
    cudaError_t e = cudaMemcpyAsync(d_sph, spheres_xyzr, (size_t)n_spheres * 16, cudaMemcpyDefault, st);
    if (e != cudaSuccess) return cfail(e, "copy of the spheres");
    e = cudaMemcpyAsync(d_rays, rays6, (size_t)n_rays * 24, cudaMemcpyDefault, st);
    if (e != cudaSuccess) return cfail(e, "copy of the rays");
    k_static_prep<<<(npad + 255u) / 256u, 256, 0, st>>>(n_spheres, npad);
    e = cudaGetLastError();
    if (e != cudaSuccess) return cfail(e, "k_static_prep");

    RtHitArgs a{};
    a.sc.sph = d_sph; a.sc.sph_rad = d_rad;
    a.sc.n = n_spheres; a.sc.npad = npad;
    a.rays = d_rays; a.nrays = n_rays;
    a.idx = d_idx; a.t = d_t; a.normal = d_n;
    a.stats = nullptr;                       // only the LBVH instantiation counts
    e = rtx::launch_hits(a, st, false);
    if (e != cudaSuccess) return cfail(e, "launch_hits");

    e = cudaMemcpyAsync(sphere_index, d_idx, (size_t)n_rays * 4, cudaMemcpyDefault, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(t, d_t, (size_t)n_rays * 4, cudaMemcpyDefault, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(normal_xyz, d_n, (size_t)n_rays * 12, cudaMemcpyDefault, st);
    if (e != cudaSuccess) return cfail(e, "copy of the results");
    if (!(flags & RT_STATIC_NO_SYNC)) {
        e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) return cfail(e, "cudaStreamSynchronize");
    }
    return RT_OK;
}
