// rt_api.cu -- the C ABI of include/rt_api.h: context, scene/camera upload, render scheduling,
// readback, statistics, PPM writer, FP32 probe and the node-level tile queue.
// Host code only (compiled with -Xcompiler -ffp-contract=off so the camera maths keeps the reference's
// float rounding); all device work is in rt_kernels_{exact,fast}.cu, rt_common.cu and rt_lbvh.cu.
#include <algorithm>
#include <atomic>
#include <cerrno>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <new>
#include <string>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "../../include/rt_api.h"
#include "rt_common.h"
#include "rt_lbvh.h"

namespace {

thread_local std::string g_err;

int fail(int code, const std::string &msg)
{
    g_err = msg;
    return code;
}
int cuda_fail(cudaError_t e, const char *what)
{
    return fail(RT_ERR_CUDA, std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")");
}
#define CK(call)                                              \
    do {                                                      \
        cudaError_t e_ = (call);                              \
        if (e_ != cudaSuccess) return cuda_fail(e_, #call);   \
    } while (0)

struct EventPair { cudaEvent_t a, b; };

}  // namespace

constexpr int kAux = 2;    // auxiliary streams for RT_FLAG_OVERLAP chunks (as many chunks can run beside each other)

struct rt_ctx {
    int device = 0;
    int sm_count = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr;

    // scene
    float4 *d_sph = nullptr; float *d_rad = nullptr; uint32_t *d_meta = nullptr; float4 *d_mat = nullptr;
    float4 *d_filt = nullptr;               // conservative-filter records of the throughput build
    float4 *d_tcb = nullptr;                // tensor-filter sphere matrix (rt_tensor.cuh), nullptr above RT_TC_MAX_SPHERES
    uint32_t tc_ns = 0;
    float tc_g[3] = {0.f, 0.f, 0.f};
    uint32_t n = 0, npad = 0, m = 0, nfilt = 0;
    float sigma = 1.f, S = 1.f, o_limit2 = 0.f, scene_radius = 1.f;
    rt_lbvh *bvh = nullptr;
    bool bvh_valid = false;

    RtDeviceCamera cam{};
    bool have_cam = false;

    uint32_t W = 0, H = 0;
    unsigned long long *d_accum = nullptr;
    unsigned char *d_rgb8 = nullptr;
    float *d_meanf = nullptr;

    uint32_t *d_counter = nullptr;          // ring of work counters, one per launch
    uint32_t counter_slot = 0;
    uint4 *d_tile_ring = nullptr;           // ring of tile lists, one slice per launch (no stream-ordered allocation per launch: a block
    uint32_t tile_ring_pos = 0;             // freed on one stream and reused on another would chain the overlapped launches together)
    unsigned long long *d_stats = nullptr;  // 8 x u64
    float *d_scratch = nullptr;

    void *h_pin = nullptr; size_t h_pin_bytes = 0;   // pinned staging for uploads / readbacks
    // device-side tile queue (rt_render_queue): the feed of the most recent launch and what its resolve needs
    RtTileFeed *d_feed = nullptr; uint32_t feed_slots = 0;
    const uint4 *feed_tiles = nullptr; uint32_t feed_TP = 0; bool feed_valid = false;
    uint32_t *d_self_cursors = nullptr;     // rt_render's own one-range queue (large rectangles are rendered tile by tile)
    void *d_hits = nullptr; size_t d_hits_bytes = 0; // grow-only scratch of rt_primary_hits / rt_closest_hits (index, t, normal, rays)

    // RT_FLAG_OVERLAP: chunk launches rotate over kAux auxiliary streams so that the tail of one persistent launch is
    // filled by the next; aux_tail[i] marks the end of the work queued on aux[i], aux_dirty[i] = c->stream has not waited for it
    cudaStream_t aux[kAux] = {};
    cudaEvent_t aux_tail[kAux] = {}, aux_head = nullptr;
    bool aux_dirty[kAux] = {};
    int aux_next = 0;

    std::vector<EventPair> free_events, pending;
    rt_stats hs{};                          // host-side part of the statistics
    int occ[2][2] = {{0, 0}, {0, 0}};       // cached CTAs/SM [exact][lbvh]
    int occ_small = 0;                      // cached CTAs/SM of k_render_small
    size_t occ_smem[2][2] = {{0, 0}, {0, 0}};
};

namespace {

const uint32_t kCounterRing = 4096;
const uint32_t kTileRing = 1u << 16;       // tiles (1 MiB)

// Every entry point that queues work on c->stream (or waits for it) goes through here: selects the device and makes
// c->stream wait for the overlapped chunk launches still running on the auxiliary streams (RT_FLAG_OVERLAP), so that
// clear / resolve / readback / upload are ordered after them without the caller doing anything.
cudaError_t enter_ctx(rt_ctx *c)
{
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) return e;
    for (int i = 0; i < kAux; ++i)
        if (c->aux_dirty[i]) {
            e = cudaStreamWaitEvent(c->stream, c->aux_tail[i], 0);
            if (e != cudaSuccess) return e;
            c->aux_dirty[i] = false;
        }
    return cudaSuccess;
}
// stream for the next overlapped launch: ordered after everything queued on c->stream so far, not after the other aux stream
cudaError_t overlap_stream(rt_ctx *c, cudaStream_t *out, int *slot)
{
    cudaError_t e;
    if (!c->aux[0]) {
        for (int i = 0; i < kAux; ++i) {
            if ((e = cudaStreamCreateWithFlags(&c->aux[i], cudaStreamNonBlocking)) != cudaSuccess) return e;
            if ((e = cudaEventCreateWithFlags(&c->aux_tail[i], cudaEventDisableTiming)) != cudaSuccess) return e;
        }
        if ((e = cudaEventCreateWithFlags(&c->aux_head, cudaEventDisableTiming)) != cudaSuccess) return e;
    }
    const int i = c->aux_next;
    c->aux_next = (i + 1) % kAux;
    if ((e = cudaEventRecord(c->aux_head, c->stream)) != cudaSuccess) return e;
    if ((e = cudaStreamWaitEvent(c->aux[i], c->aux_head, 0)) != cudaSuccess) return e;
    *out = c->aux[i];
    *slot = i;
    return cudaSuccess;
}

int ensure_pinned(rt_ctx *c, size_t bytes)
{
    if (bytes <= c->h_pin_bytes) return RT_OK;
    if (c->h_pin) { cudaFreeHost(c->h_pin); c->h_pin = nullptr; c->h_pin_bytes = 0; }
    size_t want = bytes + bytes / 4 + 4096;
    CK(cudaMallocHost(&c->h_pin, want));
    c->h_pin_bytes = want;
    return RT_OK;
}

int drain_events(rt_ctx *c)
{
    for (auto &p : c->pending) {
        CK(cudaEventSynchronize(p.b));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, p.a, p.b));
        c->hs.render_ms += ms;
        c->hs.last_render_ms = ms;
        c->free_events.push_back(p);
    }
    c->pending.clear();
    return RT_OK;
}

int get_events(rt_ctx *c, EventPair *out)
{
    if (c->pending.size() > 256) { int r = drain_events(c); if (r) return r; }
    if (!c->free_events.empty()) { *out = c->free_events.back(); c->free_events.pop_back(); return RT_OK; }
    CK(cudaEventCreate(&out->a));
    CK(cudaEventCreate(&out->b));
    return RT_OK;
}

void free_scene(rt_ctx *c)
{
    cudaFree(c->d_sph); cudaFree(c->d_rad); cudaFree(c->d_meta); cudaFree(c->d_mat); cudaFree(c->d_filt); cudaFree(c->d_tcb);
    c->d_tcb = nullptr; c->tc_ns = 0;
    c->d_sph = nullptr; c->d_rad = nullptr; c->d_meta = nullptr; c->d_mat = nullptr; c->d_filt = nullptr;
    c->n = c->npad = c->m = c->nfilt = 0;
    c->bvh_valid = false;
}

void free_framebuffer(rt_ctx *c)
{
    cudaFree(c->d_accum); cudaFree(c->d_rgb8); cudaFree(c->d_meanf);
    c->d_accum = nullptr; c->d_rgb8 = nullptr; c->d_meanf = nullptr;
    c->W = c->H = 0;
}

RtDeviceScene device_scene(const rt_ctx *c, bool lbvh)
{
    RtDeviceScene s{};
    s.sph = c->d_sph; s.sph_rad = c->d_rad; s.sph_meta = c->d_meta; s.mat = c->d_mat;
    s.n = c->n; s.npad = c->npad; s.m = c->m;
    s.filt = c->d_filt; s.nfilt = c->nfilt; s.sigma = c->sigma; s.S = c->S; s.o_limit2 = c->o_limit2;
    s.tc_b = c->d_tcb; s.tc_ns = c->tc_ns; s.tc_gx = c->tc_g[0]; s.tc_gy = c->tc_g[1]; s.tc_gz = c->tc_g[2];
    s.tc_eps = (float)RT_TC_EPS;
    if (lbvh && c->bvh_valid) {
        rt_lbvh_fill(c->bvh, &s);
        s.bvh_pad0 = 16.f * 5.9604644775390625e-8f * c->scene_radius;
    }
    return s;
}

// brute force unless asked for / forced by size
int pick_accel(rt_ctx *c, uint32_t flags, bool *lbvh, bool hits = false)
{
    bool want_bvh = (flags & RT_FLAG_ACCEL_LBVH) != 0;
    bool want_brute = (flags & RT_FLAG_ACCEL_BRUTE) != 0;
    if (want_bvh && want_brute) return fail(RT_ERR_INVALID, "both RT_FLAG_ACCEL_BRUTE and RT_FLAG_ACCEL_LBVH set");
    // automatic choice: the parity build brute-forces whatever fits in shared memory; the throughput build switches
    // to the LBVH where it overtakes the tensor-filter brute force (measured on synthetic grids, tools/crossover.sh:
    // 730 spheres 1.96 vs 1.74 Gsamples/s, 1025 spheres 0.64 (FMA filter) vs 2.35)
    if (!want_bvh && !want_brute)
        want_bvh = c->n > ((flags & RT_FLAG_EXACT) ? RT_BRUTE_MAX_SPHERES : RT_AUTO_LBVH_SPHERES);
    if (!want_bvh && c->n > RT_BRUTE_MAX_SPHERES)
        return fail(RT_ERR_INVALID, "brute force supports at most RT_BRUTE_MAX_SPHERES spheres (shared-memory staging)");
    if (!want_bvh) {
        // the brute-force kernels stage the scene in shared memory: the parity kernels take RT_BRUTE_MAX_SPHERES, the
        // throughput kernel with the FMA-chain filter (sphere + filter records + its mask words) about 5.8 K spheres
        const bool exact = (flags & RT_FLAG_EXACT) != 0;
        const bool tensor = c->d_tcb && c->n >= RT_TC_MIN_SPHERES && !(flags & RT_FLAG_FILTER_FMA);
        const size_t need = hits ? rt_hits_smem_bytes(c->npad, c->nfilt, exact, false) : rt_render_smem_bytes(c->npad, c->nfilt, exact, false);
        if (!tensor && need > RT_TC_SMEM_LIMIT) {
            if (want_brute)
                return fail(RT_ERR_INVALID, "brute force: the scene does not fit in shared memory for this build "
                                            "(throughput build: about 5 800 spheres; RT_FLAG_EXACT: RT_BRUTE_MAX_SPHERES); use the LBVH");
            want_bvh = true;
        }
    }
    if (want_bvh && c->n < 2) return fail(RT_ERR_INVALID, "the LBVH needs at least 2 spheres");
    if (want_bvh && !c->bvh_valid) {
        if (!c->bvh) { int r = rt_lbvh_create(&c->bvh); if (r) return fail(RT_ERR_CUDA, "rt_lbvh_create failed"); }
        float ms = 0.f; unsigned launches = 0;
        cudaError_t e = rt_lbvh_build(c->bvh, c->d_sph, c->d_rad, c->n, c->stream, &ms, &launches);
        if (e != cudaSuccess) return cuda_fail(e, "rt_lbvh_build");
        c->hs.bvh_build_ms = ms;
        c->hs.launches += launches;
        c->bvh_valid = true;
    }
    *lbvh = want_bvh;
    return RT_OK;
}

int occupancy(rt_ctx *c, bool exact, bool lbvh, size_t smem, int *ctas)
{
    int &slot = c->occ[exact][lbvh];
    if (slot == 0 || c->occ_smem[exact][lbvh] != smem) {
        int v = 0;
        cudaError_t e = exact ? rtx::render_occupancy(smem, lbvh, &v) : lbvh ? rtb::render_occupancy(smem, lbvh, &v)
                                                                           : rtf::render_occupancy(smem, lbvh, &v);
        if (e != cudaSuccess) return cuda_fail(e, "render_occupancy");
        if (v < 1) return fail(RT_ERR_INVALID, "render kernel does not fit on an SM (scene too large for shared memory)");
        slot = v;
        c->occ_smem[exact][lbvh] = smem;
    }
    *ctas = slot;
    return RT_OK;
}

}  // namespace

extern "C" {

// rt_multi.cu reports through the same thread-local error string
int rt_internal_set_error(int code, const char *msg) { return fail(code, msg ? msg : ""); }

const char *rt_last_error(void) { return g_err.c_str(); }
const char *rt_version(void) { return "rtiow-b200 0.1 (sm_100a)"; }

int rt_create(rt_ctx **out, int device)
{
    if (!out) return fail(RT_ERR_INVALID, "rt_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(RT_ERR_NO_DEVICE, std::string("no usable CUDA device (") + cudaGetErrorString(e) +
                                          "); this backend has no CPU fallback");
    if (device < 0 || device >= count) return fail(RT_ERR_INVALID, "rt_create: device ordinal out of range");
    CK(cudaSetDevice(device));
    cudaDeviceProp prop{};
    CK(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        return fail(RT_ERR_NO_DEVICE, std::string("device '") + prop.name + "' is not sm_100-class; kernels are built for sm_100a only");
    rt_ctx *c = new (std::nothrow) rt_ctx();
    if (!c) return fail(RT_ERR_SYS, "out of host memory");
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    CK(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
    c->stream = c->own_stream;
    CK(cudaMalloc(&c->d_counter, kCounterRing * sizeof(uint32_t)));
    CK(cudaMalloc(&c->d_tile_ring, (size_t)kTileRing * sizeof(uint4)));
    CK(cudaMemsetAsync(c->d_counter, 0, kCounterRing * sizeof(uint32_t), c->stream));
    CK(cudaMalloc(&c->d_stats, 8 * sizeof(unsigned long long)));
    CK(cudaMemsetAsync(c->d_stats, 0, 8 * sizeof(unsigned long long), c->stream));
    CK(cudaMalloc(&c->d_scratch, 256));
    *out = c;
    return RT_OK;
}

int rt_destroy(rt_ctx *c)
{
    if (!c) return RT_OK;
    cudaSetDevice(c->device);
    for (int i = 0; i < kAux; ++i) if (c->aux[i]) cudaStreamSynchronize(c->aux[i]);
    cudaStreamSynchronize(c->stream);
    for (int i = 0; i < kAux; ++i) {
        if (c->aux[i]) cudaStreamDestroy(c->aux[i]);
        if (c->aux_tail[i]) cudaEventDestroy(c->aux_tail[i]);
    }
    if (c->aux_head) cudaEventDestroy(c->aux_head);
    free_scene(c);
    free_framebuffer(c);
    if (c->bvh) rt_lbvh_destroy(c->bvh);
    cudaFree(c->d_counter); cudaFree(c->d_tile_ring); cudaFree(c->d_stats); cudaFree(c->d_scratch);
    if (c->h_pin) cudaFreeHost(c->h_pin);
    cudaFree(c->d_hits); cudaFree(c->d_feed); cudaFree(c->d_self_cursors);
    for (auto &p : c->pending) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
    for (auto &p : c->free_events) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
    return RT_OK;
}

int rt_set_stream(rt_ctx *c, void *cuda_stream)
{
    if (!c) return fail(RT_ERR_INVALID, "ctx is NULL");
    CK(enter_ctx(c));
    CK(cudaStreamSynchronize(c->stream));
    c->stream = cuda_stream ? (cudaStream_t)cuda_stream : c->own_stream;
    return RT_OK;
}

int rt_get_stream(rt_ctx *c, void **cuda_stream)
{
    if (!c || !cuda_stream) return fail(RT_ERR_INVALID, "rt_get_stream: NULL argument");
    *cuda_stream = (void *)c->stream;
    return RT_OK;
}

int rt_sync(rt_ctx *c)
{
    if (!c) return fail(RT_ERR_INVALID, "ctx is NULL");
    CK(enter_ctx(c));
    CK(cudaStreamSynchronize(c->stream));
    return drain_events(c);
}

int rt_throttle(rt_ctx *c, uint32_t max_in_flight)
{
    if (!c) return fail(RT_ERR_INVALID, "ctx is NULL");
    CK(cudaSetDevice(c->device));      // a host-side wait only: no join, or the next overlapped chunk would queue behind this one
    // render calls are retired in submission order (overlapped chunks may finish out of order; waiting for the oldest is enough)
    while (c->pending.size() > max_in_flight) {
        EventPair p = c->pending.front();
        CK(cudaEventSynchronize(p.b));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, p.a, p.b));
        c->hs.render_ms += ms;
        c->hs.last_render_ms = ms;
        c->free_events.push_back(p);
        c->pending.erase(c->pending.begin());
    }
    return RT_OK;
}

int rt_scene_upload(rt_ctx *c, const rt_sphere *spheres, uint32_t n, const rt_material *materials, uint32_t m)
{
    if (!c || !spheres || !materials || n == 0 || m == 0) return fail(RT_ERR_INVALID, "rt_scene_upload: empty scene or NULL argument");
    if (m >= (1u << 30)) return fail(RT_ERR_INVALID, "rt_scene_upload: too many materials");
    CK(enter_ctx(c));
    for (uint32_t i = 0; i < m; ++i)
        if (materials[i].kind > 2u) return fail(RT_ERR_INVALID, "rt_scene_upload: unknown material kind");
    for (uint32_t i = 0; i < n; ++i)
        if (spheres[i].mat >= m) return fail(RT_ERR_INVALID, "rt_scene_upload: sphere references a missing material");

    uint32_t npad = (n + 31u) & ~31u;
    if (npad != c->npad || m != c->m || (n + RT_FGROUP - 1) / RT_FGROUP * RT_FGROUP != c->nfilt) {
        CK(cudaStreamSynchronize(c->stream));
        free_scene(c);
        CK(cudaMalloc(&c->d_sph, (size_t)npad * sizeof(float4)));
        CK(cudaMalloc(&c->d_rad, (size_t)npad * sizeof(float)));
        CK(cudaMalloc(&c->d_meta, (size_t)npad * sizeof(uint32_t)));
        CK(cudaMalloc(&c->d_mat, (size_t)m * sizeof(float4)));
        CK(cudaMalloc(&c->d_filt, (size_t)((n + RT_FGROUP - 1) / RT_FGROUP * RT_FGROUP) * sizeof(float4)));
    }
    const uint32_t nfilt = (n + RT_FGROUP - 1) / RT_FGROUP * RT_FGROUP;
    const uint32_t tc_tile = RT_TC_CHUNK * RT_TC_NBUF;   // whole tiles, a multiple of NBUF of them
    const uint32_t tc_ns = (n <= RT_TC_MAX_SPHERES) ? (n + tc_tile - 1u) / tc_tile * tc_tile : 0u;
    if (tc_ns != c->tc_ns || (tc_ns && !c->d_tcb)) {
        CK(cudaStreamSynchronize(c->stream));
        cudaFree(c->d_tcb); c->d_tcb = nullptr; c->tc_ns = 0;
        if (tc_ns) CK(cudaMalloc(&c->d_tcb, (size_t)tc_ns * 8 * sizeof(float4)));
        c->tc_ns = tc_ns;
    }
    size_t b_sph = (size_t)npad * 16, b_rad = (size_t)npad * 4, b_meta = (size_t)npad * 4, b_mat = (size_t)m * 16;
    size_t b_filt = (size_t)nfilt * 16, b_tcb = (size_t)tc_ns * 8 * 16;
    int r = ensure_pinned(c, b_sph + b_rad + b_meta + b_mat + b_filt + b_tcb);
    if (r) return r;
    CK(cudaStreamSynchronize(c->stream));   // staging buffer may still be in flight
    char *h = (char *)c->h_pin;
    float4 *hs = (float4 *)h;
    float *hr = (float *)(h + b_sph);
    uint32_t *hm = (uint32_t *)(h + b_sph + b_rad);
    float4 *hmat = (float4 *)(h + b_sph + b_rad + b_meta);
    float4 *hf = (float4 *)(h + b_sph + b_rad + b_meta + b_mat);
    // Scale of the conservative filter (rt_device.cuh): S = sigma^2, the largest power of four that keeps
    // S * 4 * Rs^2 * 4 below FLT_MAX, Rs = the farthest surface point from the origin (ray origins live there),
    // capped at 2^100.  Rays that start farther out than 2*Rs take the slow everything-is-a-candidate route.
    double Rs = 1.0;
    for (uint32_t i = 0; i < n; ++i) {
        const rt_sphere &sp = spheres[i];
        double d = std::sqrt((double)sp.cx * sp.cx + (double)sp.cy * sp.cy + (double)sp.cz * sp.cz) + std::fabs((double)sp.r);
        if (d > Rs) Rs = d;
    }
    int e2 = (int)std::floor((120.0 - std::ceil(std::log2(Rs * Rs))) / 2.0);
    if (e2 > 50) e2 = 50;
    if (e2 < -60) e2 = -60;
    const double sigma = std::ldexp(1.0, e2), S = sigma * sigma;
    for (uint32_t i = 0; i < nfilt; ++i) {
        if (i < n) {
            const rt_sphere &sp = spheres[i];
            double c2 = (double)sp.cx * sp.cx + (double)sp.cy * sp.cy + (double)sp.cz * sp.cz, r2 = (double)sp.r * sp.r;
            double nk = (r2 - c2) + RT_FILTER_EPS * (c2 + r2);
            hf[i] = make_float4(sp.cx, sp.cy, sp.cz, (float)(S * nk));
        } else {
            hf[i] = make_float4(0.f, 0.f, 0.f, -std::numeric_limits<float>::infinity());
        }
    }
    for (uint32_t i = 0; i < npad; ++i) {
        if (i < n) {
            const rt_sphere &s = spheres[i];
            hs[i] = make_float4(s.cx, s.cy, s.cz, s.r * s.r);   // radius*radius, src/raytracer.hxx:59
            hr[i] = s.r;
            hm[i] = s.mat | (materials[s.mat].kind << 30);
        } else {
            hs[i] = make_float4(0.f, 0.f, 0.f, -std::numeric_limits<float>::infinity());   // never a candidate
            hr[i] = 1.f;
            hm[i] = 0u;
        }
    }
    for (uint32_t i = 0; i < m; ++i) hmat[i] = make_float4(materials[i].r, materials[i].g, materials[i].b, materials[i].param);
    if (tc_ns) {
        // Sphere side of the bilinear form of rt_tensor.cuh, expanded about g = the component-wise median of the
        // centres (keeps |c - g| small for the many small spheres whatever the ground sphere does).
        // Slots: 9 terms {xx,yy,zz,xy,xz,yz,x,y,z} x (f_h, f_l, f_h) | w_h, w_m, w_l | 1, 1 with
        // w = r^2 - |c|^2 + eps (|c|^2 + r^2); hi parts are TF32 (10-bit mantissa, round to nearest), evaluated in double.
        // Column pos = 32 b + v holds sphere 32 b + (v < 16 ? 2 v : 2 (v - 16) + 1): the bit order of the I2IP.U2 sign masks.
        float *hb = (float *)(h + b_sph + b_rad + b_meta + b_mat + b_filt);
        std::memset(hb, 0, b_tcb);
        std::vector<float> med(n);
        float g[3];
        for (int ax = 0; ax < 3; ++ax) {
            for (uint32_t i = 0; i < n; ++i) med[i] = ax == 0 ? spheres[i].cx : ax == 1 ? spheres[i].cy : spheres[i].cz;
            std::nth_element(med.begin(), med.begin() + n / 2, med.end());
            g[ax] = std::isfinite(med[n / 2]) ? med[n / 2] : 0.f;
        }
        auto tf32 = [](double x) {
            float f = (float)x; uint32_t u; std::memcpy(&u, &f, 4);
            if ((u & 0x7f800000u) == 0x7f800000u) return f;
            u += 0x1000u; u &= 0xffffe000u; std::memcpy(&f, &u, 4); return f;
        };
        auto put = [&](uint32_t pos, uint32_t slot, float v) { hb[((size_t)(slot >> 2) * tc_ns + pos) * 4 + (slot & 3u)] = v; };
        for (uint32_t pos = 0; pos < tc_ns; ++pos) {
            const uint32_t v = pos & 31u, i = (pos & ~31u) + (v < 16u ? 2u * v : 2u * (v - 16u) + 1u);   // I2IP.U2 bit order
            if (i >= n) { put(pos, 27, -3.0e38f); continue; }           // padding: never a candidate
            const rt_sphere &sp = spheres[i];
            const double x = (double)sp.cx - g[0], y = (double)sp.cy - g[1], z = (double)sp.cz - g[2], r2 = (double)sp.r * sp.r;
            const double f[9] = {x * x, y * y, z * z, x * y, x * z, y * z, x, y, z};
            for (uint32_t t = 0; t < 9; ++t) {
                float fh = tf32(f[t]);
                float fl = (float)(f[t] - (double)fh);
                put(pos, 3 * t + 0, fh); put(pos, 3 * t + 1, fl); put(pos, 3 * t + 2, fh);
            }
            const double c2 = x * x + y * y + z * z;
            const double w = (r2 - c2) + RT_TC_EPS * (c2 + r2);
            float wh = tf32(w), wm = tf32(w - (double)wh);
            float wl = (float)(w - (double)wh - (double)wm);
            put(pos, 27, wh); put(pos, 28, wm); put(pos, 29, wl);
            put(pos, 30, 1.f); put(pos, 31, 1.f);
        }
        CK(cudaMemcpyAsync(c->d_tcb, hb, b_tcb, cudaMemcpyHostToDevice, c->stream));
        c->tc_g[0] = g[0]; c->tc_g[1] = g[1]; c->tc_g[2] = g[2];
    }
    CK(cudaMemcpyAsync(c->d_sph, hs, b_sph, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->d_rad, hr, b_rad, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->d_meta, hm, b_meta, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->d_mat, hmat, b_mat, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->d_filt, hf, b_filt, cudaMemcpyHostToDevice, c->stream));
    c->n = n; c->npad = npad; c->m = m; c->nfilt = nfilt;
    c->sigma = (float)sigma; c->S = (float)S; c->o_limit2 = (float)(4.0 * Rs * Rs); c->scene_radius = (float)Rs;
    c->bvh_valid = false;
    return RT_OK;
}

// src/camera.hxx:24-44, float arithmetic in the same order (this TU is built with -ffp-contract=off).
int rt_camera_set(rt_ctx *c, const float pos[3], const float lookat[3], const float up[3],
                  float aspect, float vfov_deg, float aperture, float focus_dist)
{
    if (!c || !pos || !lookat || !up) return fail(RT_ERR_INVALID, "rt_camera_set: NULL argument");
    struct v3 { float x, y, z; };
    auto sub = [](v3 a, v3 b) { return v3{a.x - b.x, a.y - b.y, a.z - b.z}; };
    auto add = [](v3 a, v3 b) { return v3{a.x + b.x, a.y + b.y, a.z + b.z}; };
    auto muls = [](v3 a, float s) { return v3{a.x * s, a.y * s, a.z * s}; };
    auto cross = [](v3 a, v3 b) { return v3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; };
    auto normalize = [](v3 a) {
        float len = std::sqrt(a.x * a.x + a.y * a.y + a.z * a.z);
        if (std::fabs(len) > FLT_MIN) { a.x /= len; a.y /= len; a.z /= len; }
        return a;
    };
    v3 position{pos[0], pos[1], pos[2]}, look{lookat[0], lookat[1], lookat[2]}, upv{up[0], up[1], up[2]};
    float theta = (vfov_deg * static_cast<float>(0.01745329251994329576923690768489)) / 2.f;
    float height = std::tan(theta);
    float width = height * aspect;
    v3 w = normalize(sub(position, look));
    v3 u = normalize(cross(upv, w));
    v3 v = normalize(cross(w, u));
    v3 llc = sub(position, muls(add(add(muls(u, width), muls(v, height)), w), focus_dist));
    v3 hor = muls(muls(muls(u, width), focus_dist), 2.f);
    v3 ver = muls(muls(muls(v, height), focus_dist), 2.f);
    RtDeviceCamera &d = c->cam;
    d.ox = position.x; d.oy = position.y; d.oz = position.z;
    d.llx = llc.x; d.lly = llc.y; d.llz = llc.z;
    d.hx = hor.x; d.hy = hor.y; d.hz = hor.z;
    d.vx = ver.x; d.vy = ver.y; d.vz = ver.z;
    d.lens_radius = aperture / 2.f;
    d.bux = u.x; d.buy = u.y; d.buz = u.z;
    d.bvx = v.x; d.bvy = v.y; d.bvz = v.z;
    c->have_cam = true;
    return RT_OK;
}

int rt_camera_get(rt_ctx *c, float out[13])
{
    if (!c || !out || !c->have_cam) return fail(RT_ERR_INVALID, "rt_camera_get: no camera set");
    const RtDeviceCamera &d = c->cam;
    const float v[13] = {d.ox, d.oy, d.oz, d.llx, d.lly, d.llz, d.hx, d.hy, d.hz, d.vx, d.vy, d.vz, d.lens_radius};
    std::memcpy(out, v, sizeof v);
    return RT_OK;
}

int rt_framebuffer(rt_ctx *c, uint32_t width, uint32_t height)
{
    if (!c || width == 0 || height == 0) return fail(RT_ERR_INVALID, "rt_framebuffer: zero size");
    if ((unsigned long long)width * height >= (1ull << 31)) return fail(RT_ERR_INVALID, "rt_framebuffer: image too large");
    CK(enter_ctx(c));
    if (width != c->W || height != c->H) {
        CK(cudaStreamSynchronize(c->stream));
        free_framebuffer(c);
        size_t npix = (size_t)width * height;
        CK(cudaMalloc(&c->d_accum, npix * 4 * sizeof(unsigned long long)));
        CK(cudaMalloc(&c->d_rgb8, npix * 3 + 16));
        CK(cudaMalloc(&c->d_meanf, npix * 3 * sizeof(float) + 64));
        c->W = width; c->H = height;
    }
    return rt_clear(c);
}

int rt_clear(rt_ctx *c)
{
    if (!c || !c->d_accum) return fail(RT_ERR_INVALID, "rt_clear: no framebuffer");
    CK(enter_ctx(c));
    CK(cudaMemsetAsync(c->d_accum, 0, (size_t)c->W * c->H * 4 * sizeof(unsigned long long), c->stream));
    return RT_OK;
}

// rt_render_queue: the launch pulls its tiles from the device-side queue instead of rendering the whole list
struct QueueMode { uint32_t *cursors; uint32_t rank, nranks; };

// Common scheduling of one or more launches over a tile list.  `tiles` all fit in tw x th pixel slots.
static int render_tiles_impl(rt_ctx *c, const rt_tile *tiles, uint32_t ntiles, uint32_t tw, uint32_t th,
                             const rt_render_params *p, const QueueMode *qm = nullptr)
{
    if (!c->d_sph) return fail(RT_ERR_INVALID, "rt_render: no scene uploaded");
    if (!c->have_cam) return fail(RT_ERR_INVALID, "rt_render: no camera set");
    if (!c->d_accum) return fail(RT_ERR_INVALID, "rt_render: no framebuffer");
    if (p->max_depth == 0) return fail(RT_ERR_INVALID, "rt_render: max_depth must be >= 1");
    if ((unsigned long long)p->sample_base + p->spp > 0xffffffffull) return fail(RT_ERR_INVALID, "rt_render: sample index overflow");
    if (ntiles == 0 || p->spp == 0 || tw == 0 || th == 0) return RT_OK;   // nothing to add
    for (uint32_t i = 0; i < ntiles; ++i) {
        const rt_tile &t = tiles[i];
        if ((unsigned long long)t.x0 + t.w > c->W || (unsigned long long)t.y0 + t.h > c->H)
            return fail(RT_ERR_INVALID, "rt_render: tile outside the framebuffer");
        if (t.w > tw || t.h > th) return fail(RT_ERR_INVALID, "rt_render_tiles: tile larger than the nominal tile size");
    }
    const unsigned long long TP = (unsigned long long)tw * th;
    if (TP * p->spp >= (1ull << 31)) return fail(RT_ERR_INVALID, "rt_render: tile_pixels * spp must stay below 2^31; use smaller tiles or split spp with sample_base");
    const bool overlap = !qm && (p->flags & RT_FLAG_OVERLAP) != 0;
    if (overlap) CK(cudaSetDevice(c->device));        // no join: this launch may run beside the chunks already queued
    else CK(enter_ctx(c));
    if (qm) {
        if (TP % RT_BATCH != 0) return fail(RT_ERR_INVALID, "rt_render_queue: tile_w * tile_h must be a multiple of 256 (one work batch never straddles two tiles)");
        if (ntiles > kTileRing) return fail(RT_ERR_INVALID, "rt_render_queue: too many tiles");
        if (p->flags & RT_FLAG_FILTER_CHECK) return fail(RT_ERR_INVALID, "rt_render_queue: RT_FLAG_FILTER_CHECK is a single-GPU self-check");
        const uint32_t want = ntiles + 4096u + 64u;   // tiles this GPU can acquire + the dry slots its warps touch at the end
        if (want > c->feed_slots) {
            CK(cudaStreamSynchronize(c->stream));
            cudaFree(c->d_feed); c->d_feed = nullptr; c->feed_slots = 0;
            CK(cudaMalloc((void **)&c->d_feed, sizeof(RtTileFeed) + 3u * (size_t)want * sizeof(uint32_t)));
            c->feed_slots = want;
        }
    }

    const bool exact = (p->flags & RT_FLAG_EXACT) != 0;
    bool lbvh = false;
    int r = pick_accel(c, p->flags, &lbvh);
    if (r) return r;
    // throughput build, brute force: the candidate filter runs on the tensor cores unless the scene is too large
    // for the shared-memory sphere matrix or the caller asks for the FMA-pipe filter
    // (below RT_TC_MIN_SPHERES one 128-column tile is mostly padding and the FMA-pipe loop is shorter than the handshake)
    // (RT_FLAG_EXACT takes the same kernel with the parity shading; with RT_FLAG_FILTER_FMA it takes the rtx kernels)
    const bool tensor = !lbvh && c->d_tcb && c->n >= RT_TC_MIN_SPHERES && !(p->flags & RT_FLAG_FILTER_FMA);
    bool smem_mats = false;
    size_t smem = tensor ? rt_render_tc_smem_bytes(c->npad, c->tc_ns, c->m, &smem_mats) : rt_render_smem_bytes(c->npad, c->nfilt, exact, lbvh);
    // throughput build, a handful of spheres: the one-path-per-thread kernel without filter code (k_render_small)
    const bool small = !exact && !lbvh && c->n <= RT_SMALL_SCENE && !(p->flags & RT_FLAG_FILTER_CHECK);
    int ctas = 1;                      // the tensor kernel owns the SM (all 512 TMEM columns): one CTA per SM
    if (small) {
        if (c->occ_small == 0) CK(rtf::small_occupancy(&c->occ_small));
        ctas = c->occ_small > 0 ? c->occ_small : 1;
    } else if (!tensor) {
        r = occupancy(c, exact, lbvh, smem, &ctas);
        if (r) return r;
    }

    EventPair ev;
    r = get_events(c, &ev);
    if (r) return r;
    cudaStream_t st = c->stream;
    int aux_slot = -1;
    if (overlap) {
        const uint32_t band = (uint32_t)((((1ull << 31) - 1) / p->spp) / TP), nbands = (ntiles + band - 1u) / band;
        if (c->counter_slot + nbands > kCounterRing) {   // the ring would be cleared inside this call: clear it now, behind every chunk still counting in it
            CK(enter_ctx(c));
            CK(cudaMemsetAsync(c->d_counter, 0, kCounterRing * sizeof(uint32_t), c->stream));
            c->counter_slot = 0;
        }
        if (ntiles > 1 && ntiles <= kTileRing && c->tile_ring_pos + ntiles > kTileRing) {   // same for the tile-list ring
            CK(enter_ctx(c));
            c->tile_ring_pos = 0;
        }
        CK(overlap_stream(c, &st, &aux_slot));
    }
    CK(cudaEventRecord(ev.a, st));

    uint4 *d_tiles = nullptr;
    uint4 tile0 = make_uint4(tiles[0].x0, tiles[0].y0, tiles[0].w, tiles[0].h);
    bool tiles_from_pool = false;
    if (ntiles > 1 || qm) {
        static_assert(sizeof(rt_tile) == sizeof(uint4), "rt_tile must be 16 bytes");
        if (ntiles <= kTileRing) {
            // (without RT_FLAG_OVERLAP everything is on one stream and a wrapped slice is only rewritten behind its readers)
            if (c->tile_ring_pos + ntiles > kTileRing) c->tile_ring_pos = 0;
            d_tiles = c->d_tile_ring + c->tile_ring_pos;
            c->tile_ring_pos += ntiles;
        } else {
            CK(cudaMallocAsync((void **)&d_tiles, (size_t)ntiles * sizeof(uint4), st));
            tiles_from_pool = true;
        }
        // pageable source: the runtime stages the bytes before returning, so `tiles` may be reused by the caller
        CK(cudaMemcpyAsync(d_tiles, tiles, (size_t)ntiles * sizeof(uint4), cudaMemcpyHostToDevice, st));
    }
    if (qm) {                                         // fresh feed: item counter 0, every slot RT_FEED_EMPTY
        CK(cudaMemsetAsync(c->d_feed, 0, sizeof(RtTileFeed), st));
        CK(cudaMemsetAsync(c->d_feed->slot, 0xff, (size_t)c->feed_slots * sizeof(uint32_t), st));
        CK(cudaMemsetAsync(c->d_feed->slot + c->feed_slots, 0, (size_t)c->feed_slots * sizeof(uint32_t), st));
    } else {
        CK(rt_launch_add_count(c->d_accum, c->W, tile0, d_tiles, ntiles, (uint32_t)TP, p->spp, st));
        c->hs.launches += 1;
    }

    RtRenderArgs a{};
    a.sc = device_scene(c, lbvh);
    a.cam = c->cam;
    a.accum = c->d_accum;
    a.W = c->W; a.H = c->H;
    a.tile0 = tile0;
    a.tiles = d_tiles;
    a.TP = (uint32_t)TP;
    a.TPS = (uint32_t)(TP * p->spp);
    a.mTP = (uint32_t)(0x100000000ull / a.TP); a.mTPS = (uint32_t)(0x100000000ull / a.TPS);
    a.mW0 = (uint32_t)(0x100000000ull / (tile0.z ? tile0.z : 1u));
    a.spp = p->spp; a.depth = p->max_depth; a.sample_base = p->sample_base;
    a.key0 = (uint32_t)p->seed; a.key1 = (uint32_t)(p->seed >> 32);
    a.stats = c->d_stats;
    a.check = (p->flags & RT_FLAG_FILTER_CHECK) ? 1u : 0u;
    a.book = (p->flags & RT_FLAG_BOOK_OPTICS) ? 1u : 0u;
    a.smem_mats = smem_mats ? 1u : 0u;
    const bool test_stragglers = (p->flags & RT_FLAG_TEST_STRAGGLERS) != 0;
    a.coop_lanes = test_stragglers ? 32u : (uint32_t)RT_BVH_COOP;
    a.coop_after = test_stragglers ? 1u : (uint32_t)RT_BVH_COOP_AFTER;
    a.coop_cap_odd = test_stragglers ? 8u : (uint32_t)RT_BVH_COOP_CAP;
    if (qm) {
        a.feed = c->d_feed; a.feed_slots = c->feed_slots;
        a.q_cursors = qm->cursors; a.q_rank = qm->rank; a.q_nranks = qm->nranks; a.q_ntiles = ntiles;

    }

    // bands of consecutive tiles so that items = nt * TP * spp stays below 2^31 (32-bit item arithmetic);
    // queue mode: ONE launch, the items are tile-relative and the feed's counter is 64 bits wide
    uint32_t per_band = qm ? ntiles : (uint32_t)((((1ull << 31) - 1) / p->spp) / TP);
    for (uint32_t t0 = 0; t0 < ntiles; t0 += per_band) {
        uint32_t nt = (ntiles - t0 < per_band) ? (ntiles - t0) : per_band;
        a.t0 = t0;
        a.P = nt * (uint32_t)TP;
        a.total_items = qm ? 0x7fffffffu : a.P * p->spp;
        if (c->counter_slot == kCounterRing) {
            CK(cudaMemsetAsync(c->d_counter, 0, kCounterRing * sizeof(uint32_t), st));
            c->counter_slot = 0;
        }
        a.work_counter = c->d_counter + c->counter_slot++;
        const unsigned long long per_cta = tensor ? (unsigned long long)(RT_TC_WGS * 128)
                                           : small ? (unsigned long long)RT_THREADS : (unsigned long long)RT_THREADS * RT_RPT;
        unsigned long long want_ctas = qm ? ~0ull : (a.total_items + per_cta - 1) / per_cta;
        int grid = c->sm_count * ctas;
        if ((unsigned long long)grid > want_ctas) grid = (int)want_ctas;
        if (grid < 1) grid = 1;
        const bool q = qm != nullptr;
        cudaError_t e = small ? rtf::launch_render_small(a, grid, st, q)
                        : tensor ? (exact ? rtx::launch_render_tc(a, grid, smem, st, q) : rtf::launch_render_tc(a, grid, smem, st, q))
                        : exact ? rtx::launch_render(a, grid, smem, st, lbvh, q)
                        : lbvh ? rtb::launch_render(a, grid, smem, st, lbvh, q)       // throughput LBVH kernel: its own unit, no FMA contraction
                               : rtf::launch_render(a, grid, smem, st, lbvh, q);
        if (e != cudaSuccess) return cuda_fail(e, "launch_render");
        c->hs.launches += 1;
        c->hs.last_kernel = small ? RT_KERNEL_RENDER_SMALL
                            : tensor ? (exact ? RT_KERNEL_RENDER_TC_PARITY : RT_KERNEL_RENDER_TC)
                            : lbvh ? (exact ? RT_KERNEL_RENDER_BVH_PARITY : RT_KERNEL_RENDER_BVH)
                                   : (exact ? RT_KERNEL_RENDER_PARITY : RT_KERNEL_RENDER_FMA);
    }
    if (tiles_from_pool) CK(cudaFreeAsync(d_tiles, st));
    if (qm) {                                         // the sample count of the tiles this GPU ended up rendering
        CK(rt_launch_add_count_feed(c->d_accum, c->W, d_tiles, c->d_feed->slot, c->feed_slots, (uint32_t)TP, p->spp, st));
        c->hs.launches += 1;
        c->feed_tiles = d_tiles; c->feed_TP = (uint32_t)TP; c->feed_valid = true;
    }
    CK(cudaEventRecord(ev.b, st));
    if (aux_slot >= 0) { CK(cudaEventRecord(c->aux_tail[aux_slot], st)); c->aux_dirty[aux_slot] = true; }
    c->pending.push_back(ev);
    return RT_OK;
}

int rt_render(rt_ctx *c, const rt_tile *tile, const rt_render_params *p)
{
    if (!c || !tile || !p) return fail(RT_ERR_INVALID, "rt_render: NULL argument");
    if (tile->w == 0 || tile->h == 0 || p->spp == 0) {
        // still validate the call order so errors do not depend on the tile being empty
        if (!c->d_sph || !c->have_cam || !c->d_accum) return render_tiles_impl(c, tile, 1, 1, 1, p);
        return RT_OK;
    }
    const unsigned long long npx = (unsigned long long)tile->w * tile->h;
    // A large rectangle is rendered TILE BY TILE through the context's own one-range queue (the path the multi-GPU frames
    // take, with this GPU as rank 0 of 1): one persistent launch whatever the size (64-bit item counter), and tile-major
    // work order -- a tile's accumulators stay cache-resident while all of its samples are traced (measured on the huge
    // scene, 1280x720x100 spp: 1.60 vs 1.46 Gsamples/s for the sample-major order of a single W x H tile).
    constexpr uint32_t kTw = 128, kTh = 64;
#ifndef RT_RENDER_SELF_QUEUE
#define RT_RENDER_SELF_QUEUE 1
#endif
#ifndef RT_RENDER_SELF_QUEUE_LBVH
#define RT_RENDER_SELF_QUEUE_LBVH 1     // the LBVH kernel too, since it finishes its stragglers cooperatively (rt_bvh_render.cuh: bvh_coop_query):
                                       // 1M spheres, 1920x1080x256 spp: 569-577 ms through the queue, 600 ms as one sample-major tile
                                       // (before: 606-640 vs 595-603 -- the tile-major frame happened to end on a few very long paths)
#endif
#ifndef RT_RENDER_SELF_QUEUE_LBVH_SPP
#define RT_RENDER_SELF_QUEUE_LBVH_SPP 128   // ... from this many samples per pixel on: at 32 spp the same frame is 8 % faster as one tile
                                            // (850 vs 784 Msamples/s; 488 spheres through the LBVH at 100 spp: 1340 vs 1275)
#endif
    const bool will_lbvh = (p->flags & RT_FLAG_ACCEL_LBVH) ||
                           (!(p->flags & RT_FLAG_ACCEL_BRUTE) && c->n > ((p->flags & RT_FLAG_EXACT) ? RT_BRUTE_MAX_SPHERES : RT_AUTO_LBVH_SPHERES));
    if (RT_RENDER_SELF_QUEUE && (!will_lbvh || (RT_RENDER_SELF_QUEUE_LBVH && p->spp >= RT_RENDER_SELF_QUEUE_LBVH_SPP)) && npx >= 4ull * kTw * kTh &&
        !(p->flags & (RT_FLAG_OVERLAP | RT_FLAG_FILTER_CHECK)) &&
        (unsigned long long)kTw * kTh * p->spp < (1ull << 31) &&
        ((unsigned long long)(tile->w + kTw - 1) / kTw) * ((tile->h + kTh - 1) / kTh) <= kTileRing) {
        std::vector<rt_tile> tl;
        for (uint32_t y = 0; y < tile->h; y += kTh)
            for (uint32_t x = 0; x < tile->w; x += kTw)
                tl.push_back(rt_tile{tile->x0 + x, tile->y0 + y, (tile->w - x < kTw) ? tile->w - x : kTw, (tile->h - y < kTh) ? tile->h - y : kTh});
        if (!c->d_sph || !c->have_cam || !c->d_accum) return render_tiles_impl(c, tile, 1, 1, 1, p);   // call-order errors first
        CK(enter_ctx(c));
        if (!c->d_self_cursors) {
            CK(cudaMalloc((void **)&c->d_self_cursors, (size_t)RT_QUEUE_SLOTS * RT_QUEUE_STRIDE * sizeof(uint32_t)));
        }
        CK(rt_launch_queue_reset(c->d_self_cursors, (uint32_t)tl.size(), 1, c->stream));
        c->hs.launches += 1;
        QueueMode qm{c->d_self_cursors, 0, 1};
        return render_tiles_impl(c, tl.data(), (uint32_t)tl.size(), kTw, kTh, p, &qm);
    }
    if (npx * p->spp < (1ull << 31)) return render_tiles_impl(c, tile, 1, tile->w, tile->h, p);
    // big rectangle: cut it into row strips whose item count fits 31 bits and run them as one tile list
    unsigned long long rows = (((1ull << 31) - 1) / p->spp) / tile->w;
    if (rows == 0) return fail(RT_ERR_INVALID, "rt_render: width * spp must stay below 2^31; split spp with sample_base");
    if (rows > tile->h) rows = tile->h;
    std::vector<rt_tile> strips;
    for (uint32_t y = 0; y < tile->h; y += (uint32_t)rows) {
        uint32_t h = (tile->h - y < rows) ? (tile->h - y) : (uint32_t)rows;
        strips.push_back(rt_tile{tile->x0, tile->y0 + y, tile->w, h});
    }
    return render_tiles_impl(c, strips.data(), (uint32_t)strips.size(), tile->w, (uint32_t)rows, p);
}

int rt_render_tiles(rt_ctx *c, const rt_tile *tiles, uint32_t ntiles, uint32_t tile_w, uint32_t tile_h,
                    const rt_render_params *p)
{
    if (!c || !p || (ntiles && !tiles)) return fail(RT_ERR_INVALID, "rt_render_tiles: NULL argument");
    if (ntiles == 0) return RT_OK;
    return render_tiles_impl(c, tiles, ntiles, tile_w, tile_h, p);
}

// grow-only device scratch of the hit passes: [index i32 | t f32 | normal 3 x f32 | rays 6 x f32] x total
static int hits_scratch(rt_ctx *c, size_t bytes)
{
    if (bytes <= c->d_hits_bytes) return RT_OK;
    CK(cudaStreamSynchronize(c->stream));
    cudaFree(c->d_hits); c->d_hits = nullptr; c->d_hits_bytes = 0;
    CK(cudaMalloc(&c->d_hits, bytes + bytes / 4));
    c->d_hits_bytes = bytes + bytes / 4;
    return RT_OK;
}

static int hits_common(rt_ctx *c, RtHitArgs &a, const float *rays6, unsigned long long total, uint32_t flags,
                       int32_t *sphere_index, float *t, float *normal_xyz)
{
    const bool exact = (flags & RT_FLAG_EXACT) != 0;
    bool lbvh = false;
    int r = pick_accel(c, flags, &lbvh, true);
    if (r) return r;
    r = hits_scratch(c, (size_t)total * (rays6 ? 44u : 20u));
    if (r) return r;
    int32_t *d_idx = (int32_t *)c->d_hits;
    float *d_t = (float *)(d_idx + total), *d_n = d_t + total, *d_rays = d_n + 3 * total;
    if (rays6) {
        CK(cudaMemcpyAsync(d_rays, rays6, (size_t)total * 24, cudaMemcpyHostToDevice, c->stream));
        a.rays = d_rays; a.nrays = total;
    }
    a.sc = device_scene(c, lbvh);
    a.idx = d_idx; a.t = d_t; a.normal = d_n;
    a.stats = c->d_stats;
    // throughput build on a tensor-eligible scene: the closest hit of k_render_tc itself (k_hits_tc, rt_tensor.cuh)
    const bool tensor = !exact && !lbvh && c->d_tcb && c->n >= RT_TC_MIN_SPHERES && !(flags & RT_FLAG_FILTER_FMA);
    cudaError_t e;
    if (tensor) {
        const size_t smem = rt_render_tc_smem_bytes(c->npad, c->tc_ns, c->m, nullptr);
        unsigned long long want = (total + RT_TC_WGS * 128ull - 1) / (RT_TC_WGS * 128ull);
        int grid = (unsigned long long)c->sm_count > want ? (int)want : c->sm_count;
        e = rtf::launch_hits_tc(a, grid < 1 ? 1 : grid, smem, c->stream);
    } else {
        e = exact ? rtx::launch_hits(a, c->stream, lbvh) : rtf::launch_hits(a, c->stream, lbvh);
    }
    if (e != cudaSuccess) return cuda_fail(e, "launch_hits");
    c->hs.launches += 1;
    CK(cudaMemcpyAsync(sphere_index, d_idx, (size_t)total * 4, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(t, d_t, (size_t)total * 4, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(normal_xyz, d_n, (size_t)total * 12, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return RT_OK;
}

int rt_primary_hits(rt_ctx *c, const rt_tile *tile, uint32_t flags, int32_t *sphere_index, float *t, float *normal_xyz)
{
    if (!c || !tile || !sphere_index || !t || !normal_xyz) return fail(RT_ERR_INVALID, "rt_primary_hits: NULL argument");
    if (!c->d_sph || !c->have_cam) return fail(RT_ERR_INVALID, "rt_primary_hits: scene and camera must be set");
    if (c->W == 0) return fail(RT_ERR_INVALID, "rt_primary_hits: call rt_framebuffer first (it defines W x H)");
    if (tile->w == 0 || tile->h == 0) return RT_OK;
    if ((unsigned long long)tile->x0 + tile->w > c->W || (unsigned long long)tile->y0 + tile->h > c->H)
        return fail(RT_ERR_INVALID, "rt_primary_hits: tile outside the framebuffer");
    CK(enter_ctx(c));
    RtHitArgs a{};
    a.cam = c->cam;
    a.W = c->W; a.H = c->H; a.x0 = tile->x0; a.y0 = tile->y0; a.w = tile->w; a.h = tile->h;
    a.rays = nullptr; a.nrays = 0;
    return hits_common(c, a, nullptr, (unsigned long long)tile->w * tile->h, flags, sphere_index, t, normal_xyz);
}

int rt_closest_hits(rt_ctx *c, const float *rays6, uint64_t n, uint32_t flags, int32_t *sphere_index, float *t, float *normal_xyz)
{
    if (!c || !sphere_index || !t || !normal_xyz) return fail(RT_ERR_INVALID, "rt_closest_hits: NULL argument");
    if (!c->d_sph) return fail(RT_ERR_INVALID, "rt_closest_hits: no scene uploaded");
    if (n == 0) return RT_OK;
    if (!rays6) return fail(RT_ERR_INVALID, "rt_closest_hits: rays is NULL");
    if (n >= (1ull << 32)) return fail(RT_ERR_INVALID, "rt_closest_hits: at most 2^32 - 1 rays per call");
    CK(enter_ctx(c));
    RtHitArgs a{};
    return hits_common(c, a, rays6, n, flags, sphere_index, t, normal_xyz);
}

int rt_resolve_device(rt_ctx *c)
{
    if (!c || !c->d_accum) return fail(RT_ERR_INVALID, "rt_resolve_device: no framebuffer");
    CK(enter_ctx(c));
    CK(rt_launch_resolve(c->d_accum, (unsigned long long)c->W * c->H, c->d_rgb8, c->d_meanf, c->stream));
    c->hs.launches += 1;
    return RT_OK;
}

int rt_readback_rgb8(rt_ctx *c, uint8_t *dst)
{
    if (!c || !dst) return fail(RT_ERR_INVALID, "rt_readback_rgb8: NULL argument");
    int r = rt_resolve_device(c);
    if (r) return r;
    size_t bytes = (size_t)c->W * c->H * 3;
    r = ensure_pinned(c, bytes);
    if (r) return r;
    CK(cudaMemcpyAsync(c->h_pin, c->d_rgb8, bytes, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    std::memcpy(dst, c->h_pin, bytes);
    return drain_events(c);
}

int rt_readback_f32(rt_ctx *c, float *dst)
{
    if (!c || !dst) return fail(RT_ERR_INVALID, "rt_readback_f32: NULL argument");
    int r = rt_resolve_device(c);
    if (r) return r;
    CK(cudaMemcpyAsync(dst, c->d_meanf, (size_t)c->W * c->H * 12, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return drain_events(c);
}

int rt_readback_accum(rt_ctx *c, uint64_t *dst)
{
    if (!c || !dst || !c->d_accum) return fail(RT_ERR_INVALID, "rt_readback_accum: NULL argument or no framebuffer");
    CK(enter_ctx(c));
    CK(cudaMemcpyAsync(dst, c->d_accum, (size_t)c->W * c->H * 32, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return drain_events(c);
}

int rt_device_accum(rt_ctx *c, void **dptr, size_t *bytes)
{
    if (!c || !dptr || !c->d_accum) return fail(RT_ERR_INVALID, "rt_device_accum: no framebuffer");
    *dptr = c->d_accum;
    if (bytes) *bytes = (size_t)c->W * c->H * 32;
    return RT_OK;
}

int rt_device_rgb8(rt_ctx *c, void **dptr, size_t *bytes)
{
    if (!c || !dptr || !c->d_rgb8) return fail(RT_ERR_INVALID, "rt_device_rgb8: no framebuffer");
    *dptr = c->d_rgb8;
    if (bytes) *bytes = (size_t)c->W * c->H * 3;
    return RT_OK;
}

int rt_render_frame(rt_ctx *c, const rt_sphere *spheres, uint32_t n, const rt_material *materials, uint32_t m,
                    const float cam[13], uint32_t width, uint32_t height, const rt_render_params *params, uint8_t *rgb8_out)
{
    if (!c || !cam || !params || !rgb8_out) return fail(RT_ERR_INVALID, "rt_render_frame: NULL argument");
    int r = rt_scene_upload(c, spheres, n, materials, m);
    if (r) return r;
    r = rt_camera_set(c, cam, cam + 3, cam + 6, cam[9], cam[10], cam[11], cam[12]);
    if (r) return r;
    r = rt_framebuffer(c, width, height);
    if (r) return r;
    rt_tile full = {0, 0, width, height};
    r = rt_render(c, &full, params);
    if (r) return r;
    return rt_readback_rgb8(c, rgb8_out);
}

// Progressive rendering: the accumulation of src/main.cxx:205-207 split into passes, with previews
// streamed to the host while the next pass renders.
//   render stream : [pass k] [resolve -> preview buffer i] (event R_i) [pass k+1] ...
//   copy stream   : wait R_i, D2H preview buffer i -> pinned buffer i (event C_i)
//   host          : enqueue pass k+1 first, then wait C_i and hand pinned buffer i to the callback
// The resolve runs in stream order between two passes, so a preview is exactly the image of the samples so far (never a
// torn mix with the pass that follows); two device + two pinned buffers let copy and callback overlap the rendering.
int rt_render_progressive(rt_ctx *c, const rt_render_params *p, uint32_t spp_per_pass, uint32_t preview_every,
                          rt_preview_fn cb, void *user, uint8_t *rgb8_out, uint32_t *spp_done_out)
{
    if (!c || !p) return fail(RT_ERR_INVALID, "rt_render_progressive: NULL argument");
    if (spp_per_pass == 0) return fail(RT_ERR_INVALID, "rt_render_progressive: spp_per_pass must be > 0");
    if (!c->d_sph || !c->have_cam || !c->d_accum)
        return fail(RT_ERR_INVALID, "rt_render_progressive: scene, camera and framebuffer must be set first");
    CK(enter_ctx(c));
    const size_t bytes = (size_t)c->W * c->H * 3;
    const bool previews = cb && preview_every > 0;
    constexpr int kPrev = 2;                 // preview buffers in flight (device + pinned host)
    unsigned char *d_prev[kPrev] = {nullptr, nullptr};
    unsigned char *h_prev[kPrev] = {nullptr, nullptr};
    cudaStream_t copy = nullptr;
    cudaEvent_t R[kPrev] = {nullptr, nullptr}, Cv[kPrev] = {nullptr, nullptr};
    int rc = RT_OK;
    cudaError_t e = cudaSuccess;
    auto cleanup = [&]() {
        if (copy) { cudaStreamSynchronize(copy); cudaStreamDestroy(copy); }
        for (int i = 0; i < kPrev; ++i) {
            if (d_prev[i]) cudaFree(d_prev[i]);
            if (h_prev[i]) cudaFreeHost(h_prev[i]);
            if (R[i]) cudaEventDestroy(R[i]);
            if (Cv[i]) cudaEventDestroy(Cv[i]);
        }
    };
    if (previews) {
        e = cudaStreamCreateWithFlags(&copy, cudaStreamNonBlocking);
        for (int i = 0; i < kPrev && e == cudaSuccess; ++i) {
            e = cudaMalloc((void **)&d_prev[i], bytes);
            if (e == cudaSuccess) e = cudaMallocHost((void **)&h_prev[i], bytes);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&R[i], cudaEventDisableTiming);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&Cv[i], cudaEventDisableTiming);
        }
        if (e != cudaSuccess) { cleanup(); return cuda_fail(e, "rt_render_progressive: preview buffers"); }
    }
    const rt_tile full = {0, 0, c->W, c->H};
    uint32_t done = 0, pass = 0, nprev = 0;
    int pending = -1; uint32_t pending_pass = 0, pending_spp = 0;     // preview in flight: buffer index, its pass and sample count
    bool stop = false;
    auto deliver = [&]() -> int {                                    // wait for the copy of the pending preview and call back
        if (pending < 0) return RT_OK;
        cudaError_t ee = cudaEventSynchronize(Cv[pending]);
        if (ee != cudaSuccess) return cuda_fail(ee, "rt_render_progressive: preview copy");
        if (cb(user, pending_pass, pending_spp, h_prev[pending]) != 0) stop = true;
        pending = -1;
        return RT_OK;
    };
    while (done < p->spp && !stop) {
        rt_render_params q = *p;
        q.flags &= ~(uint32_t)RT_FLAG_OVERLAP;        // passes cover the same pixels and previews are taken between them: keep stream order
        q.spp = (p->spp - done < spp_per_pass) ? (p->spp - done) : spp_per_pass;
        q.sample_base = p->sample_base + done;
        rc = rt_render(c, &full, &q);
        if (rc) break;
        done += q.spp; ++pass;
        // the previous preview is delivered while this pass runs on the GPU
        rc = deliver();
        if (rc) break;
        if (previews && !stop && (pass % preview_every == 0) && done < p->spp) {
            const int i = (int)(nprev % kPrev);
            e = rt_launch_resolve(c->d_accum, (unsigned long long)c->W * c->H, d_prev[i], nullptr, c->stream);
            if (e == cudaSuccess) e = cudaEventRecord(R[i], c->stream);
            if (e == cudaSuccess) e = cudaStreamWaitEvent(copy, R[i], 0);
            if (e == cudaSuccess) e = cudaMemcpyAsync(h_prev[i], d_prev[i], bytes, cudaMemcpyDeviceToHost, copy);
            if (e == cudaSuccess) e = cudaEventRecord(Cv[i], copy);
            if (e != cudaSuccess) { rc = cuda_fail(e, "rt_render_progressive: preview"); break; }
            c->hs.launches += 1;
            pending = i; pending_pass = pass; pending_spp = done; ++nprev;
        }
    }
    if (rc == RT_OK && !stop) rc = deliver();                        // a preview taken before the last pass
    if (spp_done_out) *spp_done_out = done;
    if (rc == RT_OK && rgb8_out) rc = rt_readback_rgb8(c, rgb8_out); // final image of everything accumulated
    else if (rc == RT_OK) { cudaError_t ee = cudaStreamSynchronize(c->stream); if (ee != cudaSuccess) rc = cuda_fail(ee, "rt_render_progressive"); }
    if (rc == RT_OK && cb && !stop && rgb8_out) cb(user, pass, done, rgb8_out);   // the final image also goes through the callback
    cleanup();
    return rc;
}

// ---------------------------------------------------------------- device-side tile queue (multi-GPU)
int rt_queue_dev_create(rt_ctx *c, void **cursors)
{
    if (!c || !cursors) return fail(RT_ERR_INVALID, "rt_queue_dev_create: NULL argument");
    CK(cudaSetDevice(c->device));
    *cursors = nullptr;
    CK(cudaMalloc(cursors, (size_t)RT_QUEUE_SLOTS * RT_QUEUE_STRIDE * sizeof(uint32_t)));
    cudaError_t e = cudaMemset(*cursors, 0, (size_t)RT_QUEUE_SLOTS * RT_QUEUE_STRIDE * sizeof(uint32_t));
    if (e != cudaSuccess) {
        cudaFree(*cursors); *cursors = nullptr;          // nothing is handed out half-made
        return cuda_fail(e, "rt_queue_dev_create: cudaMemset");
    }
    return RT_OK;
}

int rt_queue_dev_reset(rt_ctx *c, void *cursors, uint32_t ntiles, uint32_t nranks)
{
    if (!c || !cursors || nranks == 0 || nranks > RT_QUEUE_SLOTS) return fail(RT_ERR_INVALID, "rt_queue_dev_reset: bad argument");
    CK(enter_ctx(c));
    CK(rt_launch_queue_reset((uint32_t *)cursors, ntiles, nranks, c->stream));
    c->hs.launches += 1;
    return RT_OK;
}

int rt_queue_dev_free(rt_ctx *c, void *cursors)
{
    if (!c) return fail(RT_ERR_INVALID, "ctx is NULL");
    CK(cudaSetDevice(c->device));
    CK(cudaFree(cursors));
    return RT_OK;
}

int rt_render_queue(rt_ctx *c, const rt_tile *tiles, uint32_t ntiles, uint32_t tile_w, uint32_t tile_h,
                    const rt_render_params *p, void *cursors, uint32_t rank, uint32_t nranks)
{
    if (!c || !p || !tiles || !cursors) return fail(RT_ERR_INVALID, "rt_render_queue: NULL argument");
    if (ntiles == 0 || nranks == 0 || rank >= nranks || nranks > RT_QUEUE_SLOTS)
        return fail(RT_ERR_INVALID, "rt_render_queue: bad tile count, rank or rank count");
    if (p->spp == 0) return fail(RT_ERR_INVALID, "rt_render_queue: spp must be > 0");
    QueueMode qm{(uint32_t *)cursors, rank, nranks};
    return render_tiles_impl(c, tiles, ntiles, tile_w, tile_h, p, &qm);
}

int rt_resolve_tiles_to(rt_ctx *c, void *rgb8_frame)
{
    if (!c || !c->d_accum) return fail(RT_ERR_INVALID, "rt_resolve_tiles_to: no framebuffer");
    if (!c->feed_valid) return fail(RT_ERR_INVALID, "rt_resolve_tiles_to: no rt_render_queue launch to resolve");
    CK(enter_ctx(c));
    CK(rt_launch_resolve_tiles(c->d_accum, c->W, c->feed_tiles, c->d_feed->slot, c->feed_slots, c->feed_TP,
                               rgb8_frame ? (unsigned char *)rgb8_frame : c->d_rgb8, c->stream));
    c->hs.launches += 1;
    return RT_OK;
}

int rt_queue_tiles_rendered(rt_ctx *c, uint32_t *count)
{
    if (!c || !count || !c->feed_valid) return fail(RT_ERR_INVALID, "rt_queue_tiles_rendered: no rt_render_queue launch");
    CK(enter_ctx(c));
    CK(cudaStreamSynchronize(c->stream));
    std::vector<uint32_t> h(c->feed_slots);
    CK(cudaMemcpy(h.data(), c->d_feed->slot, (size_t)c->feed_slots * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    uint32_t n = 0;
    for (uint32_t v : h) if (v < RT_FEED_NONE) ++n;
    *count = n;
    return drain_events(c);
}

int rt_queue_trace(rt_ctx *c, uint32_t *tiles, uint32_t *when_us, uint32_t cap, uint32_t *count)
{
    if (!c || !count || !c->feed_valid) return fail(RT_ERR_INVALID, "rt_queue_trace: no rt_render_queue launch");
    CK(enter_ctx(c));
    CK(cudaStreamSynchronize(c->stream));
    std::vector<uint32_t> h(3u * (size_t)c->feed_slots);
    CK(cudaMemcpy(h.data(), c->d_feed->slot, h.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    uint32_t n = 0;
    for (uint32_t k = 0; k < c->feed_slots; ++k)
        if (h[k] < RT_FEED_NONE) {
            if (n < cap) { if (tiles) tiles[n] = h[k]; if (when_us) when_us[n] = h[2u * c->feed_slots + k]; }
            ++n;
        }
    *count = n;
    return drain_events(c);
}

int rt_copy_rgb8(rt_ctx *c, uint8_t *dst)
{
    if (!c || !dst || !c->d_rgb8) return fail(RT_ERR_INVALID, "rt_copy_rgb8: NULL argument or no framebuffer");
    CK(enter_ctx(c));
    size_t bytes = (size_t)c->W * c->H * 3;
    int r = ensure_pinned(c, bytes);
    if (r) return r;
    CK(cudaMemcpyAsync(c->h_pin, c->d_rgb8, bytes, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    std::memcpy(dst, c->h_pin, bytes);
    return drain_events(c);
}

int rt_ipc_export(rt_ctx *c, void *devptr, unsigned char handle[RT_IPC_HANDLE_BYTES])
{
    static_assert(sizeof(cudaIpcMemHandle_t) == RT_IPC_HANDLE_BYTES, "CUDA IPC handle size");
    if (!c || !devptr || !handle) return fail(RT_ERR_INVALID, "rt_ipc_export: NULL argument");
    CK(cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, devptr));
    std::memcpy(handle, &h, sizeof h);
    return RT_OK;
}

int rt_ipc_import(rt_ctx *c, const unsigned char handle[RT_IPC_HANDLE_BYTES], void **devptr)
{
    if (!c || !devptr || !handle) return fail(RT_ERR_INVALID, "rt_ipc_import: NULL argument");
    CK(cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof h);
    CK(cudaIpcOpenMemHandle(devptr, h, cudaIpcMemLazyEnablePeerAccess));
    return RT_OK;
}

int rt_ipc_close(rt_ctx *c, void *devptr)
{
    if (!c || !devptr) return fail(RT_ERR_INVALID, "rt_ipc_close: NULL argument");
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaIpcCloseMemHandle(devptr));
    return RT_OK;
}

int rt_stats_get(rt_ctx *c, rt_stats *out)
{
    if (!c || !out) return fail(RT_ERR_INVALID, "rt_stats_get: NULL argument");
    int r = rt_sync(c);
    if (r) return r;
    unsigned long long h[8];
    CK(cudaMemcpy(h, c->d_stats, sizeof h, cudaMemcpyDeviceToHost));
    *out = c->hs;
    out->samples = h[0]; out->segments = h[1]; out->sky = h[2]; out->absorbed = h[3]; out->capped = h[4];
    out->node_visits = h[5];
    out->sphere_tests = h[6];
    out->filter_misses = h[7];
    return RT_OK;
}

int rt_stats_reset(rt_ctx *c)
{
    if (!c) return fail(RT_ERR_INVALID, "ctx is NULL");
    int r = rt_sync(c);
    if (r) return r;
    CK(cudaMemsetAsync(c->d_stats, 0, 8 * sizeof(unsigned long long), c->stream));
    c->hs = rt_stats{};
    return RT_OK;
}

// src/main.cxx:87-101: "P6\n" W " " H "\n255\n" then the raw texels.
int rt_write_ppm(const char *path, uint32_t width, uint32_t height, const uint8_t *rgb8)
{
    if (!path || !rgb8) return fail(RT_ERR_INVALID, "rt_write_ppm: NULL argument");
    FILE *f = std::fopen(path, "wb");
    if (!f) return fail(RT_ERR_IO, std::string("bad file: ") + path);
    std::fprintf(f, "P6\n%u %u\n255\n", width, height);
    size_t bytes = (size_t)width * height * 3;
    size_t w = std::fwrite(rgb8, 1, bytes, f);
    int rc = std::fclose(f);
    if (w != bytes || rc != 0) return fail(RT_ERR_IO, std::string("short write: ") + path);
    return RT_OK;
}

int rt_fp32_peak_probe(rt_ctx *c, double out[3])
{
    if (!c || !out) return fail(RT_ERR_INVALID, "rt_fp32_peak_probe: NULL argument");
    CK(enter_ctx(c));
    const int blocks = c->sm_count * 8, iters = 4096;
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    CK(rt_launch_fp32_probe(c->d_scratch, blocks, 256, c->stream));   // warm-up
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        CK(cudaEventRecord(a, c->stream));
        CK(rt_launch_fp32_probe(c->d_scratch, blocks, iters, c->stream));
        CK(cudaEventRecord(b, c->stream));
        CK(cudaEventSynchronize(b));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, a, b));
        if (ms < best) best = ms;
    }
    c->hs.launches += 6;
    cudaEventDestroy(a); cudaEventDestroy(b);
    double flop = (double)blocks * 256.0 * iters * 64.0 * 2.0;
    out[0] = flop / (best * 1e-3) * 1e-12;
    out[1] = best;
    out[2] = c->sm_count;
    return RT_OK;
}

int rt_tf32_peak_probe(rt_ctx *c, double out[3])
{
    if (!c || !out) return fail(RT_ERR_INVALID, "rt_tf32_peak_probe: NULL argument");
    CK(enter_ctx(c));
    const int blocks = c->sm_count, iters = 4096;
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    CK(rt_launch_tf32_probe(blocks, 16, c->stream));   // warm-up
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        CK(cudaEventRecord(a, c->stream));
        CK(rt_launch_tf32_probe(blocks, iters, c->stream));
        CK(cudaEventRecord(b, c->stream));
        CK(cudaEventSynchronize(b));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, a, b));
        if (ms < best) best = ms;
    }
    c->hs.launches += 6;
    cudaEventDestroy(a); cudaEventDestroy(b);
    double flop = (double)blocks * iters * 4.0 * (128.0 * 256.0 * 8.0) * 2.0;
    out[0] = flop / (best * 1e-3) * 1e-12;
    out[1] = best;
    out[2] = c->sm_count;
    return RT_OK;
}

int rt_filter_info(rt_ctx *c, uint32_t out[4])
{
    if (!c || !out) return fail(RT_ERR_INVALID, "rt_filter_info: NULL argument");
    // which candidate filter the throughput build runs for the uploaded scene, and the GEMM shape per 128 rays
    // (default dispatch: above RT_AUTO_LBVH_SPHERES the throughput build takes the LBVH unless RT_FLAG_ACCEL_BRUTE is set)
    const bool tensor = c->d_tcb && c->n >= RT_TC_MIN_SPHERES && c->n <= RT_AUTO_LBVH_SPHERES;
    out[0] = tensor ? 1u : 0u;
    out[1] = tensor ? c->tc_ns : c->nfilt;      // sphere columns per ray (padded)
    out[2] = tensor ? 32u : 0u;                 // K (tf32 slots)
    out[3] = tensor ? (uint32_t)(RT_TC_WGS * 128) : (uint32_t)(RT_THREADS * RT_RPT);   // rays resident per CTA
    return RT_OK;
}

int rt_debug_sample_ball(rt_ctx *c, uint64_t n, uint64_t seed, int lens, float *out_xyz)
{
    if (!c || (n && !out_xyz)) return fail(RT_ERR_INVALID, "rt_debug_sample_ball: NULL argument");
    if (n == 0) return RT_OK;
    if (n > (1ull << 28)) return fail(RT_ERR_INVALID, "rt_debug_sample_ball: n too large");
    CK(enter_ctx(c));
    float *d = nullptr;
    CK(cudaMallocAsync((void **)&d, (size_t)n * 3u * sizeof(float), c->stream));
    cudaError_t e = rtf::launch_sample_ball(d, n, (uint32_t)seed, (uint32_t)(seed >> 32), lens, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_xyz, d, (size_t)n * 3u * sizeof(float), cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFreeAsync(d, c->stream);
    if (e != cudaSuccess) return cuda_fail(e, "rt_debug_sample_ball");
    c->hs.launches += 1;
    return RT_OK;
}

// ---------------------------------------------------------------- node-level tile queue (POSIX shm)
struct rt_queue {
    std::atomic<uint64_t> *ctr;   // RT_QUEUE_SLOTS counters, 64-byte stride
    std::string name;
    int fd;
};
static const size_t kQueueBytes = 64 * RT_QUEUE_SLOTS;

int rt_queue_open(rt_queue **out, const char *name, int create)
{
    if (!out || !name || name[0] != '/') return fail(RT_ERR_INVALID, "rt_queue_open: name must start with '/'");
    static_assert(sizeof(std::atomic<uint64_t>) == 8, "lock-free 64-bit atomic expected");
    int fd = shm_open(name, create ? (O_CREAT | O_RDWR) : O_RDWR, 0600);
    if (fd < 0) return fail(RT_ERR_SYS, std::string("shm_open(") + name + "): " + std::strerror(errno));
    if (create && ftruncate(fd, (off_t)kQueueBytes) != 0) { int e = errno; close(fd); return fail(RT_ERR_SYS, std::string("ftruncate: ") + std::strerror(e)); }
    void *p = mmap(nullptr, kQueueBytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (p == MAP_FAILED) { int e = errno; close(fd); return fail(RT_ERR_SYS, std::string("mmap: ") + std::strerror(e)); }
    rt_queue *q = new (std::nothrow) rt_queue();
    if (!q) { munmap(p, kQueueBytes); close(fd); return fail(RT_ERR_SYS, "out of host memory"); }
    q->ctr = reinterpret_cast<std::atomic<uint64_t> *>(p);
    q->name = name;
    q->fd = fd;
    if (create) for (uint32_t s = 0; s < RT_QUEUE_SLOTS; ++s) q->ctr[8 * s].store(0, std::memory_order_seq_cst);
    *out = q;
    return RT_OK;
}

int rt_queue_reset(rt_queue *q, uint32_t slot, uint64_t value)
{
    if (!q || slot >= RT_QUEUE_SLOTS) return fail(RT_ERR_INVALID, "rt_queue_reset: bad queue or slot");
    q->ctr[8 * slot].store(value, std::memory_order_seq_cst);
    return RT_OK;
}

int rt_queue_next(rt_queue *q, uint32_t slot, uint64_t n, uint64_t *previous)
{
    if (!q || !previous || slot >= RT_QUEUE_SLOTS) return fail(RT_ERR_INVALID, "rt_queue_next: bad queue, slot or NULL argument");
    *previous = q->ctr[8 * slot].fetch_add(n, std::memory_order_seq_cst);
    return RT_OK;
}

int rt_queue_close(rt_queue *q, int unlink_name)
{
    if (!q) return RT_OK;
    munmap((void *)q->ctr, kQueueBytes);
    close(q->fd);
    if (unlink_name) shm_unlink(q->name.c_str());
    delete q;
    return RT_OK;
}

}  // extern "C"
