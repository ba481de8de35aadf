// rt_multi.cu -- several GPUs behind one host call (include/rt_api.h, "rt_multi"): one process, one context + stream +
// host worker thread per GPU, peer access to the first device, the device-side tile queue (cursors in the first
// device's memory, pulled by the persistent render kernels over NVLink) and resolve stores straight into the first
// device's frame.  NCCL (loaded at run time: ncclCommInitAll) acknowledges the end of a frame and can do the gather
// itself (RT_MULTI_GATHER_NCCL: ncclReduce(SUM, uint8) of per-GPU frames whose foreign tiles are zero).
// The reference has no multi-GPU path; its seam is the C++ call cuda_impl(width, height, texels)
// (src/main.cxx:18,114), which host/cuda_impl_shim.cpp maps onto rt_multi_render_frame when asked for several devices.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include <dlfcn.h>
#include <cuda_runtime.h>

#include "../../include/rt_api.h"

// the handful of NCCL entry points used, resolved with dlsym (no link-time dependency: a single-GPU user never loads NCCL)
typedef struct ncclComm *ncclComm_t;
enum { kNcclSuccess = 0, kNcclInt8 = 0, kNcclUint8 = 1, kNcclInt32 = 2, kNcclSum = 0 };
struct NcclApi {
    void *lib = nullptr;
    int (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Reduce)(const void *, void *, size_t, int, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    bool load()
    {
        if (lib) return true;
        lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) return false;
        *(void **)&CommInitAll = dlsym(lib, "ncclCommInitAll");
        *(void **)&CommDestroy = dlsym(lib, "ncclCommDestroy");
        *(void **)&AllReduce = dlsym(lib, "ncclAllReduce");
        *(void **)&Reduce = dlsym(lib, "ncclReduce");
        *(void **)&GroupStart = dlsym(lib, "ncclGroupStart");
        *(void **)&GroupEnd = dlsym(lib, "ncclGroupEnd");
        *(void **)&GetErrorString = dlsym(lib, "ncclGetErrorString");
        return CommInitAll && CommDestroy && AllReduce && Reduce && GroupStart && GroupEnd;
    }
};

extern "C" int rt_internal_set_error(int code, const char *msg);   // rt_api.cu: sets the calling thread's rt_last_error

struct rt_multi {
    int ndev = 0;
    std::vector<int> dev;
    std::vector<rt_ctx *> ctx;
    void *cursors = nullptr;                 // range cursors of the tile queue, on dev[0]
    NcclApi nccl;
    std::vector<ncclComm_t> comms;
    std::vector<int *> d_ack;                // one int per device: payload of the end-of-frame all-reduce
    bool nccl_ok = false;
    // current frame geometry
    uint32_t W = 0, H = 0, tw = 0, th = 0;
    std::vector<rt_tile> tiles;              // interleaved: range r holds every ndev-th tile of the row-major grid
    bool have_scene = false;
    // last frame
    std::vector<uint32_t> last_tiles;
    double last_ms = 0.0;
    int last_nccl = 0;
};

namespace {

int mfail(int code, const std::string &msg) { return rt_internal_set_error(code, msg.c_str()); }

void make_tiles(rt_multi *m)
{
    std::vector<rt_tile> grid;
    for (uint32_t y = 0; y < m->H; y += m->th)
        for (uint32_t x = 0; x < m->W; x += m->tw)
            grid.push_back(rt_tile{x, y, (m->W - x < m->tw) ? m->W - x : m->tw, (m->H - y < m->th) ? m->H - y : m->th});
    m->tiles.clear();
    for (int r = 0; r < m->ndev; ++r)
        for (size_t i = (size_t)r; i < grid.size(); i += (size_t)m->ndev) m->tiles.push_back(grid[i]);
}

// run fn(i) on one host thread per device; the first failing status wins, its message is re-raised on the caller's thread
template <typename F>
int for_each_device(rt_multi *m, F fn)
{
    std::vector<int> rc(m->ndev, RT_OK);
    std::vector<std::string> err(m->ndev);
    std::vector<std::thread> th;
    for (int i = 0; i < m->ndev; ++i)
        th.emplace_back([&, i] {
            rc[i] = fn(i);
            if (rc[i] != RT_OK) err[i] = rt_last_error();
        });
    for (auto &t : th) t.join();
    for (int i = 0; i < m->ndev; ++i)
        if (rc[i] != RT_OK) return mfail(rc[i], "device " + std::to_string(m->dev[i]) + ": " + err[i]);
    return RT_OK;
}

int render_impl(rt_multi *m, const rt_render_params *p, uint32_t gather_flags, uint8_t *rgb8_out)
{
    if (!m->have_scene) return mfail(RT_ERR_INVALID, "rt_multi: no frame set up (call rt_multi_render_frame first)");
    const bool nccl_gather = (gather_flags & RT_MULTI_GATHER_NCCL) != 0;
    if (nccl_gather && !m->nccl_ok) return mfail(RT_ERR_INVALID, "rt_multi: RT_MULTI_GATHER_NCCL asked for, but NCCL is not available");
    const auto t0 = std::chrono::steady_clock::now();
    // rewind the range cursors (device 0) before any kernel pulls
    int r = rt_queue_dev_reset(m->ctx[0], m->cursors, (uint32_t)m->tiles.size(), (uint32_t)m->ndev);
    if (r == RT_OK) r = rt_sync(m->ctx[0]);
    if (r != RT_OK) return r;
    void *frame0 = nullptr; size_t frame_bytes = 0;
    r = rt_device_rgb8(m->ctx[0], &frame0, &frame_bytes);
    if (r != RT_OK) return r;
    rt_render_params q = *p;
    q.flags &= ~(uint32_t)RT_FLAG_OVERLAP;
    r = for_each_device(m, [&](int i) -> int {
        rt_ctx *c = m->ctx[i];
        int rc = rt_clear(c);
        if (rc) return rc;
        rc = rt_render_queue(c, m->tiles.data(), (uint32_t)m->tiles.size(), m->tw, m->th, &q, m->cursors, (uint32_t)i, (uint32_t)m->ndev);
        if (rc) return rc;
        void *stream = nullptr;
        rt_get_stream(c, &stream);
        if (nccl_gather) {
            // every GPU resolves its tiles into its OWN zeroed frame; one reduce(SUM) over NVLink lands the union on device 0
            void *mine = nullptr; size_t nb = 0;
            rc = rt_device_rgb8(c, &mine, &nb);
            if (rc) return rc;
            if (cudaSetDevice(m->dev[i]) != cudaSuccess || cudaMemsetAsync(mine, 0, nb, (cudaStream_t)stream) != cudaSuccess)
                return mfail(RT_ERR_CUDA, "rt_multi: clearing the local frame failed");
            rc = rt_resolve_tiles_to(c, nullptr);
            if (rc) return rc;
            int e = m->nccl.Reduce(mine, mine, nb, kNcclUint8, kNcclSum, 0, m->comms[i], (cudaStream_t)stream);
            if (e != kNcclSuccess) return mfail(RT_ERR_CUDA, std::string("ncclReduce: ") + (m->nccl.GetErrorString ? m->nccl.GetErrorString(e) : "failed"));
        } else {
            // the gather IS the resolve: texels of this GPU's tiles go straight into device 0's frame over NVLink
            rc = rt_resolve_tiles_to(c, frame0);
            if (rc) return rc;
            if (m->nccl_ok) {                      // end-of-frame acknowledgement (one int, all-reduce)
                int e = m->nccl.AllReduce(m->d_ack[i], m->d_ack[i], 1, kNcclInt32, kNcclSum, m->comms[i], (cudaStream_t)stream);
                if (e != kNcclSuccess) return mfail(RT_ERR_CUDA, std::string("ncclAllReduce: ") + (m->nccl.GetErrorString ? m->nccl.GetErrorString(e) : "failed"));
            }
        }
        return rt_sync(c);
    });
    if (r != RT_OK) return r;
    m->last_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    m->last_nccl = m->nccl_ok ? 1 : 0;
    for (int i = 0; i < m->ndev; ++i) {
        uint32_t n = 0;
        r = rt_queue_tiles_rendered(m->ctx[i], &n);
        if (r != RT_OK) return r;
        m->last_tiles[i] = n;
    }
    if (rgb8_out) return rt_copy_rgb8(m->ctx[0], rgb8_out);
    return RT_OK;
}

}  // namespace

extern "C" {

int rt_multi_create(rt_multi **out, const int *devices, int ndev)
{
    if (!out || !devices || ndev < 1 || ndev > (int)RT_QUEUE_SLOTS) return mfail(RT_ERR_INVALID, "rt_multi_create: bad argument");
    *out = nullptr;
    rt_multi *m = new (std::nothrow) rt_multi();
    if (!m) return mfail(RT_ERR_SYS, "out of host memory");
    m->ndev = ndev;
    m->dev.assign(devices, devices + ndev);
    m->ctx.assign(ndev, nullptr);
    m->last_tiles.assign(ndev, 0);
    for (int i = 0; i < ndev; ++i) {
        int r = rt_create(&m->ctx[i], devices[i]);
        if (r != RT_OK) { std::string e = rt_last_error(); rt_multi_destroy(m); return mfail(r, e); }
    }
    // every device must reach device 0's memory (cursors + frame): NVLink peer access
    bool distinct = true;                    // the same ordinal twice is allowed (tests on a 1-GPU box); NCCL needs distinct devices
    for (int i = 0; i < ndev; ++i)
        for (int j = 0; j < i; ++j) distinct = distinct && devices[i] != devices[j];
    for (int i = 1; i < ndev; ++i) {
        if (devices[i] == devices[0]) continue;
        int can = 0;
        cudaError_t e = cudaDeviceCanAccessPeer(&can, devices[i], devices[0]);
        if (e == cudaSuccess && can) {
            cudaSetDevice(devices[i]);
            e = cudaDeviceEnablePeerAccess(devices[0], 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
        }
        if (e != cudaSuccess || !can) {
            rt_multi_destroy(m);
            return mfail(RT_ERR_NO_DEVICE, "rt_multi_create: device " + std::to_string(devices[i]) + " has no peer access to device " +
                                               std::to_string(devices[0]) + " (the device-side tile queue lives there)");
        }
    }
    int r = rt_queue_dev_create(m->ctx[0], &m->cursors);
    if (r != RT_OK) { std::string e = rt_last_error(); rt_multi_destroy(m); return mfail(r, e); }
    // NCCL: one communicator per device in this process (ncclCommInitAll); optional -- without it the frame is still
    // complete when the host threads have synchronised their streams
    const char *no = std::getenv("RT_MULTI_NO_NCCL");
    if (!(no && no[0] == '1') && distinct && m->nccl.load()) {
        m->comms.assign(ndev, nullptr);
        if (m->nccl.CommInitAll(m->comms.data(), ndev, devices) == kNcclSuccess) {
            m->nccl_ok = true;
            m->d_ack.assign(ndev, nullptr);
            for (int i = 0; i < ndev; ++i) {
                cudaSetDevice(devices[i]);
                if (cudaMalloc((void **)&m->d_ack[i], sizeof(int)) != cudaSuccess || cudaMemset(m->d_ack[i], 0, sizeof(int)) != cudaSuccess) m->nccl_ok = false;
            }
        } else {
            m->comms.clear();
        }
    }
    *out = m;
    return RT_OK;
}

int rt_multi_destroy(rt_multi *m)
{
    if (!m) return RT_OK;
    for (size_t i = 0; i < m->comms.size(); ++i)
        if (m->comms[i]) m->nccl.CommDestroy(m->comms[i]);
    for (size_t i = 0; i < m->d_ack.size(); ++i)
        if (m->d_ack[i]) { cudaSetDevice(m->dev[i]); cudaFree(m->d_ack[i]); }
    if (m->cursors && !m->ctx.empty() && m->ctx[0]) rt_queue_dev_free(m->ctx[0], m->cursors);
    for (rt_ctx *c : m->ctx)
        if (c) rt_destroy(c);
    delete m;
    return RT_OK;
}

int rt_multi_device_count(rt_multi *m) { return m ? m->ndev : 0; }
rt_ctx *rt_multi_ctx(rt_multi *m, int i) { return (m && i >= 0 && i < m->ndev) ? m->ctx[i] : nullptr; }

int rt_multi_render_frame(rt_multi *m, const rt_sphere *spheres, uint32_t n, const rt_material *materials, uint32_t nmat,
                          const float cam[13], uint32_t width, uint32_t height, const rt_render_params *params,
                          uint32_t tile_w, uint32_t tile_h, uint32_t gather_flags, uint8_t *rgb8_out)
{
    if (!m || !cam || !params) return mfail(RT_ERR_INVALID, "rt_multi_render_frame: NULL argument");
    if (tile_w == 0 || tile_h == 0 || ((unsigned long long)tile_w * tile_h) % 256ull != 0)
        return mfail(RT_ERR_INVALID, "rt_multi_render_frame: tile_w * tile_h must be a positive multiple of 256");
    // scene, camera and framebuffer are replicated: every GPU converts and uploads its copy on its own host thread
    int r = for_each_device(m, [&](int i) -> int {
        rt_ctx *c = m->ctx[i];
        int rc = rt_scene_upload(c, spheres, n, materials, nmat);
        if (rc == RT_OK) rc = rt_camera_set(c, cam, cam + 3, cam + 6, cam[9], cam[10], cam[11], cam[12]);
        if (rc == RT_OK) rc = rt_framebuffer(c, width, height);
        return rc;
    });
    if (r != RT_OK) return r;
    if (width != m->W || height != m->H || tile_w != m->tw || tile_h != m->th) {
        m->W = width; m->H = height; m->tw = tile_w; m->th = tile_h;
        make_tiles(m);
    }
    m->have_scene = true;
    return render_impl(m, params, gather_flags, rgb8_out);
}

int rt_multi_render_again(rt_multi *m, const rt_render_params *params, uint32_t gather_flags, uint8_t *rgb8_out)
{
    if (!m || !params) return mfail(RT_ERR_INVALID, "rt_multi_render_again: NULL argument");
    return render_impl(m, params, gather_flags, rgb8_out);
}

int rt_multi_last_frame(rt_multi *m, uint32_t *tiles_per_device, int *nccl_used, double *frame_ms)
{
    if (!m) return mfail(RT_ERR_INVALID, "rt_multi_last_frame: NULL argument");
    if (tiles_per_device) std::memcpy(tiles_per_device, m->last_tiles.data(), sizeof(uint32_t) * (size_t)m->ndev);
    if (nccl_used) *nccl_used = m->last_nccl;
    if (frame_ms) *frame_ms = m->last_ms;
    return RT_OK;
}

}  // extern "C"
