/* include/rt_api.h -- C ABI of the B200-native path-tracing backend (librtiow_b200.so).
 *
 * This is the drop-in boundary for the ONE hot path of Alabuta/RaytracingInOneWeekend:
 *   camera rays -> sphere closest-hit -> material bounce loop -> accumulate / gamma / 8-bit.
 * Plain pointers and sizes only; no C++/torch types; no exception crosses this boundary and the
 * library never calls cudaDeviceReset (the reference does: src/CUDA/cuda_impl.cu:101-114).
 *
 * Reference interface each entry point replaces (paths relative to the reference repo):
 *   rt_scene_upload    <- raytracer::data{spheres,materials}            src/raytracer.hxx:19-30
 *                         primitives::sphere{center,radius,material_index} src/primitives.hxx:6-17
 *                         material::{lambert,metal,dielectric}           src/material.hxx:12-41
 *                         (hard-coded device copies in src/CUDA/cuda_impl.cu:404-417)
 *   rt_camera_set      <- raytracer::camera ctor                        src/camera.hxx:24-44
 *   rt_render          <- the pixel/sample loop + app::color            src/main.cxx:52-75,191-207
 *                         (GPU twin: cuda::data::render                 src/CUDA/cuda_impl.cu:361-380)
 *   rt_primary_hits    <- raytracer::hit_world / intersect              src/raytracer.hxx:52-118
 *   rt_readback_rgb8   <- gamma_correction + normalize_rgb_to_8bit      src/main.cxx:39-45,77-85,207-213
 *                         (GPU twin: rgb32_to_rgb8 + thrust::copy       src/CUDA/cuda_impl.cu:85-99,449-452)
 *   rt_render_frame    <- void cuda_impl(width, height, image_texels)   src/main.cxx:18,114;
 *                                                                       src/CUDA/cuda_impl.cu:384-453
 *   rt_write_ppm       <- app::save_to_file                             src/main.cxx:87-101
 *
 * Threading: one rt_ctx per GPU per process; calls on one ctx are serialised by the caller.
 * Every function returns RT_OK (0) or a negative rt_status; rt_last_error() gives the text.
 * There is NO CPU fallback: without a usable CUDA device rt_create fails.
 */
#ifndef RT_API_H
#define RT_API_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define RT_EXPORT __declspec(dllexport)
#else
#define RT_EXPORT __attribute__((visibility("default")))
#endif

typedef enum rt_status {
    RT_OK = 0,
    RT_ERR_INVALID = -1,   /* bad argument / call order */
    RT_ERR_CUDA = -2,      /* a CUDA runtime call failed */
    RT_ERR_NO_DEVICE = -3, /* no usable CUDA device: the backend refuses to run */
    RT_ERR_IO = -4,
    RT_ERR_SYS = -5
} rt_status;

enum { RT_LAMBERT = 0, RT_METAL = 1, RT_DIELECTRIC = 2 };

/* rt_render_params.flags / rt_primary_hits flags */
enum {
    RT_FLAG_EXACT = 1u << 0,      /* parity mode: the reference's arithmetic and rejection sampler; renders are bit-identical to the
                                     oracle.  48..1024 spheres, brute force: the tensor-core kernel with parity shading
                                     (rtx::k_render_tc); otherwise, or with RT_FLAG_FILTER_FMA, the straightforward parity kernels.
                                     Without this flag the CLOSEST HIT is still the oracle's bit for bit (index, t, normal); only
                                     the shading differs (unit-ball sampling by inversion, rsqrt / reciprocal normalisations) */
    RT_FLAG_ACCEL_BRUTE = 1u << 1,/* force brute-force closest hit (spheres staged in shared memory) */
    RT_FLAG_ACCEL_LBVH = 1u << 2, /* force the GPU-built LBVH */
    RT_FLAG_FILTER_FMA = 1u << 3  /* brute force: run the candidate filter on the FP32 pipe instead of the tensor cores (tcgen05 TF32x3
                                     product, the default for 48..1024 spheres); same accumulators, bit for bit */,
    RT_FLAG_FILTER_CHECK = 1u << 4,/* self-check: after the filtered closest hit, run the candidate test over ALL spheres and count
                                     the segments whose result differs in rt_stats.filter_misses (must stay 0; slow) */
    RT_FLAG_BOOK_OPTICS = 1u << 5 /* opt-in, NOT the reference's picture: textbook optics in place of the reference's quirks --
                                     Snell refraction with total internal reflection (vs src/math.hxx:300-309), camera
                                     direction measured from the eye and a disk lens in the camera basis (vs
                                     src/camera.hxx:46-57), sky blend .5*y + .5 (vs src/main.cxx:71).  Everything else
                                     (closest hit, lambert/metal sampling, Schlick, accumulate, resolve) is unchanged.
                                     rt_render / rt_render_tiles / rt_render_progressive / rt_render_frame only. */
    /* neither accel flag: parity build -- brute force up to RT_BRUTE_MAX_SPHERES, LBVH above; throughput build -- brute
       force (tensor-core filter) up to RT_AUTO_LBVH_SPHERES, LBVH above (same image: the LBVH returns the brute-force hit) */,
    RT_FLAG_OVERLAP = 1u << 6,    /* rt_render_tiles only: this launch may run BESIDE the rt_render_tiles launches queued before it
                                     (they alternate between two internal streams, so the tail of one persistent launch is
                                     filled by the next).  The caller promises disjoint tiles within a frame, as the multi-GPU
                                     tile scheduler does; every other call on the context is ordered after all of them. */
    RT_FLAG_TEST_STRAGGLERS = 1u << 7 /* test hook, LBVH render kernels only: a warp enters its straggler mode (the remaining closest-hit
                                     queries are traversed by the whole warp) the moment its work pool is dry instead of
                                     waiting until at most four lanes are left, and the odd warps of a CTA get a work list of 8 entries,
                                     so that their queries overflow into the fallback.  The image must not change by a bit. */
};
#define RT_BRUTE_MAX_SPHERES 8192u   /* parity kernels; the throughput kernel with the FMA-chain filter fits about 5 800 spheres in
                                        shared memory -- beyond that RT_FLAG_ACCEL_BRUTE fails with a clear message and the
                                        automatic choice is the LBVH (which it already is above RT_AUTO_LBVH_SPHERES) */
#define RT_AUTO_LBVH_SPHERES 768u

typedef struct rt_ctx rt_ctx;

/* primitives::sphere (src/primitives.hxx:6-17); a negative radius flips the normal (hollow glass). */
typedef struct rt_sphere { float cx, cy, cz, r; uint32_t mat; } rt_sphere;

/* material::types (src/material.hxx:12-51).  param = roughness (metal) | refraction_index (dielectric). */
typedef struct rt_material { float r, g, b, param; uint32_t kind; } rt_material;

typedef struct rt_tile { uint32_t x0, y0, w, h; } rt_tile;

typedef struct rt_render_params {
    uint32_t spp;          /* camera paths per pixel added by this call */
    uint32_t max_depth;    /* bounce cap (reference: 64 on CPU, src/raytracer.hxx:20; BASELINE: 50) */
    uint64_t seed;         /* Philox key */
    uint32_t sample_base;  /* first sample index; samples [sample_base, sample_base+spp) are traced */
    uint32_t flags;        /* RT_FLAG_* */
} rt_render_params;

typedef struct rt_stats {
    uint64_t samples;      /* camera paths traced since rt_stats_reset */
    uint64_t segments;     /* closest-hit queries (hit_world calls) */
    uint64_t sphere_tests; /* brute force: segments * n_spheres; LBVH: leaf sphere tests */
    uint64_t node_visits;  /* LBVH inner-node visits */
    uint64_t sky;          /* paths that reached the background */
    uint64_t absorbed;     /* metal scatter below the surface (returns black) */
    uint64_t capped;       /* paths cut by max_depth (returns black) */
    uint64_t launches;     /* kernels launched by this library */
    double render_ms;      /* device time of the render kernels (CUDA events), cumulative */
    double last_render_ms; /* device time of the most recent rt_render */
    double bvh_build_ms;   /* device time of the most recent LBVH build */
    uint64_t filter_misses;/* RT_FLAG_FILTER_CHECK: segments where the filtered closest hit differs from the unfiltered one */
    uint32_t last_kernel;  /* rt_kernel_id of the most recent render launch (which kernel the dispatch chose) */
    uint32_t reserved;
} rt_stats;

/* rt_stats.last_kernel */
enum rt_kernel_id {
    RT_KERNEL_NONE = 0,
    RT_KERNEL_RENDER_PARITY = 1,     /* rtx::k_render: straightforward parity kernel, sign-mask filter on the ALU pipe */
    RT_KERNEL_RENDER_FMA = 2,        /* rtf::k_render: FMA-chain candidate filter, throughput shading */
    RT_KERNEL_RENDER_SMALL = 3,      /* rtf::k_render_small: <= 47 spheres, every sphere tested */
    RT_KERNEL_RENDER_TC = 4,         /* rtf::k_render_tc: tensor-core candidate filter, throughput shading (the timed default) */
    RT_KERNEL_RENDER_TC_PARITY = 5,  /* rtx::k_render_tc: the same kernel with the parity shading (bit-identical to the oracle) */
    RT_KERNEL_RENDER_BVH_PARITY = 6, /* rtx::k_render_bvh */
    RT_KERNEL_RENDER_BVH = 7         /* rtb::k_render_bvh: throughput shading, compiled without FMA contraction */
};

RT_EXPORT const char *rt_last_error(void);
RT_EXPORT const char *rt_version(void);

/* device = CUDA ordinal (e.g. LOCAL_RANK).  Fails with RT_ERR_NO_DEVICE if CUDA is unusable. */
RT_EXPORT int rt_create(rt_ctx **out, int device);
RT_EXPORT int rt_destroy(rt_ctx *ctx);

/* Use an existing CUDA stream (e.g. torch's current stream); NULL restores the ctx's own stream. */
RT_EXPORT int rt_set_stream(rt_ctx *ctx, void *cuda_stream);
RT_EXPORT int rt_get_stream(rt_ctx *ctx, void **cuda_stream);   /* the stream the context currently queues its work on */
RT_EXPORT int rt_sync(rt_ctx *ctx);
/* Block until at most `max_in_flight` rt_render / rt_render_tiles calls are still running. */
RT_EXPORT int rt_throttle(rt_ctx *ctx, uint32_t max_in_flight);

/* Scene: converted on the host to the device layout (float4{c,r*r} + radius + material index;
 * float4{albedo,param} + kind) and uploaded; builds the LBVH when n > RT_BRUTE_MAX_SPHERES or
 * when it is later requested by flag. */
RT_EXPORT int rt_scene_upload(rt_ctx *ctx, const rt_sphere *spheres, uint32_t n,
                              const rt_material *materials, uint32_t m);

/* Camera: the constructor arguments of src/camera.hxx:24; derived vectors computed on the host with
 * the reference's float operation order.  rt_camera_get returns
 * {origin[3], lower_left_corner[3], horizontal[3], vertical[3], lens_radius}. */
RT_EXPORT int rt_camera_set(rt_ctx *ctx, const float pos[3], const float lookat[3], const float up[3],
                            float aspect, float vfov_deg, float aperture, float focus_dist);
RT_EXPORT int rt_camera_get(rt_ctx *ctx, float out13[13]);

/* Framebuffer: W x H accumulators of 4 x uint64 {R,G,B (32.32 fixed point sums), sample count}.
 * (Re)allocates if the size changes and clears. */
RT_EXPORT int rt_framebuffer(rt_ctx *ctx, uint32_t width, uint32_t height);
RT_EXPORT int rt_clear(rt_ctx *ctx);

/* Trace params->spp camera paths for every pixel of `tile` and ADD them to the accumulators.
 * Asynchronous on the ctx stream; rt_sync / any readback waits. */
RT_EXPORT int rt_render(rt_ctx *ctx, const rt_tile *tile, const rt_render_params *params);

/* Same, for a LIST of tiles in ONE persistent launch (multi-GPU: the batch of tiles a rank owns or
 * stole).  Every tile must fit in tile_w x tile_h (edge tiles may be smaller); tiles must not overlap.
 * tile_w * tile_h * spp must stay below 2^31. */
RT_EXPORT int rt_render_tiles(rt_ctx *ctx, const rt_tile *tiles, uint32_t ntiles, uint32_t tile_w,
                              uint32_t tile_h, const rt_render_params *params);

/* Parity pass (BASELINE config 2): primary rays, u = x/W, v = y/H, no jitter, no lens offset.
 * Host outputs, tile-shaped (h rows of w): sphere index (-1 = miss), t, normal xyz (0 on miss). */
RT_EXPORT int rt_primary_hits(rt_ctx *ctx, const rt_tile *tile, uint32_t flags,
                              int32_t *sphere_index, float *t, float *normal_xyz);

/* Closest hit for caller-supplied rays (n x {o[3], d[3]}), host in / host out. */
RT_EXPORT int rt_closest_hits(rt_ctx *ctx, const float *rays6, uint64_t n, uint32_t flags,
                              int32_t *sphere_index, float *t, float *normal_xyz);

/* The same closest-hit pass for callers that may not allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  It has NO counterpart in the reference.
 * No context and no allocation of any kind: the sphere records, the rays and the results are staged
 * in one file-scope device array of fixed size (csrc/rt_static.cu; about 2.9 MB that every GPU which loads the library
 * holds until exit), hence the limits below -- larger inputs are rejected with RT_ERR_INVALID and a message, use
 * rt_closest_hits for them.  spheres_xyzr = n_spheres x {cx, cy, cz, r}; rays6 = n_rays x {o[3], d[3]}; the three outputs
 * as for rt_closest_hits.  Every pointer may be host OR device memory (unified addressing).  Runs on the CURRENT device, on
 * cuda_stream (a cudaStream_t of that device; NULL = the default stream), with the parity kernel of RT_FLAG_EXACT: index,
 * t and normal are bit-identical to rt_closest_hits.  The call returns after the stream has drained unless
 * RT_STATIC_NO_SYNC is set (then pageable host outputs are not safe to read until the caller synchronises).
 * All calls on one GPU share the array: calls that overlap on the GPU (two streams or two host threads) race. */
#define RT_STATIC_MAX_SPHERES 4096u
#define RT_STATIC_MAX_RAYS 65536u
#define RT_STATIC_NO_SYNC 1u
RT_EXPORT int rt_closest_hits_static(const float *spheres_xyzr, uint32_t n_spheres, const float *rays6, uint32_t n_rays,
                                     int32_t *sphere_index, float *t, float *normal_xyz, void *cuda_stream, uint32_t flags);

/* Resolve: mean = sum / count -> powf(mean, 1/2.2f) -> (uint8)(255*c), tightly packed RGB, row-major,
 * top row first (the reference's texel layout, src/main.cxx:96-100).  dst is HOST memory, W*H*3. */
RT_EXPORT int rt_readback_rgb8(rt_ctx *ctx, uint8_t *dst);
/* Mean linear radiance, W*H*3 float, HOST memory. */
RT_EXPORT int rt_readback_f32(rt_ctx *ctx, float *dst);
/* Raw accumulators, W*H*4 uint64 {R,G,B,count}, HOST memory (bit-exact parity tests). */
RT_EXPORT int rt_readback_accum(rt_ctx *ctx, uint64_t *dst);

/* Device-side handles for host plumbing (NCCL gather of the framebuffer through torch.distributed):
 *   accum : W*H*4 uint64 accumulators;  rgb8 : W*H*3 bytes, written by rt_resolve_device. */
RT_EXPORT int rt_device_accum(rt_ctx *ctx, void **dptr, size_t *bytes);
RT_EXPORT int rt_device_rgb8(rt_ctx *ctx, void **dptr, size_t *bytes);
/* Resolve into the device rgb8 buffer without copying to the host.  Pixels with count == 0 give 0. */
RT_EXPORT int rt_resolve_device(rt_ctx *ctx);

/* The whole reference-facing call in one: upload scene + camera, render the full W x H image at
 * spp / max_depth, resolve, copy RGB8 to the HOST buffer.  This is cuda_impl() with the scene made an
 * argument; host->device and device->host copies are inside. */
RT_EXPORT int rt_render_frame(rt_ctx *ctx, const rt_sphere *spheres, uint32_t n,
                              const rt_material *materials, uint32_t m, const float camera13[13],
                              uint32_t width, uint32_t height, const rt_render_params *params,
                              uint8_t *rgb8_out);

/* Progressive rendering: trace params->spp paths per pixel over the whole framebuffer in passes of spp_per_pass
 * (samples [sample_base + k*spp_per_pass, ...), so the accumulators end bit-identical to one rt_render of params->spp --
 * the split accumulation of src/main.cxx:205-207).  After every `preview_every`-th pass the image of the samples so far
 * is resolved, copied to pinned host memory on a second stream and handed to `preview` WHILE the next pass renders;
 * the final image is written to rgb8_out (HOST, W*H*3; may be NULL) and passed to the callback too.  The callback's
 * rgb8 pointer is valid only during the call; returning non-zero stops after the pass in flight (RT_OK is returned,
 * *spp_done tells how far it got).  preview == NULL or preview_every == 0: no previews. */
typedef int (*rt_preview_fn)(void *user, uint32_t pass, uint32_t spp_done, const uint8_t *rgb8);
RT_EXPORT int rt_render_progressive(rt_ctx *ctx, const rt_render_params *params, uint32_t spp_per_pass,
                                    uint32_t preview_every, rt_preview_fn preview, void *user, uint8_t *rgb8_out,
                                    uint32_t *spp_done);

RT_EXPORT int rt_stats_get(rt_ctx *ctx, rt_stats *out);
RT_EXPORT int rt_stats_reset(rt_ctx *ctx);

/* P6 writer with the reference's exact header format (src/main.cxx:96). */
RT_EXPORT int rt_write_ppm(const char *path, uint32_t width, uint32_t height, const uint8_t *rgb8);

/* FP32 FMA-chain microbenchmark: the measured denominator of the FP32 roofline.
 * out = {TFLOP/s achieved (FMA = 2 flop), kernel ms, SM count}. */
RT_EXPORT int rt_fp32_peak_probe(rt_ctx *ctx, double out3[3]);
/* TF32 tensor-core peak of this device measured with a tcgen05.mma stream (M128 N256 K8, one CTA per SM):
 * out = {TFLOP/s, best ms, SM count}.  Denominator of the tensor roofline of the candidate filter (bench.py). */
RT_EXPORT int rt_tf32_peak_probe(rt_ctx *ctx, double out[3]);
/* Candidate filter the throughput build runs for the uploaded scene: out = {1 tensor cores | 0 FMA pipe,
 * padded sphere columns per ray, K of the filter product (tf32 slots), rays resident per CTA}. */
RT_EXPORT int rt_filter_info(rt_ctx *ctx, uint32_t out[4]);

/* Test hook: n points of the throughput build's ball sampler (lens = 0: x,y,z uniform in the unit ball, the law of
 * raytracer::random_in_unit_sphere, src/raytracer.hxx:32-43) or lens sampler (lens = 1: x,y of such a point, z = 0,
 * as used by camera::ray, src/camera.hxx:51-53), drawn exactly as the render kernels draw them.  HOST output, n*3 floats. */
RT_EXPORT int rt_debug_sample_ball(rt_ctx *ctx, uint64_t n, uint64_t seed, int lens, float *out_xyz);

/* ---- node-level tile queue (multi-GPU dynamic tile stealing, one process per GPU) -------------
 * RT_QUEUE_SLOTS lock-free counters (one cache line each) in POSIX shared memory, shared by the ranks of
 * one node: slot r is the cursor of rank r's own tile range; a rank that runs dry steals by advancing
 * another rank's cursor.  Rank 0 creates, the others open.  rt_queue_next returns the previous value
 * after adding `n` (fetch_add; n = 0 reads). */
typedef struct rt_queue rt_queue;
#define RT_QUEUE_SLOTS 64u
RT_EXPORT int rt_queue_open(rt_queue **out, const char *name, int create);
RT_EXPORT int rt_queue_reset(rt_queue *q, uint32_t slot, uint64_t value);
RT_EXPORT int rt_queue_next(rt_queue *q, uint32_t slot, uint64_t n, uint64_t *previous);
RT_EXPORT int rt_queue_close(rt_queue *q, int unlink_name);

/* ---- device-side tile queue (multi-GPU) ---------------------------------------------------------
 * The frame's tile list (interleaved so that every range sees the image's cost mix) is split into one contiguous range
 * per rank.  The cursors of all ranges live in ONE device allocation in rank 0's memory, reachable from the other GPUs
 * over NVLink (peer access in one process -- rt_multi below -- or CUDA IPC between processes).  rt_render_queue is ONE
 * persistent launch per GPU and frame: the kernel itself takes tiles from its own range (a system-scope atomicAdd on
 * the cursor) and, when that is dry, steals from the other ranges -- no host round trip, no per-chunk launch.
 * rt_resolve_tiles_to then resolves exactly the tiles this GPU rendered and stores the texels straight into rank 0's
 * frame (peer-mapped): that store IS the framebuffer gather.
 *   rank 0:  rt_queue_dev_create -> (rt_ipc_export) ; per frame rt_queue_dev_reset, then a barrier
 *   rank r:  (rt_ipc_import) ; per frame rt_render_queue + rt_resolve_tiles_to(rank 0's rgb8) ; then a barrier / ack
 * tile_w * tile_h must be a multiple of 256. */
#define RT_IPC_HANDLE_BYTES 64
RT_EXPORT int rt_queue_dev_create(rt_ctx *ctx, void **cursors_dev);
RT_EXPORT int rt_queue_dev_reset(rt_ctx *ctx, void *cursors_dev, uint32_t ntiles, uint32_t nranks);  /* on the ctx stream */
RT_EXPORT int rt_queue_dev_free(rt_ctx *ctx, void *cursors_dev);
RT_EXPORT int rt_render_queue(rt_ctx *ctx, const rt_tile *tiles, uint32_t ntiles, uint32_t tile_w, uint32_t tile_h,
                              const rt_render_params *params, void *cursors_dev, uint32_t rank, uint32_t nranks);
/* rgb8_frame_dev: device pointer to a W*H*3 frame (rank 0's, peer-mapped), or NULL for this context's own RGB8 buffer */
RT_EXPORT int rt_resolve_tiles_to(rt_ctx *ctx, void *rgb8_frame_dev);
RT_EXPORT int rt_queue_tiles_rendered(rt_ctx *ctx, uint32_t *count);   /* tiles the last rt_render_queue launch acquired */
/* the same in acquisition order: tile indices and the device's globaltimer (>> 10: about microseconds) at each acquisition;
 * at most `cap` entries are written, *count is the total (diagnosis of the stealing schedule) */
RT_EXPORT int rt_queue_trace(rt_ctx *ctx, uint32_t *tiles, uint32_t *when_us, uint32_t cap, uint32_t *count);
/* copy the context's device RGB8 buffer to the HOST as it is (no resolve): rank 0 after the peers' stores */
RT_EXPORT int rt_copy_rgb8(rt_ctx *ctx, uint8_t *dst);
/* CUDA IPC plumbing for one process per GPU: export a device allocation of this context (the cursors, the RGB8 frame
 * from rt_device_rgb8), import it in another process; handles are RT_IPC_HANDLE_BYTES opaque bytes. */
RT_EXPORT int rt_ipc_export(rt_ctx *ctx, void *devptr, unsigned char handle[RT_IPC_HANDLE_BYTES]);
RT_EXPORT int rt_ipc_import(rt_ctx *ctx, const unsigned char handle[RT_IPC_HANDLE_BYTES], void **devptr);
RT_EXPORT int rt_ipc_close(rt_ctx *ctx, void *devptr);

/* ---- several GPUs behind ONE host call (rt_create(devices, ndev); host code stays C++) --------------
 * One process, one context + stream + host worker thread per GPU, peer access to devices[0], the device-side tile
 * queue above, resolve stores into devices[0]'s frame; NCCL (ncclCommInitAll, loaded at run time) provides the
 * end-of-frame acknowledgement and, with RT_MULTI_GATHER_NCCL or when peer access is unavailable, the gather itself
 * (ncclReduce(SUM, uint8) of the per-GPU frames, whose foreign tiles are zero).
 * rt_multi_render_frame is cuda_impl(width, height, texels) on N GPUs: host scene in, host RGB8 out; the N-GPU frame
 * is bit-identical to the 1-GPU frame (integer accumulation, (pixel, sample)-keyed RNG). */
typedef struct rt_multi rt_multi;
enum { RT_MULTI_GATHER_NCCL = 1u << 0 };   /* rt_multi_render_frame gather_flags */
RT_EXPORT int rt_multi_create(rt_multi **out, const int *devices, int ndev);
RT_EXPORT int rt_multi_destroy(rt_multi *m);
RT_EXPORT int rt_multi_device_count(rt_multi *m);
RT_EXPORT rt_ctx *rt_multi_ctx(rt_multi *m, int i);          /* context of devices[i] (statistics, probes) */
RT_EXPORT int rt_multi_render_frame(rt_multi *m, const rt_sphere *spheres, uint32_t n, const rt_material *materials,
                                    uint32_t nmat, const float camera13[13], uint32_t width, uint32_t height,
                                    const rt_render_params *params, uint32_t tile_w, uint32_t tile_h,
                                    uint32_t gather_flags, uint8_t *rgb8_out);
/* the same without re-uploading the scene (rt_multi_render_frame must have run once): device-timed benchmarking */
RT_EXPORT int rt_multi_render_again(rt_multi *m, const rt_render_params *params, uint32_t gather_flags, uint8_t *rgb8_out);
/* out = {tiles rendered by device i in the last frame} for i < ndev; NCCL in use (0/1) in nccl_used */
RT_EXPORT int rt_multi_last_frame(rt_multi *m, uint32_t *tiles_per_device, int *nccl_used, double *frame_ms);

#ifdef __cplusplus
}
#endif
#endif /* RT_API_H */
