// rt_types.h -- device-side data layout shared by the kernel translation units and the C ABI.
//
// HBM layout (all library-owned, see DESIGN.md "Data layout"):
//   sph      float4[npad]   {cx, cy, cz, r*r}     npad = n rounded up to 32; padding has r*r = -inf
//   sph_rad  float [npad]   signed radius (negative = flipped normal, src/raytracer.hxx:70)
//   sph_meta uint32[npad]   material index | kind << 30
//   mat      float4[m]      {albedo.rgb, param}
//   accum    ulonglong4[W*H] {R, G, B sums in 32.32 fixed point, sample count}
//   rgb8     uchar[W*H*3], meanf float[W*H*3]  resolve outputs
//   LBVH: nodes (2 x float4 child boxes + children), sorted primitive ids
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define RT_THREADS 256          // threads per CTA of the persistent render kernel
#define RT_RPT 2                // camera paths resident per thread (ILP + LDS amortisation)
#define RT_CHUNK_WORDS 16       // exact build: candidate-mask words per chunk (16 x 32 = 512 spheres)
#define RT_FCHUNK_WORDS 22      // fast build: mask words per chunk (22 x 24 = 528 spheres)
#define RT_FGROUP 24            // fast build: spheres per mask word (bits of an exact float integer)
#define RT_FILTER_EPS 3.814697265625e-6   // 2^-18: relative safety margin of the conservative filter
#define RT_BATCH 256u           // work items a warp takes from the global counter at once
#define RT_KMIN .008f           // src/raytracer.hxx:98
#define RT_FIXED_ONE 4294967296.f

struct RtDeviceScene {
    const float4 *sph;
    const float *sph_rad;
    const uint32_t *sph_meta;
    const float4 *mat;
    uint32_t n, npad;
    uint32_t m;                  // materials
    // conservative candidate filter of the throughput build (see rt_device.cuh, closest_hit_brute_fast)
    const float4 *filt;          // float4[nfilt] {cx, cy, cz, S*(r^2 - |c|^2 + eps(|c|^2 + r^2))}; padding has w = -inf
    uint32_t nfilt;              // n rounded up to RT_FGROUP
    float sigma, S;              // power-of-two scales: b' = sigma*b, disc' = S*disc, S = sigma^2
    float o_limit2;              // rays with |o|^2 above this take every sphere as a candidate (overflow guard)
    // tensor-core candidate filter of the throughput build (rt_tensor.cuh); tc_b == nullptr when the scene is too large
    const float4 *tc_b;          // float4[8][tc_ns]: sphere feature matrix, K-major core-matrix layout, permuted columns
    uint32_t tc_ns;              // n rounded up to 128 (accumulator tile width)
    float tc_gx, tc_gy, tc_gz;   // centre the bilinear form is expanded about
    float tc_eps;                // relative inflation of the filter (2^-16)
    // LBVH (rt_lbvh.cu); nodes == nullptr when not built
    const float4 *bvh_nodes;     // 4 x float4 per inner node
    const uint32_t *bvh_prims;   // leaf order -> sphere index
    const float4 *bvh_leaf_sph;  // sphere records {cx,cy,cz,r*r} in leaf order
    uint32_t bvh_root;
    float bvh_pad0;              // absolute box padding (covers the rounding of o - c at the scene's scale)
};

struct RtDeviceCamera {          // src/camera.hxx:66-76 derived state
    float ox, oy, oz;
    float llx, lly, llz;
    float hx, hy, hz;
    float vx, vy, vz;
    float lens_radius;
    float bux, buy, buz;         // unit camera basis vectors u, v (src/camera.hxx:36-37): only the book-optics lens uses them
    float bvx, bvy, bvz;
};

// Device-side tile queue (multi-GPU).  The frame's tile list is split into one contiguous range per rank;
// the cursor of range r lives in RANK 0's memory (peer-mapped over NVLink: cudaDeviceEnablePeerAccess or CUDA IPC), one
// 64-byte line each.  A persistent render kernel acquires tiles itself: own range first (atomicAdd at system scope on
// its cursor), then it steals by advancing another range's cursor.  Locally the kernel keeps a FEED: the tiles this GPU
// acquired, in order (slot k = the k-th acquired tile, filled on demand by the first warp that reaches it), one 32-bit
// item counter per slot (TPS = tile pixels x spp items each, handed out in batches of RT_BATCH) and the index of the
// slot currently being handed out.  All 32-bit, no division: the render loops pay two local atomics per 256 items.
#define RT_FEED_EMPTY 0xffffffffu      // slot not acquired yet
#define RT_FEED_FETCHING 0xfffffffeu   // a warp is talking to the queue for this slot
#define RT_FEED_NONE 0xfffffffdu       // the queue was dry: no more tiles for this GPU
#define RT_QUEUE_STRIDE 16u            // cursor r is at cursors[r * RT_QUEUE_STRIDE] (uint32, one 64-byte line each)
struct RtTileFeed {
    uint32_t cur;                      // slot being handed out; zeroed before the launch
    uint32_t drained;                  // set once a fetch found every range dry
    uint32_t pad[2];
    uint32_t slot[1];                  // [nslots] RT_FEED_EMPTY before the launch; afterwards the tiles this GPU rendered
                                       // followed by ctr[nslots]: items handed out of each slot, zeroed before the launch,
                                       // and when[nslots]: globaltimer >> 10 at the acquisition of the slot (rt_queue_trace)
};

struct RtRenderArgs {
    RtDeviceScene sc;
    RtDeviceCamera cam;
    unsigned long long *accum;   // W*H*4
    uint32_t W, H;               // full image
    // Work domain of one launch: `nt` tiles, each allotted TP = tw*th pixel slots (edge tiles are
    // smaller and leave slots unused), times spp samples.  item -> (tile, sample, pixel-in-tile).
    uint4 tile0;                 // {x0,y0,w,h} when tiles == nullptr (single-tile launch)
    const uint4 *tiles;          // device tile list {x0,y0,w,h} or nullptr
    uint32_t t0;                 // first list entry of this launch
    uint32_t TP;                 // pixel slots per tile (nominal tile width * height)
    uint32_t TPS;                // TP * spp: work items per tile
    uint32_t mTP, mTPS, mW0;     // floor(2^32 / d) for d = TP, TPS, tile0.z: work-item decode without integer division
    uint32_t P;                  // pixel slots of this launch = nt * TP
    uint32_t spp, depth, sample_base;
    uint32_t key0, key1;
    uint32_t total_items;        // P * spp  (< 2^31)
    uint32_t *work_counter;      // zeroed before launch
    unsigned long long *stats;   // [0] samples [1] segments [2] sky [3] absorbed [4] capped [5] node_visits [6] sphere tests [7] filter misses
    uint32_t check;              // RT_FLAG_FILTER_CHECK
    uint32_t book;               // RT_FLAG_BOOK_OPTICS
    uint32_t smem_mats;          // tensor kernel: radii, material indices and materials are staged in shared memory (they fit)
    // device-side tile queue (feed == nullptr: the work domain is the tile list above)
    RtTileFeed *feed;            // local feed of this launch
    uint32_t feed_slots;
    uint32_t *q_cursors;         // RT_QUEUE_STRIDE-strided range cursors in rank 0's memory (system-scope atomics)
    uint32_t q_rank, q_nranks, q_ntiles;
    // straggler mode of k_render_bvh (rt_bvh_render.cuh); at the END of the struct: the other kernels' parameter offsets stay put
    uint32_t coop_lanes;         // enter once the pool is dry and at most this many lanes still traverse (0 = never) ...
    uint32_t coop_after;         // ... for this many iterations of the traversal loop
    uint32_t coop_cap_odd;       // work-list entries of the odd warps of a CTA (RT_FLAG_TEST_STRAGGLERS: 8; else the full list)
};

struct RtHitArgs {
    RtDeviceScene sc;
    RtDeviceCamera cam;
    uint32_t W, H, x0, y0, w, h; // primary pass
    const float *rays;           // closest_hits pass: n x 6 (nullptr for the primary pass)
    unsigned long long nrays;
    int32_t *idx; float *t; float *normal;
    unsigned long long *stats;
};

// launchers, one set per arithmetic mode (exact = --fmad=false TU, fast = FMA TU)
namespace rtx {
cudaError_t launch_render_tc(const RtRenderArgs &a, int grid, size_t smem, cudaStream_t s, bool queue);
cudaError_t launch_render(const RtRenderArgs &a, int grid, size_t smem, cudaStream_t s, bool lbvh, bool queue);
cudaError_t launch_hits(const RtHitArgs &a, cudaStream_t s, bool lbvh);
cudaError_t render_occupancy(size_t smem, bool lbvh, int *ctas_per_sm);
}
namespace rtf {
cudaError_t launch_render(const RtRenderArgs &a, int grid, size_t smem, cudaStream_t s, bool lbvh, bool queue);
cudaError_t launch_hits(const RtHitArgs &a, cudaStream_t s, bool lbvh);
cudaError_t render_occupancy(size_t smem, bool lbvh, int *ctas_per_sm);
}

namespace rtb {   // rt_kernels_bvh.cu: the throughput LBVH render kernel, compiled without FMA contraction (lbvh must be true)
cudaError_t launch_render(const RtRenderArgs &a, int grid, size_t smem, cudaStream_t s, bool lbvh, bool queue);
cudaError_t render_occupancy(size_t smem, bool lbvh, int *ctas_per_sm);
}

namespace rtf {
cudaError_t launch_render_tc(const RtRenderArgs &a, int grid, size_t smem, cudaStream_t s, bool queue);
cudaError_t launch_hits_tc(const RtHitArgs &a, int grid, size_t smem, cudaStream_t s);
cudaError_t launch_render_small(const RtRenderArgs &a, int grid, cudaStream_t s, bool queue);
cudaError_t small_occupancy(int *ctas_per_sm);
cudaError_t launch_sample_ball(float *out, unsigned long long n, uint32_t key0, uint32_t key1, int lens, cudaStream_t s);
}
// tensor-filter kernel shape (rt_tensor.cuh): RT_TC_WGS ray warpgroups x (32 + RT_TC_NBUF * RT_TC_CHUNK) TMEM columns <= 512
#ifndef RT_SMALL_SCENE
#define RT_SMALL_SCENE 47u    // throughput build: at or below this many spheres the closest hit tests every sphere directly (k_render_small;
                              // measured against the FMA-filter kernel and the tensor kernel, tools/midsize.sh: 37 spheres 3.63 vs 3.08
                              // vs 3.31 Gsamples/s, 50 spheres 4.23 vs 4.42 (tensor))
#endif
#ifndef RT_BVH_SMEM_DEPTH
#define RT_BVH_SMEM_DEPTH 40       // levels of the LBVH traversal stack kept in shared memory (k_render_bvh; rt_common.cu sizes it); deeper levels
                                   // go to a local-memory overflow, so smaller values only change speed (and test the overflow)
#endif
#ifndef RT_BVH_COOP
#define RT_BVH_COOP 4              // straggler mode (0 = off): once the warp's work pool is dry and at most this many lanes are still
                                   // traversing, and have been for RT_BVH_COOP_AFTER loop iterations, the WHOLE warp traverses their rays.
                                   // Entering as soon as the pool is dry (32 lanes / 1 iteration) serialises the many short queries of
                                   // a frame's end.  Msamples/s, 1M spheres at 32 spp | at 256 spp | 488 spheres through the LBVH
                                   // (profiles/r02_coop_ab11.log):  32/1: 765 | 930 | 1181   8/16: 800 | 924 | 1282   4/64: 804 | 926 | 1286
                                   // 2/256: 819 | 926 | 1280   off: 846 | 885 | 1345 (the short frames pay ~4 % for the mode)
#endif
#ifndef RT_BVH_COOP_AFTER
#define RT_BVH_COOP_AFTER 64
#endif
#ifndef RT_BVH_COOP_CAP
#define RT_BVH_COOP_CAP (RT_BVH_SMEM_DEPTH * 32u)   // entries of the warp's work list (test builds shrink it to exercise the overflow path)
#endif
#ifndef RT_BVH_CENTER
#define RT_BVH_CENTER 1     // 1: LBVH nodes hold child boxes as centre + half extent (+ L1 half extent): the padded slab test needs 34 instead of 42 FP32 instructions
#endif
#ifndef RT_BVH_LEAFINNODE
#define RT_BVH_LEAFINNODE 1 // 1 (needs RT_BVH_CENTER): the child slot of a leaf holds the sphere (centre, |r|) instead of its box; the leaf test reads no memory
#endif
#ifndef RT_TC_WGS
#define RT_TC_WGS 5
#endif
#ifndef RT_TC_CHUNK
#define RT_TC_CHUNK 64
#endif
#ifndef RT_TC_NBUF
#define RT_TC_NBUF 1
#endif
#ifndef RT_TC_REGROUP
#define RT_TC_REGROUP 0            // 1: material-regrouping prototype (north_star (3)): sort the 128 paths of a warpgroup by shading kind
#endif
#ifndef RT_TC_UNROLL1
#define RT_TC_UNROLL1 0            // 1: do not unroll the ray warps' tile loop (measurement option)
#endif
#ifndef RT_TC_RAY_HINT
#define RT_TC_RAY_HINT 0           // > 0: suspend-time hint (ns) of the ray warps' try_wait on an accumulator tile; 0 = bare spin
#endif
#define RT_TC_EPS 1.52587890625e-5   // 2^-16: relative inflation of the tensor filter (measured error < 2^-20.4)
#ifndef RT_TC_MIN_SPHERES
#define RT_TC_MIN_SPHERES 48u
#endif
#define RT_TC_MAX_SPHERES 1024u   // tensor-filter path: B matrix (128 B per sphere) + records + masks fit in shared memory
size_t rt_render_tc_smem_bytes(uint32_t npad, uint32_t tc_ns, uint32_t m, bool *mats_in_smem);
#define RT_TC_SMEM_LIMIT 232448u    // 227 KB of dynamic shared memory per CTA on sm_100
size_t rt_render_smem_bytes(uint32_t npad, uint32_t nfilt, bool exact, bool lbvh);
size_t rt_hits_smem_bytes(uint32_t npad, uint32_t nfilt, bool exact, bool lbvh);
