// rt_lbvh.cu -- GPU-built linear BVH over spheres (no reference counterpart: the
// reference only has the brute-force hit_world, src/raytracer.hxx:94-118, whose result the traversal must
// reproduce exactly: closest t in (kMIN, inf), ties to the LOWEST sphere index).
//
// Build (all on the device, one stream):
//   k_bvh_bounds     AABB of the sphere centres (float atomics on order-preserving integer keys)
//   k_bvh_morton     63-bit Morton code of the normalised centre (21 bits per axis) + sphere index
//   cub::DeviceRadixSort::SortPairs   (key = code, value = index; library call -- the sort is build-time only)
//   k_bvh_hierarchy  Karras 2012: one thread per internal node finds its range and split by longest common
//                    prefix (duplicates broken by position)
//   k_bvh_refit      one thread per leaf computes the sphere's box and walks up; the second child to arrive
//                    at a node (atomic flag) merges the two child boxes and continues
// Node layout (64 B, 4 x float4): the boxes of BOTH children live in the parent, so one node visit decides
// both subtrees:  {l.min.xyz, l.max.x} {l.max.yz, r.min.xy} {r.min.z, r.max.xyz} {left, right, -, -}
// A child reference is an internal-node index, or 0x80000000 | leaf position.  Leaf data: the sphere records
// {cx,cy,cz,r*r} and signed radii re-ordered by leaf position (contiguous for neighbouring leaves) plus the
// original sphere index (for the tie rule and the material lookup).
#include <cub/device/device_radix_sort.cuh>

#include "rt_lbvh.h"

struct rt_lbvh {
    uint32_t n = 0, cap = 0;
    float4 *nodes = nullptr;          // (n-1) * 4
    float4 *leaf_sph = nullptr;       // n, sorted order
    uint32_t *prims = nullptr;        // n, sorted order -> sphere index
    unsigned long long *keys = nullptr, *keys_alt = nullptr;
    uint32_t *vals_alt = nullptr;
    uint32_t *parent = nullptr;       // (n-1) internal + n leaf parents; bit 31 = "I am the right child"
    uint32_t *flags = nullptr;        // n-1 arrival counters
    int *bounds = nullptr;            // 6 ordered ints
    void *cub_tmp = nullptr;
    size_t cub_bytes = 0;
};

namespace {

__device__ __forceinline__ int f2ord(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

__global__ void k_bvh_bounds(const float4 *__restrict__ sph, uint32_t n, int *bounds)
{
    float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float4 s = sph[i];
        lo[0] = fminf(lo[0], s.x); lo[1] = fminf(lo[1], s.y); lo[2] = fminf(lo[2], s.z);
        hi[0] = fmaxf(hi[0], s.x); hi[1] = fmaxf(hi[1], s.y); hi[2] = fmaxf(hi[2], s.z);
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
        if ((threadIdx.x & 31u) == 0) { atomicMin(bounds + k, f2ord(lo[k])); atomicMax(bounds + 3 + k, f2ord(hi[k])); }
    }
}

__device__ __forceinline__ unsigned long long spread21(unsigned long long x)
{
    x &= 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void k_bvh_morton(const float4 *__restrict__ sph, uint32_t n, const int *__restrict__ bounds,
                             unsigned long long *keys, uint32_t *vals)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 s = sph[i];
    float c[3] = {s.x, s.y, s.z};
    unsigned long long code = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float lo = ord2f(bounds[k]), hi = ord2f(bounds[3 + k]);
        float ext = fmaxf(hi - lo, 1e-30f);
        float q = (c[k] - lo) / ext;
        q = fminf(fmaxf(q, 0.f), 1.f);
        unsigned long long v = (unsigned long long)fminf(q * 2097152.f, 2097151.f);
        code |= spread21(v) << (2 - k);
    }
    keys[i] = code;
    vals[i] = i;
}

__device__ __forceinline__ int delta(const unsigned long long *__restrict__ keys, int n, int i, int j)
{
    if (j < 0 || j >= n) return -1;
    unsigned long long a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz(i ^ j);
    return __clzll((long long)(a ^ b));
}

__global__ void k_bvh_hierarchy(const unsigned long long *__restrict__ keys, int n, float4 *nodes, uint32_t *parent)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + min(d, 0);
    uint32_t left = (min(i, j) == gamma) ? (0x80000000u | (uint32_t)gamma) : (uint32_t)gamma;
    uint32_t right = (max(i, j) == gamma + 1) ? (0x80000000u | (uint32_t)(gamma + 1)) : (uint32_t)(gamma + 1);
    nodes[4 * (size_t)i + 3] = make_float4(__uint_as_float(left), __uint_as_float(right), 0.f, 0.f);
    // parent table: internal nodes first (n-1 entries), then leaves (n entries); bit 31 marks a right child
    if (left & 0x80000000u) parent[(n - 1) + gamma] = (uint32_t)i; else parent[gamma] = (uint32_t)i;
    if (right & 0x80000000u) parent[(n - 1) + gamma + 1] = 0x80000000u | (uint32_t)i; else parent[gamma + 1] = 0x80000000u | (uint32_t)i;
    if (i == 0) parent[0] = 0xffffffffu;   // root (overwritten above only if node 0 were somebody's child: it is not)
}

__device__ __forceinline__ void store_child_box(float4 *node, bool right, const float (&lo)[3], const float (&hi)[3])
{
    float *f = reinterpret_cast<float *>(node);
    if (!right) { f[0] = lo[0]; f[1] = lo[1]; f[2] = lo[2]; f[3] = hi[0]; f[4] = hi[1]; f[5] = hi[2]; }
    else { f[6] = lo[0]; f[7] = lo[1]; f[8] = lo[2]; f[9] = hi[0]; f[10] = hi[1]; f[11] = hi[2]; }
}

__global__ void k_bvh_refit(const float4 *__restrict__ sph, const float *__restrict__ rad, const uint32_t *__restrict__ prims,
                            uint32_t n, float4 *nodes, float4 *leaf_sph, const uint32_t *__restrict__ parent, uint32_t *flags)
{
    uint32_t pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= n) return;
    uint32_t j = prims[pos];
    float4 s = sph[j];
    leaf_sph[pos] = s;
    float r = fabsf(rad[j]);
    float lo[3] = {s.x - r, s.y - r, s.z - r}, hi[3] = {s.x + r, s.y + r, s.z + r};
    uint32_t p = parent[(n - 1) + pos];
    for (;;) {
        uint32_t node = p & 0x7fffffffu;
        store_child_box(nodes + 4 * (size_t)node, (p & 0x80000000u) != 0, lo, hi);
        __threadfence();
        if (atomicAdd(flags + node, 1u) == 0u) return;         // the sibling subtree is not finished yet
        const volatile float *f = reinterpret_cast<const volatile float *>(nodes + 4 * (size_t)node);
#pragma unroll
        for (int k = 0; k < 3; ++k) { lo[k] = fminf(f[k], f[6 + k]); hi[k] = fmaxf(f[3 + k], f[9 + k]); }
        if (node == 0) return;                                  // root done
        p = parent[node];
    }
}

#if RT_BVH_CENTER
// Child boxes (lo, hi) -> (centre, half extent), in place, plus the L1 half extent hx + hy + hz in the two spare words.
// The half extents are rounded outwards so that [c - h, c + h] contains [lo, hi] in float arithmetic.
__global__ void k_bvh_center_form(float4 *nodes, uint32_t ninner, const float4 *__restrict__ leaf_sph,
                                  const float *__restrict__ rad, const uint32_t *__restrict__ prims)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ninner) return;
    float *f = reinterpret_cast<float *>(nodes + 4 * (size_t)i);
    float v[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) v[k] = f[k];
#pragma unroll
    for (int ch = 0; ch < 2; ++ch) {
        float hs = 0.f;
#if RT_BVH_LEAFINNODE
        const uint32_t ref = __float_as_uint(f[12 + ch]);
        if (ref & 0x80000000u) {
            // a leaf: the sphere itself.  Centre and |r| exact -- the box [c - |r|, c + |r|] holds the sphere without any rounding,
            // and |r| * |r| reproduces the record's r*r bit for bit (the leaf test squares it)
            const uint32_t pos = ref & 0x7fffffffu;
            const float4 sp = leaf_sph[pos];
            const float r = fabsf(rad[prims[pos]]);
            f[6 * ch + 0] = sp.x; f[6 * ch + 1] = sp.y; f[6 * ch + 2] = sp.z;
            f[6 * ch + 3] = r; f[6 * ch + 4] = r; f[6 * ch + 5] = r;
            f[14 + ch] = 3.f * r * 1.000001f;
            continue;
        }
#endif
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float lo = v[6 * ch + k], hi = v[6 * ch + 3 + k];
            const float c = 0.5f * lo + 0.5f * hi;
            float h = fmaxf(hi - c, c - lo);
            h = fmaf(fabsf(h) + fabsf(c), 2.4e-7f, h);          // two ulps of the larger magnitude: covers the rounding of c and of h
            f[6 * ch + k] = c; f[6 * ch + 3 + k] = h;
            hs += h;
        }
        f[14 + ch] = hs * 1.000001f;
    }
}
#endif

template <typename T>
cudaError_t grow(T **p, size_t count)
{
    if (*p) cudaFree(*p);
    *p = nullptr;
    return cudaMalloc(p, count * sizeof(T));
}

}  // namespace

int rt_lbvh_create(rt_lbvh **out)
{
    *out = new (std::nothrow) rt_lbvh();
    return *out ? 0 : 1;
}

void rt_lbvh_destroy(rt_lbvh *b)
{
    if (!b) return;
    cudaFree(b->nodes); cudaFree(b->leaf_sph); cudaFree(b->prims); cudaFree(b->keys); cudaFree(b->keys_alt);
    cudaFree(b->vals_alt); cudaFree(b->parent); cudaFree(b->flags); cudaFree(b->bounds); cudaFree(b->cub_tmp);
    delete b;
}

cudaError_t rt_lbvh_build(rt_lbvh *b, const float4 *sph, const float *rad, uint32_t n, cudaStream_t s, float *ms, unsigned *launches)
{
    *ms = 0.f; *launches = 0;
    if (n < 2) return cudaErrorInvalidValue;
    cudaError_t e;
#define LB(call) do { e = (call); if (e != cudaSuccess) return e; } while (0)
    if (n > b->cap) {
        LB(grow(&b->nodes, (size_t)(n - 1) * 4));
        LB(grow(&b->leaf_sph, (size_t)n));
        LB(grow(&b->prims, (size_t)n));
        LB(grow(&b->keys, (size_t)n));
        LB(grow(&b->keys_alt, (size_t)n));
        LB(grow(&b->vals_alt, (size_t)n));
        LB(grow(&b->parent, (size_t)2 * n));
        LB(grow(&b->flags, (size_t)n));
        if (!b->bounds) LB(cudaMalloc(&b->bounds, 6 * sizeof(int)));
        size_t bytes = 0;
        LB(cub::DeviceRadixSort::SortPairs(nullptr, bytes, b->keys_alt, b->keys, b->vals_alt, b->prims, (int)n, 0, 63, s));
        if (bytes > b->cub_bytes) {
            if (b->cub_tmp) cudaFree(b->cub_tmp);
            b->cub_tmp = nullptr;
            LB(cudaMalloc(&b->cub_tmp, bytes));
            b->cub_bytes = bytes;
        }
        b->cap = n;
    }
    b->n = n;
    cudaEvent_t e0, e1;
    LB(cudaEventCreate(&e0));
    LB(cudaEventCreate(&e1));
    LB(cudaEventRecord(e0, s));
    // order-preserving integer encodings of +FLT_MAX (min slots) and -FLT_MAX (max slots)
    const int h_init[6] = {0x7f7fffff, 0x7f7fffff, 0x7f7fffff, (int)0x80800000u, (int)0x80800000u, (int)0x80800000u};
    LB(cudaMemcpyAsync(b->bounds, h_init, sizeof h_init, cudaMemcpyHostToDevice, s));
    LB(cudaMemsetAsync(b->flags, 0, (size_t)n * sizeof(uint32_t), s));
    const unsigned blocks = (n + 255u) / 256u;
    k_bvh_bounds<<<min(blocks, 148u * 8u), 256, 0, s>>>(sph, n, b->bounds);
    k_bvh_morton<<<blocks, 256, 0, s>>>(sph, n, b->bounds, b->keys_alt, b->vals_alt);
    size_t bytes = b->cub_bytes;
    LB(cub::DeviceRadixSort::SortPairs(b->cub_tmp, bytes, b->keys_alt, b->keys, b->vals_alt, b->prims, (int)n, 0, 63, s));
    k_bvh_hierarchy<<<blocks, 256, 0, s>>>(b->keys, (int)n, b->nodes, b->parent);
    k_bvh_refit<<<blocks, 256, 0, s>>>(sph, rad, b->prims, n, b->nodes, b->leaf_sph, b->parent, b->flags);
#if RT_BVH_CENTER
    k_bvh_center_form<<<(n - 1u + 255u) / 256u, 256, 0, s>>>(b->nodes, n - 1u, b->leaf_sph, rad, b->prims);
#endif
    LB(cudaGetLastError());
    LB(cudaEventRecord(e1, s));
    LB(cudaEventSynchronize(e1));
    LB(cudaEventElapsedTime(ms, e0, e1));
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    *launches = 4 + 8;   // four build kernels + the radix-sort passes of the library call (approximate)
#undef LB
    return cudaSuccess;
}

void rt_lbvh_fill(const rt_lbvh *b, RtDeviceScene *sc)
{
    sc->bvh_nodes = b->nodes;
    sc->bvh_prims = b->prims;
    sc->bvh_leaf_sph = b->leaf_sph;
    sc->bvh_root = 0;
}
