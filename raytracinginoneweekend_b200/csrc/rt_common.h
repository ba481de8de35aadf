// rt_common.h -- launchers of the mode-independent kernels (rt_common.cu) and the LBVH (rt_lbvh.cu).
#pragma once
#include "rt_types.h"

cudaError_t rt_launch_add_count(unsigned long long *accum, uint32_t W, uint4 tile0, const uint4 *tiles, uint32_t nt,
                                uint32_t TP, unsigned long long spp, cudaStream_t s);
cudaError_t rt_launch_resolve(const unsigned long long *accum, unsigned long long npix, unsigned char *rgb8, float *meanf,
                              cudaStream_t s);
// device-side tile queue (multi-GPU): cursor reset, sample-count update and resolve of the tiles a feed acquired
cudaError_t rt_launch_queue_reset(uint32_t *cursors, uint32_t ntiles, uint32_t nranks, cudaStream_t s);
cudaError_t rt_launch_add_count_feed(unsigned long long *accum, uint32_t W, const uint4 *tiles, const uint32_t *slots,
                                     uint32_t nslots, uint32_t TP, unsigned long long spp, cudaStream_t s);
cudaError_t rt_launch_resolve_tiles(const unsigned long long *accum, uint32_t W, const uint4 *tiles, const uint32_t *slots,
                                    uint32_t nslots, uint32_t TP, unsigned char *dst, cudaStream_t s);
cudaError_t rt_launch_fp32_probe(float *scratch, int blocks, int iters, cudaStream_t s);
// streams iters x 4 tcgen05.mma kind::tf32 (M128 N256 K8) per CTA: 128*256*8 MACs each
cudaError_t rt_launch_tf32_probe(int blocks, int iters, cudaStream_t s);
