// Throughput LBVH render kernel (namespace rtb): the throughput shading of rtf (unit ball by inversion, rsqrt / reciprocal
// normalisations), compiled WITHOUT FMA contraction (--fmad=false, see Makefile).
// Why a unit of its own: k_render_bvh shades in two places -- in its main loop and in the straggler loop that finishes a warp's
// last paths cooperatively (rt_bvh_render.cuh) -- and with contraction on the compiler fused the two copies of the same source
// differently: radiance sums that differed in the last bit depending on WHEN a warp's work pool ran dry (caught by
// test_overlapped_chunks_equal_one_launch).  Without contraction every float operation is the one that is written, in both
// copies and in both instantiations, so an LBVH frame of the throughput build is reproducible bit for bit -- on one GPU, on N
// GPUs, through the tile queue or not.  The box tests keep their FMAs: they are written as fmaf.
#define RT_EXACT 0
#define RT_NS_OVERRIDE rtb
#define RT_BVH_ONLY 1
#include "rt_device.cuh"
#include "rt_launch.inc"
