// Throughput build of the hot path: FMA contraction on, explicit FMA chains in the sphere loop.
// Namespace rtf.
#define RT_EXACT 0
#include "rt_device.cuh"
#include "rt_launch.inc"
