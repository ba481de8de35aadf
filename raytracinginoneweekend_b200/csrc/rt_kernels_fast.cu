// Throughput kernels of the hot path (namespace rtf): tensor-core / FMA-chain candidate filters, the IEEE reference test
// on the candidates (explicit round-to-nearest intrinsics, so the closest hit is bit-identical to the oracle although
// this unit is compiled with FMA contraction on), throughput shading.
#define RT_EXACT 0
#include "rt_device.cuh"
#include "rt_launch.inc"
