// Throughput kernels of the hot path (namespace rtf): tensor-core / FMA-chain candidate filters, the IEEE reference test
// on the candidates, parity and throughput shading.  Compiled with --fmad=false like the parity unit (see Makefile): the
// compiler never contracts on its own, so the closest hit is bit-identical to the oracle; every FMA here is an explicit fmaf.
#define RT_EXACT 0
#include "rt_device.cuh"
#include "rt_launch.inc"
