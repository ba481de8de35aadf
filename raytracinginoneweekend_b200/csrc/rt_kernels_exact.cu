// Parity kernels of the hot path (namespace rtx): compiled with --fmad=false (see Makefile) so every float operation is a
// separate IEEE-754 multiply/add in the reference's order; includes the tensor-core kernel with the parity shading.
#define RT_EXACT 1
#include "rt_device.cuh"
#include "rt_launch.inc"
