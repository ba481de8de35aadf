// Parity build of the hot path: compiled with --fmad=false (see Makefile) so every float operation
// is a separate IEEE-754 multiply/add in the reference's order.  Namespace rtx.
#define RT_EXACT 1
#include "rt_device.cuh"
#include "rt_launch.inc"
