// rt_lbvh.h -- GPU-built linear BVH over spheres (Morton codes -> radix sort -> Karras hierarchy -> refit).
#pragma once
#include "rt_types.h"

struct rt_lbvh;
int rt_lbvh_create(rt_lbvh **out);
void rt_lbvh_destroy(rt_lbvh *b);
// sph = {cx,cy,cz,r*r}, rad = signed radius.  ms / launches report the build cost.
cudaError_t rt_lbvh_build(rt_lbvh *b, const float4 *sph, const float *rad, uint32_t n, cudaStream_t s,
                          float *ms, unsigned *launches);
void rt_lbvh_fill(const rt_lbvh *b, RtDeviceScene *sc);
