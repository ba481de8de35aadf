// rt_lbvh.cu -- placeholder until the LBVH lands (next milestone); building reports "not supported".
#include "rt_lbvh.h"
struct rt_lbvh { int unused; };
int rt_lbvh_create(rt_lbvh **out) { *out = new rt_lbvh(); return 0; }
void rt_lbvh_destroy(rt_lbvh *b) { delete b; }
cudaError_t rt_lbvh_build(rt_lbvh *, const float4 *, const float *, uint32_t, cudaStream_t, float *ms, unsigned *launches)
{
    *ms = 0.f; *launches = 0;
    return cudaErrorNotSupported;
}
void rt_lbvh_fill(const rt_lbvh *, RtDeviceScene *) {}
