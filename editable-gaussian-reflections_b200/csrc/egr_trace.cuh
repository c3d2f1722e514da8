// Context shared by the forward (egr_trace.cu) and backward (egr_backward.cu) kernels.
#pragma once
#include "egr_common.cuh"

namespace egr {

struct TraceCtx {
    EgrParams p;
    BvhView bvh;
    float *planes;             // pixel state planes [SP_COUNT][S][P]
    uint32_t *steps_executed;  // [P] executed steps | (spawned << 8)
    uint32_t *warp_tail;       // [num_warps][S][2] (tail chunk, batches)
    uint32_t P, S, tiles_x, num_warps;
    uint32_t hitlog_chunks;    // capacity in chunks
};

__device__ __forceinline__ float &plane(const TraceCtx &c, int pl, int step, uint32_t pix) {
    return c.planes[((size_t)pl * c.S + step) * c.P + pix];
}
__device__ __forceinline__ f3 plane3(const TraceCtx &c, int pl, int step, uint32_t pix) {
    return mk3(plane(c, pl, step, pix), plane(c, pl + 1, step, pix), plane(c, pl + 2, step, pix));
}
__device__ __forceinline__ void set_plane3(const TraceCtx &c, int pl, int step, uint32_t pix, f3 v) {
    plane(c, pl, step, pix) = v.x, plane(c, pl + 1, step, pix) = v.y, plane(c, pl + 2, step, pix) = v.z;
}


}  // namespace egr
