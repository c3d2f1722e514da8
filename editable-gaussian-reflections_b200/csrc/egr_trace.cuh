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
    uint32_t *ray_pix;         // [S][P]
    uint32_t *warp_mask;       // [S][num_warps]
    uint32_t *warp_off;        // [S][num_warps]
    uint32_t *ray_count;       // [S]
    uint32_t *work_counter;    // [S] next 32-ray group of the persistent forward launch
    char *hitlist;             // [resident warps][HITLIST_WARP_BYTES] hit lists of the rays being traced (egr_common.cuh)
    uint32_t P, S, tiles_x, num_warps;
    uint32_t tile_first, tile_count;  // step 0 traces the 8x4 tiles [tile_first, tile_first + tile_count) (a band of tile rows)
    uint32_t hitlog_chunks;    // capacity in chunks
    int force_walk;            // 0 = auto (warp-uniform walk for step 0, per-lane for bounces)
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

// Which pixel does lane `lane` of warp `warp` trace at `step`?  Step 0: the 8x4 tile of the warp.  Later steps: the
// compacted ray list written by the forward.  Returns false for lanes without a ray.
__device__ __forceinline__ bool lane_pixel(const TraceCtx &c, int step, uint32_t warp, uint32_t lane, uint32_t &pix) {
    if (step == 0) {
        const uint32_t px = (warp % c.tiles_x) * 8 + (lane & 7), py = (warp / c.tiles_x) * 4 + (lane >> 3);
        const bool valid = px < c.p.image_width && py < c.p.image_height;
        pix = valid ? py * c.p.image_width + px : 0;
        return valid;
    }
    const uint32_t r = warp * 32 + lane;
    const bool valid = r < c.ray_count[step];
    pix = valid ? c.ray_pix[(size_t)step * c.P + r] : 0;
    return valid;
}

}  // namespace egr
