// Headless stand-in for the reference's denoise() (csrc/optix/denoiser_wrapper.h:76-104: OptiX AI denoiser, HDR
// model, guided by the normal layer, output_final -> output_denoised).  The OptiX network is a proprietary model
// inside the driver and is unavailable here; this is NOT that network.  It is a classical edge-avoiding a-trous
// wavelet filter (Dammertz et al. 2010) over the same inputs: three passes of a 5x5 B3-spline kernel with hole sizes
// 1, 2, 4 (a 29x29 footprint), every tap weighted by edge-stopping functions of the colour difference, of the
// primary-step normal (output_normal[0], the guide layer the reference hands to OptiX) and of the relative
// primary-step depth (output_depth[0]).  Constants: include/egr_b200.h (EGR_DENOISE_*); CPU restatement:
// oracle/denoise_oracle.py.
//
// Intermediate passes ping-pong through six planes of the per-pixel state block that are dead between two raytrace()
// calls (the "remaining" planes of step 0, which only the backward reads), so the library still allocates nothing.
// HBM-bound image filter: algorithmic bytes per pass = 25 taps x (12 B colour + 12 B normal + 4 B depth) from L1/L2
// (unique DRAM traffic: 28 B read + 12 B written per pixel).
#include "egr_trace.cuh"

namespace egr {

struct DenoiseArgs {
    const float *in_aos;   // [P,3] or null
    const float *in_soa[3];
    float *out_aos;        // [P,3] or null
    float *out_soa[3];
    const float *normal;   // [P,3] primary-step normal
    const float *depth;    // [P]   primary-step depth
    int W, H, hole;
    float inv_sigma_c2;    // 1 / sigma_c^2 of this pass
};

__device__ __forceinline__ f3 dn_load(const DenoiseArgs &a, int q) {
    return a.in_aos ? ld3(a.in_aos, q) : mk3(a.in_soa[0][q], a.in_soa[1][q], a.in_soa[2][q]);
}

__global__ void __launch_bounds__(256) k_denoise_atrous(const DenoiseArgs a) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= a.W || y >= a.H) return;
    const int p = y * a.W + x;
    const f3 cp = dn_load(a, p), np_ = ld3(a.normal, p);
    const float dp = a.depth[p];
    const float inv_sn2 = 1.0f / (EGR_DENOISE_SIGMA_NORMAL * EGR_DENOISE_SIGMA_NORMAL);
    const float inv_sd2 = 1.0f / (EGR_DENOISE_SIGMA_DEPTH * EGR_DENOISE_SIGMA_DEPTH);
    const float kern[5] = {1.0f / 16, 1.0f / 4, 3.0f / 8, 1.0f / 4, 1.0f / 16};  // B3 spline
    f3 sum = mk3(0, 0, 0);
    float wsum = 0.0f;
#pragma unroll
    for (int j = -2; j <= 2; ++j)
#pragma unroll
        for (int i = -2; i <= 2; ++i) {
            const int xx = x + i * a.hole, yy = y + j * a.hole;
            if (xx < 0 || xx >= a.W || yy < 0 || yy >= a.H) continue;
            const int q = yy * a.W + xx;
            const f3 cq = dn_load(a, q), nq = ld3(a.normal, q);
            const float dq = a.depth[q];
            const f3 dc = cq - cp, dn = nq - np_;
            const float dd = (dq - dp) / (fabsf(dp) + EGR_DENOISE_DEPTH_EPS);
            const float e = dotp(dc, dc) * a.inv_sigma_c2 + dotp(dn, dn) * inv_sn2 + dd * dd * inv_sd2;
            const float w = kern[i + 2] * kern[j + 2] * expf(-e);
            sum = sum + cq * w;
            wsum += w;
        }
    const f3 r = sum * (1.0f / wsum);  // the centre tap has weight (3/8)^2: wsum > 0
    if (a.out_aos) st3(a.out_aos, p, r);
    else a.out_soa[0][p] = r.x, a.out_soa[1][p] = r.y, a.out_soa[2][p] = r.z;
}

}  // namespace egr

using namespace egr;

extern "C" int egr_denoise(const EgrParams *p, void *stream) {
    if (!p) {
        set_error("params is null");
        return EGR_EINVAL;
    }
    const EgrFramebuffer &fb = p->framebuffer;
    if (!fb.output_final || !fb.output_denoised || !fb.output_normal || !fb.output_depth) {
        set_error("egr_denoise: framebuffer planes missing");
        return EGR_EINVAL;
    }
    StateLayout L = state_layout(p->image_width, p->image_height, p->num_steps);
    if (!p->pixel_state || p->pixel_state_bytes < L.total) {
        set_error("pixel_state too small: %zu < %zu", p->pixel_state_bytes, L.total);
        return EGR_ENOMEM;
    }
    NvtxRange r_("egr:denoise");
    const size_t P = L.P, S = L.S;
    float *planes = (float *)((char *)p->pixel_state + L.planes);
    auto plane_ptr = [&](int pl) { return planes + ((size_t)pl * S + 0) * P; };  // step-0 plane `pl`
    float *A[3] = {plane_ptr(SP_REM_RGB), plane_ptr(SP_REM_RGB + 1), plane_ptr(SP_REM_RGB + 2)};
    float *B[3] = {plane_ptr(SP_REM_NORMAL), plane_ptr(SP_REM_NORMAL + 1), plane_ptr(SP_REM_NORMAL + 2)};
    DenoiseArgs a{};
    a.normal = fb.output_normal, a.depth = fb.output_depth;  // step 0 = the first [H,W,C] slab of the [S,H,W,C] planes
    a.W = (int)p->image_width, a.H = (int)p->image_height;
    const dim3 grid((a.W + 31) / 32, (a.H + 7) / 8);
    float sigma_c = EGR_DENOISE_SIGMA_COLOR;
    for (int pass = 0; pass < EGR_DENOISE_PASSES; ++pass) {
        a.in_aos = pass == 0 ? fb.output_final : nullptr;
        float **src = (pass & 1) ? A : B, **dst = (pass & 1) ? B : A;  // final -> A -> B -> A ... -> denoised
        for (int c = 0; c < 3; ++c) a.in_soa[c] = src[c], a.out_soa[c] = dst[c];
        a.out_aos = pass == EGR_DENOISE_PASSES - 1 ? fb.output_denoised : nullptr;
        a.hole = 1 << pass;
        a.inv_sigma_c2 = 1.0f / (sigma_c * sigma_c);
        k_denoise_atrous<<<grid, 256, 0, (cudaStream_t)stream>>>(a);
        sigma_c *= 0.5f;  // the colour edge-stopping function tightens as the footprint grows (Dammertz et al.)
    }
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}
