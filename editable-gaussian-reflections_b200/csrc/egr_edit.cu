// Row f2: selection masks and in-place edit transforms (include/egr_b200.h).  Elementwise, HBM-bound:
// 12 B read + 1 B written per Gaussian for a selection; up to 44 B read + written per SELECTED Gaussian for an edit.
#include "egr_common.cuh"

namespace egr {

struct Box3 {
    float lo[3], hi[3];
};

// every operation is written with explicit round-to-nearest intrinsics: the reference runs them as separate torch
// kernels, so nothing may be contracted into an fma
__global__ void __launch_bounds__(256) k_select_box(const float *__restrict__ mean, uint32_t n, Box3 b, uint8_t *mask) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    bool in = true;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float v = mean[3 * (size_t)i + c];
        in = in && (__fsub_rn(v, b.lo[c]) >= 0.0f) && (__fsub_rn(v, b.hi[c]) <= 0.0f);
    }
    mask[i] = in ? 1 : 0;
}

__global__ void __launch_bounds__(256) k_select_cyl(const float *__restrict__ mean, uint32_t n, Box3 b, uint8_t *mask) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float cx = __fmul_rn(0.5f, __fadd_rn(b.lo[0], b.hi[0])), cy = __fmul_rn(0.5f, __fadd_rn(b.lo[1], b.hi[1]));
    const float hx = __fmul_rn(0.5f, __fsub_rn(b.hi[0], b.lo[0])), hy = __fmul_rn(0.5f, __fsub_rn(b.hi[1], b.lo[1]));
    const float x = mean[3 * (size_t)i], y = mean[3 * (size_t)i + 1], z = mean[3 * (size_t)i + 2];
    const float nx = __fdiv_rn(__fsub_rn(x, cx), hx), ny = __fdiv_rn(__fsub_rn(y, cy), hy);
    const bool ell = __fadd_rn(__fmul_rn(nx, nx), __fmul_rn(ny, ny)) <= 1.0f;
    mask[i] = (ell && z >= b.lo[2] && z <= b.hi[2]) ? 1 : 0;
}

__global__ void __launch_bounds__(256)
k_edit_apply(const uint8_t *__restrict__ mask, uint32_t n, float *mean, float *scale, float *rotation, float *opacity,
             EgrEdit e) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !mask[i]) return;
    if (mean) {
        float p[3], oc[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            oc[c] = __fadd_rn(e.center[c], e.translate[c]);                    // :218-221
            p[c] = __fadd_rn(mean[3 * (size_t)i + c], e.translate[c]);         // :208-211
            p[c] = __fadd_rn(__fmul_rn(__fsub_rn(p[c], oc[c]), e.scale), oc[c]);  // :222
            p[c] = __fsub_rn(p[c], oc[c]);                                     // :229
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) {  // (p - oc) @ R^T  (:227-233)
            const float r = fmaf(e.rotation[3 * c + 2], p[2], fmaf(e.rotation[3 * c + 1], p[1], __fmul_rn(e.rotation[3 * c], p[0])));
            mean[3 * (size_t)i + c] = __fadd_rn(r, oc[c]);
        }
    }
    if (scale) {
#pragma unroll
        for (int c = 0; c < 3; ++c)  // :242,251,253
            scale[3 * (size_t)i + c] = logf(__fmul_rn(expf(scale[3 * (size_t)i + c]), e.scale));
    }
    if (rotation) {
        float4 q = reinterpret_cast<float4 *>(rotation)[i];  // (r, x, y, z)
        const float inv = 1.0f / sqrtf(fmaf(q.w, q.w, fmaf(q.z, q.z, fmaf(q.y, q.y, q.x * q.x))));
        q.x *= inv, q.y *= inv, q.z *= inv, q.w *= inv;
        const float a = e.quaternion[0], b = e.quaternion[1], c = e.quaternion[2], d = e.quaternion[3];
        float4 o;  // Hamilton product q_edit (x) q: the edit rotation applied after the Gaussian's own
        o.x = a * q.x - b * q.y - c * q.z - d * q.w;
        o.y = a * q.y + b * q.x + c * q.w - d * q.z;
        o.z = a * q.z - b * q.w + c * q.x + d * q.y;
        o.w = a * q.w + b * q.z - c * q.y + d * q.x;
        const float io = 1.0f / sqrtf(fmaf(o.w, o.w, fmaf(o.z, o.z, fmaf(o.y, o.y, o.x * o.x))));
        o.x *= io, o.y *= io, o.z *= io, o.w *= io;
        reinterpret_cast<float4 *>(rotation)[i] = o;
    }
    if (opacity && e.kill) opacity[i] = __fsub_rn(__fmul_rn(opacity[i], 0.0f), 100000000.0f);  // :326-327
}

static int launch_select(bool cyl, const float *mean, uint32_t n, const float *min3, const float *max3, uint8_t *mask,
                         cudaStream_t st) {
    if (n == 0) return EGR_OK;
    if (!mean || !min3 || !max3 || !mask) {
        set_error("egr_select_*: null argument");
        return EGR_EINVAL;
    }
    Box3 b;
    for (int c = 0; c < 3; ++c) b.lo[c] = min3[c], b.hi[c] = max3[c];
    if (cyl) k_select_cyl<<<(n + 255) / 256, 256, 0, st>>>(mean, n, b, mask);
    else k_select_box<<<(n + 255) / 256, 256, 0, st>>>(mean, n, b, mask);
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}

}  // namespace egr

using namespace egr;

extern "C" int egr_select_box(const float *mean, uint32_t n, const float *min3, const float *max3, uint8_t *mask,
                              void *stream) {
    return launch_select(false, mean, n, min3, max3, mask, (cudaStream_t)stream);
}
extern "C" int egr_select_cylinder(const float *mean, uint32_t n, const float *min3, const float *max3, uint8_t *mask,
                                   void *stream) {
    return launch_select(true, mean, n, min3, max3, mask, (cudaStream_t)stream);
}
extern "C" int egr_edit_apply(const uint8_t *mask, uint32_t n, float *mean, float *scale, float *rotation,
                              float *opacity, const EgrEdit *edit, void *stream) {
    if (n == 0) return EGR_OK;
    if (!mask || !edit) {
        set_error("egr_edit_apply: null mask / edit");
        return EGR_EINVAL;
    }
    if (rotation && ((uintptr_t)rotation & 15u)) {
        set_error("egr_edit_apply: rotation must be 16-byte aligned");
        return EGR_EINVAL;
    }
    k_edit_apply<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(mask, n, mean, scale, rotation, opacity, *edit);
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}
