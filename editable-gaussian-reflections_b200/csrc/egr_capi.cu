// Library-level entry points of the C ABI (include/egr_b200.h): error message, version and the compile-time
// constants the reference exports as static methods (csrc/raytracer.cpp:145-148, csrc/flags.h:5-17).
#include <stdarg.h>

#include "egr_common.cuh"

namespace egr {
static thread_local char g_err[512] = "";
void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace egr

extern "C" int egr_abi_version(void) { return EGR_ABI_VERSION; }
extern "C" const char *egr_last_error(void) { return egr::g_err; }
extern "C" double egr_max_alpha(void) { return (double)egr::A_MAX; }
extern "C" int egr_roughness_downweight_grad(void) { return 1; }
extern "C" double egr_roughness_downweight_grad_power(void) { return (double)egr::ROUGH_DOWNWEIGHT_POW; }
extern "C" int egr_buffer_size(void) { return egr::KBUF; }
extern "C" int egr_max_iterations(void) { return egr::MAX_ITERATIONS; }

namespace egr {
struct PoseArgs {
    float o[3], r[9];
    int has_o, has_r;
};
__global__ void k_set_pose(EgrCamera cam, PoseArgs a) {
    const int i = threadIdx.x;
    if (a.has_o && i < 3) const_cast<float *>(cam.origin)[i] = a.o[i];
    if (a.has_r && i < 9) {
        const_cast<float *>(cam.rotation_c2w)[i] = a.r[i];
        const_cast<float *>(cam.rotation_w2c)[(i % 3) * 3 + i / 3] = a.r[i];  // camera.h:69-76
    }
}
}  // namespace egr

extern "C" int egr_camera_set_pose(const EgrCamera *cam, const float *host_origin, const float *host_rotation_c2w,
                                   void *stream) {
    if (!cam || !cam->origin || !cam->rotation_c2w || !cam->rotation_w2c) {
        egr::set_error("egr_camera_set_pose: camera pointers are null");
        return EGR_EINVAL;
    }
    egr::PoseArgs a = {};
    a.has_o = host_origin != nullptr, a.has_r = host_rotation_c2w != nullptr;
    for (int i = 0; i < 3 && a.has_o; ++i) a.o[i] = host_origin[i];
    for (int i = 0; i < 9 && a.has_r; ++i) a.r[i] = host_rotation_c2w[i];
    egr::k_set_pose<<<1, 32, 0, (cudaStream_t)stream>>>(*cam, a);
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}

namespace egr {
// One block = a tile of 256 Gaussians: the record tile is staged in shared memory with coalesced 16-byte loads, then
// each destination array is swept with consecutive threads on consecutive elements (all global traffic coalesced).
constexpr int ACC_TILE = 256;
template <int CH>
__device__ __forceinline__ void acc_array(float *D, const float *tile, const int col, const size_t g0, const int ng) {
    if (!D) return;
    D += g0 * CH;
    for (int e = threadIdx.x; e < ng * CH; e += ACC_TILE) {
        const int gl = e / CH, ch = e - gl * CH;
        D[e] += tile[gl * EGR_GRAD_STRIDE + col + ch];
    }
}
__global__ void __launch_bounds__(ACC_TILE) k_grad_accumulate(const float4 *__restrict__ rec, uint32_t n, EgrGradTargets d) {
    __shared__ __align__(16) float tile[ACC_TILE * EGR_GRAD_STRIDE];
    const size_t g0 = (size_t)blockIdx.x * ACC_TILE;
    const int ng = (int)min((size_t)ACC_TILE, (size_t)n - g0);
    constexpr int Q = EGR_GRAD_STRIDE / 4;
    for (int i = threadIdx.x; i < ng * Q; i += ACC_TILE) reinterpret_cast<float4 *>(tile)[i] = rec[g0 * Q + i];
    __syncthreads();
    acc_array<3>(d.mean, tile, 0, g0, ng);  // column map: include/egr_b200.h, EgrGaussians
    acc_array<1>(d.opacity, tile, 3, g0, ng);
    acc_array<3>(d.scale, tile, 4, g0, ng);
    acc_array<1>(d.roughness, tile, 7, g0, ng);
    acc_array<4>(d.rotation, tile, 8, g0, ng);
    acc_array<3>(d.rgb, tile, 12, g0, ng);
    acc_array<3>(d.normal, tile, 16, g0, ng);
    acc_array<3>(d.f0, tile, 20, g0, ng);
}
}  // namespace egr

extern "C" int egr_gradients_accumulate(const float *grad_records, uint32_t count, const EgrGradTargets *dst, void *stream) {
    if (!grad_records || !dst) {
        egr::set_error("egr_gradients_accumulate: null argument");
        return EGR_EINVAL;
    }
    if (count == 0) return EGR_OK;
    egr::k_grad_accumulate<<<(count + egr::ACC_TILE - 1) / egr::ACC_TILE, egr::ACC_TILE, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const float4 *>(grad_records), count, *dst);
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}

// ---- row f1: fused import + scale decay + Adam + clamps + zero_grad (include/egr_b200.h) ------------------------
namespace egr {
struct AdamConsts {
    float lr_over_bc1[8];  // lr / (1 - beta1^t)
    float one_minus_b1, beta2, one_minus_b2, eps, bc2_sqrt, scale_decay;  // each rounded from the double torch computes
    int clamp, zero_grads, zero_tw;
};
// one element: torch.optim.Adam's single-tensor formulas (torch/optim/adam.py, foreach path)
__device__ __forceinline__ void adam1(float &p, float &m, float &v, const float g, const float step_size, const AdamConsts &c) {
    m = fmaf(c.one_minus_b1, g - m, m);                   // exp_avg.lerp_(grad, 1 - beta1)
    v = fmaf(c.one_minus_b2 * g, g, v * c.beta2);         // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1-beta2)
    const float denom = sqrtf(v) / c.bc2_sqrt + c.eps;    // (exp_avg_sq.sqrt() / bias_correction2_sqrt).add_(eps)
    p = fmaf(-step_size, m / denom, p);                   // param.addcdiv_(exp_avg, denom, value=-step_size)
}
// One block = a tile of 256 Gaussians: the 24 KB tile of the gradient record is staged in shared memory with
// coalesced 16-byte loads, then every parameter array (and its two moments) is swept with consecutive threads on
// consecutive elements, so that all global traffic is coalesced although the record interleaves the eight gradients.
constexpr int ADAM_TILE = 256;
template <int CH>
__device__ __forceinline__ void adam_array(float *P, float *M, float *V, const float *tile, const int col, const size_t g0,
                                           const int ng, const float ss, const AdamConsts &c, const float lo,
                                           const float hi, const bool clamp, const bool decay) {
    if (!P) return;
    P += g0 * CH, M += g0 * CH, V += g0 * CH;
    for (int e = threadIdx.x; e < ng * CH; e += ADAM_TILE) {
        const int gl = e / CH, ch = e - gl * CH;
        float p = P[e], m = M[e], v = V[e];
        if (decay) p = logf(expf(p) * c.scale_decay);  // train.py:224-226
        adam1(p, m, v, tile[gl * EGR_GRAD_STRIDE + col + ch], ss, c);
        if (clamp) p = fminf(fmaxf(p, lo), hi);         // train.py:251-254
        P[e] = p, M[e] = m, V[e] = v;
    }
}
__global__ void __launch_bounds__(ADAM_TILE)
k_adam_step(float4 *rec, uint32_t n, EgrGradTargets P, EgrGradTargets M, EgrGradTargets V, AdamConsts c) {
    __shared__ __align__(16) float tile[ADAM_TILE * EGR_GRAD_STRIDE];
    const size_t g0 = (size_t)blockIdx.x * ADAM_TILE;
    const int ng = (int)min((size_t)ADAM_TILE, (size_t)n - g0);
    constexpr int Q = EGR_GRAD_STRIDE / 4;
    float4 *rt = rec + g0 * Q;
    for (int i = threadIdx.x; i < ng * Q; i += ADAM_TILE) reinterpret_cast<float4 *>(tile)[i] = rt[i];
    __syncthreads();
    const float INF = __int_as_float(0x7f800000);
    const bool cl = c.clamp != 0;  // column map: include/egr_b200.h, EgrGaussians
    adam_array<3>(P.mean, M.mean, V.mean, tile, 0, g0, ng, c.lr_over_bc1[0], c, -INF, INF, false, false);
    adam_array<1>(P.opacity, M.opacity, V.opacity, tile, 3, g0, ng, c.lr_over_bc1[1], c, -INF, INF, false, false);
    adam_array<3>(P.scale, M.scale, V.scale, tile, 4, g0, ng, c.lr_over_bc1[2], c, -INF, INF, false, c.scale_decay < 1.0f);
    adam_array<1>(P.roughness, M.roughness, V.roughness, tile, 7, g0, ng, c.lr_over_bc1[3], c, 0.0f, 1.0f, cl, false);
    adam_array<4>(P.rotation, M.rotation, V.rotation, tile, 8, g0, ng, c.lr_over_bc1[4], c, -INF, INF, false, false);
    adam_array<3>(P.rgb, M.rgb, V.rgb, tile, 12, g0, ng, c.lr_over_bc1[5], c, 0.0f, INF, cl, false);
    adam_array<3>(P.normal, M.normal, V.normal, tile, 16, g0, ng, c.lr_over_bc1[6], c, -INF, INF, false, false);
    adam_array<3>(P.f0, M.f0, V.f0, tile, 20, g0, ng, c.lr_over_bc1[7], c, 0.0f, 1.0f, cl, false);
    if (c.zero_grads)
        for (int i = threadIdx.x; i < ng * Q; i += ADAM_TILE)  // quad 3 of a record ends with total_weight (column 15)
            rt[i] = make_float4(0.0f, 0.0f, 0.0f, (i % Q) == 3 && !c.zero_tw ? tile[(i / Q) * EGR_GRAD_STRIDE + 15] : 0.0f);
}
}  // namespace egr

extern "C" int egr_adam_step(float *grad_records, uint32_t count, const EgrGradTargets *params,
                             const EgrGradTargets *exp_avg, const EgrGradTargets *exp_avg_sq, const EgrAdam *opt,
                             void *stream) {
    if (!grad_records || !params || !exp_avg || !exp_avg_sq || !opt) {
        egr::set_error("egr_adam_step: null argument");
        return EGR_EINVAL;
    }
    if (opt->step == 0 || !(opt->beta1 >= 0.0 && opt->beta1 < 1.0) || !(opt->beta2 >= 0.0 && opt->beta2 < 1.0)) {
        egr::set_error("egr_adam_step: step must be >= 1 and betas in [0,1)");
        return EGR_EINVAL;
    }
    const float *const *pp = (const float *const *)params, *const *pm = (const float *const *)exp_avg,
                       *const *pv = (const float *const *)exp_avg_sq;
    for (int k = 0; k < 8; ++k)
        if ((pp[k] != nullptr) != (pm[k] != nullptr) || (pp[k] != nullptr) != (pv[k] != nullptr)) {
            egr::set_error("egr_adam_step: parameter %d needs both moment arrays (or none of the three)", k);
            return EGR_EINVAL;
        }
    if (count == 0) return EGR_OK;
    egr::AdamConsts c;
    const double bc1 = 1.0 - pow(opt->beta1, (double)opt->step), bc2 = 1.0 - pow(opt->beta2, (double)opt->step);
    for (int k = 0; k < 8; ++k) c.lr_over_bc1[k] = (float)(opt->lr[k] / bc1);  // step_size = lr / bias_correction1
    c.one_minus_b1 = (float)(1.0 - opt->beta1), c.beta2 = (float)opt->beta2, c.one_minus_b2 = (float)(1.0 - opt->beta2);
    c.eps = (float)opt->eps, c.bc2_sqrt = (float)sqrt(bc2);
    c.scale_decay = (float)opt->scale_decay, c.clamp = opt->clamp, c.zero_grads = opt->zero_grads, c.zero_tw = opt->zero_total_weight;
    egr::k_adam_step<<<(count + egr::ADAM_TILE - 1) / egr::ADAM_TILE, egr::ADAM_TILE, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<float4 *>(grad_records), count, *params, *exp_avg, *exp_avg_sq, c);
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}

// ---- GaussianRaytracer._export_param_values (renderer/gaussian_raytracer.py:41-51: eight `holder.copy_(param)`) as
//      one launch: blockIdx.y = array, 16-byte words where alignment allows ------------------------------------------
namespace egr {
struct ExportArrays {
    const float *src[8];
    float *dst[8];
    uint32_t floats[8];
};
__global__ void __launch_bounds__(256) k_export(ExportArrays a) {
    const float *s = a.src[blockIdx.y];
    float *d = a.dst[blockIdx.y];
    const uint32_t n = a.floats[blockIdx.y];
    if (!s || !d) return;
    if ((((uintptr_t)s | (uintptr_t)d) & 15u) == 0) {
        const uint32_t n4 = n / 4;
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * blockDim.x)
            reinterpret_cast<float4 *>(d)[i] = reinterpret_cast<const float4 *>(s)[i];
        for (uint32_t i = n4 * 4 + blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) d[i] = s[i];
    } else {
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) d[i] = s[i];
    }
}
}  // namespace egr

extern "C" int egr_params_export(const EgrGradTargets *src, const EgrGradTargets *dst, uint32_t count, void *stream) {
    if (!src || !dst) {
        egr::set_error("egr_params_export: null argument");
        return EGR_EINVAL;
    }
    if (count == 0) return EGR_OK;
    if (count > 0x0fffffffu) {
        egr::set_error("egr_params_export: too many rows: %u", count);
        return EGR_EINVAL;
    }
    static const uint32_t ch[8] = {3, 1, 3, 1, 4, 3, 3, 3};  // EgrGradTargets order
    egr::ExportArrays a;
    const float *const *ps = (const float *const *)src;
    float *const *pd = (float *const *)dst;
    for (int k = 0; k < 8; ++k) a.src[k] = ps[k], a.dst[k] = pd[k], a.floats[k] = count * ch[k];
    const unsigned bx = (unsigned)min((size_t)148 * 4, ((size_t)count + 255) / 256);
    egr::k_export<<<dim3(bx, 8), 256, 0, (cudaStream_t)stream>>>(a);
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}
