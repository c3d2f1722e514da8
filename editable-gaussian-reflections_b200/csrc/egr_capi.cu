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
// one thread per (Gaussian, float4 column group of the record): coalesced 16-byte reads of the record
__global__ void __launch_bounds__(256) k_grad_accumulate(const float4 *__restrict__ rec, uint32_t n, EgrGradTargets d) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)n * (EGR_GRAD_STRIDE / 4)) return;
    const uint32_t g = (uint32_t)(i / (EGR_GRAD_STRIDE / 4)), q = (uint32_t)(i % (EGR_GRAD_STRIDE / 4));
    const float4 v = rec[i];
    auto add3 = [](float *p, size_t k, float x, float y, float z) {
        if (p) p[3 * k] += x, p[3 * k + 1] += y, p[3 * k + 2] += z;
    };
    switch (q) {  // column map: include/egr_b200.h, EgrGaussians
    case 0: add3(d.mean, g, v.x, v.y, v.z); if (d.opacity) d.opacity[g] += v.w; break;
    case 1: add3(d.scale, g, v.x, v.y, v.z); if (d.roughness) d.roughness[g] += v.w; break;
    case 2:
        if (d.rotation) {
            float4 *r = reinterpret_cast<float4 *>(d.rotation) + g;
            float4 o = *r;
            o.x += v.x, o.y += v.y, o.z += v.z, o.w += v.w;
            *r = o;
        }
        break;
    case 3: add3(d.rgb, g, v.x, v.y, v.z); break;
    case 4: add3(d.normal, g, v.x, v.y, v.z); break;
    default: add3(d.f0, g, v.x, v.y, v.z); break;
    }
}
}  // namespace egr

extern "C" int egr_gradients_accumulate(const float *grad_records, uint32_t count, const EgrGradTargets *dst, void *stream) {
    if (!grad_records || !dst) {
        egr::set_error("egr_gradients_accumulate: null argument");
        return EGR_EINVAL;
    }
    if (dst->rotation && ((uintptr_t)dst->rotation & 15u)) {
        egr::set_error("egr_gradients_accumulate: the rotation target must be 16-byte aligned");
        return EGR_EINVAL;
    }
    if (count == 0) return EGR_OK;
    const size_t threads = (size_t)count * (EGR_GRAD_STRIDE / 4);
    egr::k_grad_accumulate<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const float4 *>(grad_records), count, *dst);
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}
