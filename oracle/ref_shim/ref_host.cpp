// =============================================================================================
// ORACLE 2 — TEST INFRASTRUCTURE ONLY.  Compiles the REFERENCE'S OWN device sources
//   $(REF)/editable_gauss_refl/cuda/csrc/{shaders.cu, forward_pass.cu, backward_pass.cu, params.h,
//   core/*.h, utils/*}
// as plain host C++, included IN PLACE from the read-only reference checkout (nothing is copied into
// this repository), so that oracle/egr_oracle.cpp can be pinned against the literal arithmetic of the
// reference.  Output: oracle/_ref/libref_tracer.so (git-ignored).
//
// What is NOT the reference here, and why:
//   * OptiX device intrinsics -> ref_shim/optix.h (payload registers, launch index, object-space ray);
//   * optixTraverse           -> brute-force loop over the instances with an exact segment-vs-unit-box
//                                test (the acceleration structure only enumerates candidates);
//   * the instance transform  -> restated from optix/bvh_wrapper.cu:9-59 (that file cannot be compiled
//                                by a host compiler because of its <<< >>> launch); the inverse is
//                                analytic, the object-space origin is W (o - mean);
//   * argument evaluation order of make_float2(rnd(seed), rnd(seed)) forced to nvcc's left-to-right;
//   * atomicAdd               -> real host atomics; one raygen "thread" per pixel under OpenMP, each
//                                OS thread owning a private per-pixel linked-list scratch.
// =============================================================================================
#include <math.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <limits>
#include <random>
#include <vector>

#include <vector_types.h>
#include <vector_functions.h>
#ifdef _OPENMP
#include <omp.h>
#endif

// ---- pretend to be the device compiler --------------------------------------------------------------
#undef __device__
#undef __host__
#undef __global__
#undef __forceinline__
#undef __inline__
#undef __constant__
#define __device__
#define __host__
#define __global__
#define __forceinline__ inline
#define __inline__ inline
#define __constant__ thread_local  // `params` becomes one private launch-parameter block per OS thread
#define register
#define __CUDACC__ 1

inline int max(int a, int b) { return a > b ? a : b; }
inline int min(int a, int b) { return a < b ? a : b; }
inline long long max(long long a, long long b) { return a > b ? a : b; }
inline long long min(long long a, long long b) { return a < b ? a : b; }
inline unsigned int max(unsigned int a, unsigned int b) { return a > b ? a : b; }
inline unsigned int min(unsigned int a, unsigned int b) { return a < b ? a : b; }
inline unsigned long long max(unsigned long long a, unsigned long long b) { return a > b ? a : b; }
inline unsigned long long min(unsigned long long a, unsigned long long b) { return a < b ? a : b; }
inline float max(float a, float b) { return fmaxf(a, b); }
inline float min(float a, float b) { return fminf(a, b); }
inline double max(double a, double b) { return fmax(a, b); }
inline double min(double a, double b) { return fmin(a, b); }

inline float __uint_as_float(unsigned int u) {
    float f;
    std::memcpy(&f, &u, 4);
    return f;
}
inline unsigned int __float_as_uint(float f) {
    unsigned int u;
    std::memcpy(&u, &f, 4);
    return u;
}
inline float atomicAdd(float *addr, float v) {
    uint32_t *ia = reinterpret_cast<uint32_t *>(addr);
    uint32_t old = __atomic_load_n(ia, __ATOMIC_RELAXED), nw;
    float of;
    do {
        std::memcpy(&of, &old, 4);
        float nf = of + v;
        std::memcpy(&nw, &nf, 4);
    } while (!__atomic_compare_exchange_n(ia, &old, nw, true, __ATOMIC_RELAXED, __ATOMIC_RELAXED));
    return of;
}
inline int atomicAdd(int *addr, int v) { return __atomic_fetch_add(addr, v, __ATOMIC_RELAXED); }
inline unsigned int atomicAdd(unsigned int *addr, unsigned int v) { return __atomic_fetch_add(addr, v, __ATOMIC_RELAXED); }
inline unsigned int atomicAdd(unsigned int *addr, int v) {
    return __atomic_fetch_add(addr, (unsigned int)v, __ATOMIC_RELAXED);
}

// ---- the reference's sources, in place ------------------------------------------------------------------
#include "optix.h"
#include "utils/vec_math.h"
#include "utils/random.h"
// nvcc evaluates the two rnd(seed) arguments of make_float2 left to right (visible in the reference's PTX);
// g++ would evaluate them right to left, so the three 2-argument call sites outside vec_math.h
// (core/camera.h:24,26, shaders.cu:130) are forced to the device compiler's order.
template <class FA, class FB> inline float2 ref_float2_lr(FA a, FB b) {
    float x = a();
    float y = b();
    return make_float2(x, y);
}
#define make_float2(a, b) ref_float2_lr([&]() -> float { return (a); }, [&]() -> float { return (b); })
#include "shaders.cu"  // pulls params.h, core/*, utils/*, forward_pass.cu, backward_pass.cu
#undef make_float2

// ---- stand-ins declared in optix.h ------------------------------------------------------------------------
thread_local RefThreadState ref_ts;
const RefInstance *ref_instances = nullptr;
uint32_t ref_num_instances = 0;

void optixTraverse(OptixTraversableHandle, float3 o, float3 d, float tmin, float tmax, float, OptixVisibilityMask,
                   unsigned, unsigned, unsigned, unsigned, uint32_t &p0, uint32_t &p1, uint32_t &p2, uint32_t &p3,
                   uint32_t &p4, uint32_t &p5, uint32_t &p6, uint32_t &p7) {
    uint32_t *pl[8] = {&p0, &p1, &p2, &p3, &p4, &p5, &p6, &p7};
    for (int i = 0; i < 8; ++i) ref_ts.payload[i] = pl[i];
    // OptiX treats a ray with a non-finite origin / direction as invalid: nothing is traversed
    if (!std::isfinite(o.x + o.y + o.z + d.x + d.y + d.z) || !(dot(d, d) > 0.0f)) return;
    for (uint32_t i = 0; i < ref_num_instances; ++i) {
        const RefInstance &in = ref_instances[i];
        if (!in.visible) continue;
        float3 pc = in.center - o;
        float b = dot(pc, d);
        if (dot(pc, pc) - b * b > in.radius2) continue;  // conservative bounding-sphere reject
        float3 w0 = make_float3(in.w2l[0]), w1 = make_float3(in.w2l[1]), w2 = make_float3(in.w2l[2]);
        float3 p = o - in.center;
        float3 ol = make_float3(dot(w0, p), dot(w1, p), dot(w2, p));
        float3 dl = make_float3(dot(w0, d), dot(w1, d), dot(w2, d));
        // segment [tmin,tmax] against the unit box in object space
        float tn = tmin, tf = tmax;
        const float olv[3] = {ol.x, ol.y, ol.z}, dlv[3] = {dl.x, dl.y, dl.z};
        bool miss = false;
        for (int a = 0; a < 3; ++a) {
            if (dlv[a] == 0.0f) {
                if (fabsf(olv[a]) > 1.0f) miss = true;
                continue;
            }
            float ta = (-1.0f - olv[a]) / dlv[a], tb = (1.0f - olv[a]) / dlv[a];
            tn = fmaxf(tn, fminf(ta, tb));
            tf = fminf(tf, fmaxf(ta, tb));
        }
        if (miss || tn > tf) continue;
        ref_ts.obj_origin = ol;
        ref_ts.obj_direction = dl;
        ref_ts.instance_index = i;
        __intersection__gaussian();
    }
}

// ---- driver ------------------------------------------------------------------------------------------------
extern "C" {

struct OrcConfig {
    float exp_power, alpha_threshold, transmittance_threshold, global_scale_factor;
    float loss_weight_diffuse, loss_weight_specular, loss_weight_depth, loss_weight_normal, loss_weight_f0,
        loss_weight_roughness;
    float eps_forward_normalization, eps_scale_grad, eps_ray_surface_offset, eps_min_roughness;
    float reflection_invalid_normal_threshold, backfacing_invalid_normal_threshold, backfacing_max_dist;
    int32_t num_bounces, jitter_primary_rays, accumulate_samples, max_bounces;
};
struct OrcCamera {
    float origin[3];
    float rot_c2w[9];
    float fovy, znear, zfar;
};
struct OrcScene {
    int32_t n;
    const float *mean, *scale, *rotation, *opacity, *rgb, *normal, *f0, *roughness;
};
struct OrcTargets {
    const float *diffuse, *specular, *depth, *normal, *f0, *roughness;
};
struct OrcOutputs {
    float *rgb, *depth, *normal, *f0, *roughness, *transmittance, *total_transmittance, *ray_origin, *ray_direction;
    float *final_;
    float *acc_rgb, *acc_transmittance, *acc_total_transmittance, *acc_depth, *acc_normal, *acc_f0, *acc_roughness;
    int32_t *acc_count;
    uint32_t *random_seeds;
    int32_t *num_accumulated, *num_hits_per_pixel, *num_composited_per_pixel, *num_steps_per_pixel;
    float *used_ray_origin, *used_ray_direction;
    int32_t *step_executed;
};
struct OrcGrads {
    float *dL_drgb, *dL_dnormal, *dL_df0, *dL_droughness, *dL_dopacity, *dL_dscale, *dL_dmean, *dL_drotation;
    float *total_weight;
};

int ref_max_bounces() { return MAX_BOUNCES; }
int ref_num_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

// One Raytracer::raytrace() of the reference (raytracer.cpp:81-94) on host cores.  Layouts as orc_render.
// `out->num_hits_per_pixel` receives stats.num_traversed_per_pixel (intersection-program invocations).
int ref_render(const OrcScene *sc, const OrcConfig *cfg, const OrcCamera *cam, int width, int height,
               uint32_t total_num_calls, int grads_enabled, const OrcTargets *tg, OrcOutputs *out, OrcGrads *gr,
               int pixel_begin, int pixel_end) {
    if (cfg->max_bounces != MAX_BOUNCES) return -1;  // compile-time constant of the reference (flags.h:5)
    const int n = sc->n, P = width * height, S = MAX_BOUNCES + 1;
    if (pixel_end <= 0 || pixel_end > P) pixel_end = P;
    if (pixel_begin < 0) pixel_begin = 0;

    // instance transforms: restated from optix/bvh_wrapper.cu:9-59
    std::vector<RefInstance> inst(n);
    for (int i = 0; i < n; ++i) {
        float opacity = 1.0f / (1.0f + exp(-sc->opacity[i]));  // sigmoid_act
        float sf = compute_scaling_factor(opacity, cfg->alpha_threshold, cfg->exp_power);
        float3 sizes = make_float3(exp(sc->scale[3 * i]), exp(sc->scale[3 * i + 1]), exp(sc->scale[3 * i + 2]));
        sizes = sizes * sf * cfg->global_scale_factor;
        RefInstance &in = inst[i];
        in.visible = sf > 0.0f && (sizes.x > 0.0f || sizes.y > 0.0f || sizes.z > 0.0f);
        float r = sc->rotation[4 * i], x = sc->rotation[4 * i + 1], y = sc->rotation[4 * i + 2], z = sc->rotation[4 * i + 3];
        float norm = sqrtf(r * r + x * x + y * y + z * z);
        r /= norm, x /= norm, y /= norm, z /= norm;
        float3 s = sizes, p = make_float3(sc->mean[3 * i], sc->mean[3 * i + 1], sc->mean[3 * i + 2]);
        float R[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                         {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                         {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
        in.l2w[0] = make_float4(s.x * R[0][0], s.y * R[0][1], s.z * R[0][2], p.x);
        in.l2w[1] = make_float4(s.x * R[1][0], s.y * R[1][1], s.z * R[1][2], p.y);
        in.l2w[2] = make_float4(s.x * R[2][0], s.y * R[2][1], s.z * R[2][2], p.z);
        float sv[3] = {s.x, s.y, s.z};
        for (int rr = 0; rr < 3; ++rr) {
            float3 w = make_float3(R[0][rr] / sv[rr], R[1][rr] / sv[rr], R[2][rr] / sv[rr]);
            in.w2l[rr] = make_float4(w.x, w.y, w.z, -dot(w, p));
        }
        if (!(s.x > 0.0f && s.y > 0.0f && s.z > 0.0f) || !std::isfinite(norm) || norm <= 0.0f) in.visible = 0;
        in.center = p;
        float rad = 1.7320508f * fmaxf(s.x, fmaxf(s.y, s.z)) * 1.001f + 1e-6f;
        in.radius2 = rad * rad;
    }
    ref_instances = inst.data();
    ref_num_instances = (uint32_t)n;

    // device scalars (config.h, camera.h, metadata.h)
    bool b_acc = cfg->accumulate_samples != 0, b_jit = cfg->jitter_primary_rays != 0, b_grads = grads_enabled != 0;
    int i_nb = cfg->num_bounces;
    uint32_t calls = total_num_calls;
    float3 cam_o = make_float3(cam->origin[0], cam->origin[1], cam->origin[2]);
    float3 c2w[3] = {make_float3(cam->rot_c2w[0], cam->rot_c2w[1], cam->rot_c2w[2]),
                     make_float3(cam->rot_c2w[3], cam->rot_c2w[4], cam->rot_c2w[5]),
                     make_float3(cam->rot_c2w[6], cam->rot_c2w[7], cam->rot_c2w[8])};
    float3 w2c[3] = {make_float3(c2w[0].x, c2w[1].x, c2w[2].x), make_float3(c2w[0].y, c2w[1].y, c2w[2].y),
                     make_float3(c2w[0].z, c2w[1].z, c2w[2].z)};
    std::vector<float> dummy3((size_t)P * 3 * S, 0.0f);
    std::vector<int> stat_trav(P, 0), stat_acc(P, 0);
    std::vector<float> zero_t((size_t)P * 3, 0.0f);

    Params base;
    std::memset(&base, 0, sizeof(base));
    base.image_width = width, base.image_height = height;
    base.camera = Camera{&cam_o, &cam->fovy, c2w, w2c, &cam->znear, &cam->zfar};
    base.config = Config{&cfg->exp_power, &cfg->alpha_threshold, &cfg->transmittance_threshold, &b_acc, &b_jit, &i_nb,
                         &cfg->global_scale_factor, &cfg->loss_weight_diffuse, &cfg->loss_weight_specular,
                         &cfg->loss_weight_depth, &cfg->loss_weight_normal, &cfg->loss_weight_f0,
                         &cfg->loss_weight_roughness, &cfg->eps_forward_normalization, &cfg->eps_scale_grad,
                         &cfg->eps_ray_surface_offset, &cfg->eps_min_roughness,
                         &cfg->reflection_invalid_normal_threshold, &cfg->backfacing_invalid_normal_threshold,
                         &cfg->backfacing_max_dist};
    Framebuffer &fb = base.framebuffer;
    fb.output_rgb = (float3 *)out->rgb, fb.output_depth = out->depth, fb.output_normal = (float3 *)out->normal;
    fb.output_f0 = (float3 *)out->f0, fb.output_roughness = out->roughness;
    fb.output_transmittance = out->transmittance, fb.output_total_transmittance = out->total_transmittance;
    fb.output_ray_origin = (float3 *)out->ray_origin, fb.output_ray_direction = (float3 *)out->ray_direction;
    fb.output_final = (float3 *)out->final_, fb.output_denoised = nullptr;
    fb.accumulated_rgb = (float3 *)out->acc_rgb, fb.accumulated_transmittance = out->acc_transmittance;
    fb.accumulated_total_transmittance = out->acc_total_transmittance, fb.accumulated_depth = out->acc_depth;
    fb.accumulated_normal = (float3 *)out->acc_normal, fb.accumulated_f0 = (float3 *)out->acc_f0;
    fb.accumulated_roughness = out->acc_roughness, fb.accumulated_sample_count = out->acc_count;
    const float *z = zero_t.data();
    fb.target_diffuse = (const float3 *)(tg && tg->diffuse ? tg->diffuse : z);
    fb.target_specular = (const float3 *)(tg && tg->specular ? tg->specular : z);
    fb.target_depth = tg && tg->depth ? tg->depth : z;
    fb.target_normal = (const float3 *)(tg && tg->normal ? tg->normal : z);
    fb.target_f0 = (const float3 *)(tg && tg->f0 ? tg->f0 : z);
    fb.target_roughness = tg && tg->roughness ? tg->roughness : z;
    fb.num_pixels = (uint32_t)P;
    Gaussians &g = base.gaussians;
    g.count = (uint32_t)n;
    g.rgb = (const float3 *)sc->rgb, g.normal = (const float3 *)sc->normal, g.f0 = (const float3 *)sc->f0;
    g.roughness = sc->roughness, g.opacity = sc->opacity, g.scale = (const float3 *)sc->scale;
    g.mean = (const float3 *)sc->mean, g.rotation = (const float4 *)sc->rotation;
    g.dL_drgb = (float3 *)gr->dL_drgb, g.dL_dnormal = (float3 *)gr->dL_dnormal, g.dL_df0 = (float3 *)gr->dL_df0;
    g.dL_droughness = gr->dL_droughness, g.dL_dopacity = gr->dL_dopacity, g.dL_dscale = (float3 *)gr->dL_dscale;
    g.dL_dmean = (float3 *)gr->dL_dmean, g.dL_drotation = (float4 *)gr->dL_drotation, g.total_weight = gr->total_weight;
    base.metadata = Metadata{&b_grads, &calls, out->random_seeds};
    base.stats = Stats{stat_acc.data(), stat_trav.data()};
    base.bvh_handle = 0;

    const size_t cap = (size_t)S * (size_t)n + 64;  // worst case of one pixel: every instance at every step
#pragma omp parallel
    {
        std::vector<uint32_t> head_f(P, PerPixelLinkedList::NULL_PTR), head_b(P, PerPixelLinkedList::NULL_PTR);
        struct Store {
            std::vector<uint32_t> ids, prev;
            std::vector<float> dist, alpha, trans, gv;
            std::vector<float3> lh;
            uint32_t total = 0;
            explicit Store(size_t c) : ids(c), prev(c), dist(c), alpha(c), trans(c), gv(c), lh(c) {}
            PerPixelLinkedList view(uint32_t *head) {
                PerPixelLinkedList l;
                l.head_per_pixel = head, l.total_hits = &total, l.gaussian_ids = ids.data(), l.distances = dist.data();
                l.alphas = alpha.data(), l.transmittances = trans.data(), l.local_hits = lh.data();
                l.gaussvals = gv.data(), l.previous_entries = prev.data();
                return l;
            }
        } sf_(cap), sb_(cap);
        params = base;  // this thread's private launch-parameter block
        params.ppll_forward = sf_.view(head_f.data());
        params.ppll_backward = sb_.view(head_b.data());
#pragma omp for schedule(dynamic, 16)
        for (int pix = pixel_begin; pix < pixel_end; ++pix) {
            sf_.total = 0, sb_.total = 0;
            head_f[pix] = PerPixelLinkedList::NULL_PTR, head_b[pix] = PerPixelLinkedList::NULL_PTR;
            ref_ts.launch_index = make_uint3(pix % width, pix / width, 0);
            ref_ts.launch_dim = make_uint3(width, height, 1);
            __raygen__rg();
        }
    }
    for (int i = 0; i < P; ++i) {
        if (out->num_accumulated) out->num_accumulated[i] = stat_acc[i];
        if (out->num_hits_per_pixel) out->num_hits_per_pixel[i] = stat_trav[i];
    }
    ref_instances = nullptr;
    return 0;
}
}
