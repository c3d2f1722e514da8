// =============================================================================================
// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product; nothing under the product package
// may import, link or execute this file.  Only tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py use it, and only as the checker.
//
// CPU restatement (C++17 + OpenMP, brute force over Gaussians, no BVH) of the differentiable
// Gaussian ray tracer of graphdeco-inria/editable-gaussian-reflections.  Citations are relative
// to /root/reference/editable_gauss_refl/cuda/csrc/ :
//   shaders.cu:9-75       per-candidate accept/reject, response, exact total transmittance
//   shaders.cu:77-173     step loop, bounce spawn, throughput, final, grad/no-grad outputs
//   forward_pass.cu:3-156 16-wide batch selection with strict window, compositing, remaining fill
//   backward_pass.cu:3-222 in-kernel L1 gradients, back-to-front suffix sums, geometry gradients
//   utils/kernel.cu:3-16, utils/activations.cu, utils/helpers.cu:10-58, utils/misc.cu:56-63
//   utils/random.h:34-62 (tea<4>, lcg, rnd), utils/ggx_brdf.h:32-37,63,83,134-168
//   core/camera.h:17-36, optix/bvh_wrapper.cu:9-59, core/framebuffer.h:104-152
//
// Parity status: the reference's own tests hold no golden vector for this path (its tracer is
// GPU/OptiX only).  This restatement is pinned instead against the reference's literal shader
// sources compiled as host C++ (oracle/ref_shim -> oracle/_ref, see oracle/README.md) and the
// fixtures under tests/golden/ record that agreement.
//
// Arithmetic conventions (documented in DESIGN.md, mirrored by the CUDA kernels so that hit-ID
// lists can be compared bit-exactly):
//   * dot3(a,b) = fmaf(a.z,b.z, fmaf(a.y,b.y, a.x*b.x)) on the accept/reject + distance path;
//   * exp() of the Gaussian response is evaluated by det_exp() (Cody-Waite + degree-7 Horner,
//     <= 2 ulp), integer exp_power by repeated multiplication; both are IEEE-only sequences and
//     therefore bit-identical on CPU and GPU;
//   * candidate enumeration = every visible Gaussian whose local unit box [-1,1]^3 is crossed by
//     the ray segment [tmin,tmax] (OptiX AABB semantics), tested explicitly.
// Build: g++ -O2 -fopenmp -ffp-contract=off -shared -fPIC (see oracle/Makefile).
// =============================================================================================
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

// exp / log / sin / cos as IEEE-only sequences shared with the CUDA kernels (pure arithmetic definitions; the
// tracing algorithm below is restated independently)
#include "../include/egr_detmath.h"

namespace {

struct f3 {
    float x, y, z;
};
inline f3 mk(float x, float y, float z) { return f3{x, y, z}; }
inline f3 operator+(f3 a, f3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
inline f3 operator-(f3 a, f3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
inline f3 operator*(f3 a, f3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
inline f3 operator*(f3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
inline f3 operator*(float s, f3 a) { return mk(a.x * s, a.y * s, a.z * s); }
inline f3 operator/(f3 a, f3 b) { return mk(a.x / b.x, a.y / b.y, a.z / b.z); }
inline f3 div_s(f3 a, float s) {  // vec_math.h:330-333: float3 / float multiplies by the reciprocal
    float inv = 1.0f / s;
    return a * inv;
}
inline f3 operator-(f3 a) { return mk(-a.x, -a.y, -a.z); }
// canonical dot product of the ID-critical path
inline float dot3(f3 a, f3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
// plain dot product (vec_math.h:365 order) for the non-critical shading / gradient code
inline float dotp(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline f3 cross(f3 a, f3 b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline float length(f3 a) { return sqrtf(dotp(a, a)); }
inline f3 normalize(f3 v) {  // vec_math.h:376-379
    float inv = 1.0f / sqrtf(dotp(v, v));
    return v * inv;
}
inline f3 ld3(const float *p, size_t i) { return mk(p[3 * i], p[3 * i + 1], p[3 * i + 2]); }
inline void st3(float *p, size_t i, f3 v) {
    p[3 * i] = v.x;
    p[3 * i + 1] = v.y;
    p[3 * i + 2] = v.z;
}
inline float sign1(float v) { return copysignf(1.0f, v); }  // misc.cu:56-63 (sign(0)=+1, sign(-0)=-1)
inline f3 sign3(f3 v) { return mk(sign1(v.x), sign1(v.y), sign1(v.z)); }

constexpr float A_MAX = 0.9999f;        // flags.h:8
constexpr int BUFFER_SIZE = 16;         // flags.h:16
constexpr int MAX_ITERATIONS = 99;      // flags.h:17
constexpr float ROUGH_DOWNWEIGHT_POW = 3.0f;  // flags.h:12-13

// ---- exp / integer powers (include/egr_detmath.h) ------------------------------------------------
inline float det_exp(float u) { return egr_det_exp(u); }
// d^p: exp_power 3 (the reference's default) and other small integers by left-to-right products
inline float det_pow(float d, float p) {
    if (p == 3.0f) return (d * d) * d;
    int ip = (int)p;
    if ((float)ip == p && ip >= 1 && ip <= 8) {
        float r = d;
        for (int i = 1; i < ip; ++i) r = r * d;
        return r;
    }
    return egr_det_pow(d, p);
}
// canonical explicit-fma forms (same op order as the CUDA kernels, so that the bits agree)
inline f3 fma3(f3 a, float s, f3 c) { return mk(fmaf(a.x, s, c.x), fmaf(a.y, s, c.y), fmaf(a.z, s, c.z)); }
inline f3 cross3c(f3 a, f3 b) {
    return mk(fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x)));
}
inline f3 norm3c(f3 v) {
    float inv = 1.0f / sqrtf(dot3(v, v));
    return v * inv;
}

// ---- RNG (utils/random.h:34-62) -----------------------------------------------------------
inline uint32_t tea4(uint32_t v0, uint32_t v1) {
    uint32_t s0 = 0;
    for (int n = 0; n < 4; ++n) {
        s0 += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    return v0;
}
inline float rnd(uint32_t &prev) {
    prev = 1664525u * prev + 1013904223u;
    return (float)(prev & 0x00FFFFFFu) / (float)0x01000000;
}

}  // namespace

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
    float rot_c2w[9];  // row-major; rotation_w2c is its transpose (camera.h:62-68)
    float fovy, znear, zfar;
};
struct OrcScene {
    int32_t n;
    const float *mean, *scale, *rotation, *opacity, *rgb, *normal, *f0, *roughness;  // raw (pre-activation)
};
struct OrcTargets {  // [P,C]; may be all-null when grads are off
    const float *diffuse, *specular, *depth, *normal, *f0, *roughness;
};
struct OrcOutputs {  // [S,P,C] step-major planes (framebuffer.h:108,132); any pointer may be null
    float *rgb, *depth, *normal, *f0, *roughness, *transmittance, *total_transmittance, *ray_origin, *ray_direction;
    float *final_;  // [P,3]
    float *acc_rgb, *acc_transmittance, *acc_total_transmittance, *acc_depth, *acc_normal, *acc_f0, *acc_roughness;
    int32_t *acc_count;  // [1], read only here (the host bumps it after the launch, raytracer.cpp:91-93)
    uint32_t *random_seeds;           // [P]
    int32_t *num_accumulated;         // [P] composited hits of the LAST executed forward_pass (forward_pass.cu:140)
    int32_t *num_hits_per_pixel;      // [P] accepted hits summed over steps
    int32_t *num_composited_per_pixel;  // [P] composited hits summed over steps
    int32_t *num_steps_per_pixel;     // [P] executed ray-steps
    float *used_ray_origin, *used_ray_direction;  // [S,P,3] the ray traced at each step
    int32_t *step_executed;           // [S,P]
};
struct OrcGrads {  // accumulated INTO (+=); any pointer may be null only if grads are off
    float *dL_drgb, *dL_dnormal, *dL_df0, *dL_droughness, *dL_dopacity, *dL_dscale, *dL_dmean, *dL_drotation;
    float *total_weight;
};
struct OrcHitLists {  // optional (cap = 0 disables)
    int32_t cap;
    int32_t *counts;  // [S,P] composited count (not capped)
    int32_t *ids;     // [S,P,cap]
    float *ts;        // [S,P,cap]
};
struct OrcOverride {  // optional: trace these rays instead of the oracle's own (bit-exact ID tests)
    const float *ray_origin, *ray_direction;  // [S,P,3]
    const int32_t *step_executed;             // [S,P]
    const float *records;                     // [N,16] packed geometry records, layout of orc_prepare
    const uint32_t *leaf_rank;                // [N] position of each Gaussian in the device's leaf order
};

// test hooks for the integer-exact / IEEE-only building blocks
float orc_det_exp(float u) { return det_exp(u); }
float orc_det_log(float x) { return egr_det_log(x); }
float orc_det_sin(float x) {
    float s_, c_;
    egr_det_sincos(x, &s_, &c_);
    return s_;
}
float orc_det_cos(float x) {
    float s_, c_;
    egr_det_sincos(x, &s_, &c_);
    return c_;
}
uint32_t orc_tea4(uint32_t a, uint32_t b) { return tea4(a, b); }
float orc_rnd(uint32_t *seed) { return rnd(*seed); }

int orc_num_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

// Record layout (16 floats, same as the CUDA geometry record): mean.xyz, squared bounding radius (0 = invisible) |
// W row0 xyz, sf | W row1 xyz, sigma | W row2 xyz, id bits        (W = world -> object = diag(1/S) R^T)
// bvh_wrapper.cu:9-59 (transform, visibility), kernel.cu:3-6 (scaling factor), activations.cu:5,59,67-69
void orc_prepare(const OrcScene *sc, const OrcConfig *cfg, float *rec) {
    const float k = 2.0f * cfg->exp_power;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < sc->n; ++i) {
        float sigma = 1.0f / (1.0f + egr_det_exp(-sc->opacity[i]));
        float sf = sigma <= cfg->alpha_threshold
                       ? 0.0f
                       : egr_det_pow(k * egr_det_log(sigma / cfg->alpha_threshold), 1.0f / k);
        f3 s = mk(egr_det_exp(sc->scale[3 * i]), egr_det_exp(sc->scale[3 * i + 1]), egr_det_exp(sc->scale[3 * i + 2]));
        f3 S = s * sf * cfg->global_scale_factor;
        float r = sc->rotation[4 * i], x = sc->rotation[4 * i + 1], y = sc->rotation[4 * i + 2],
              z = sc->rotation[4 * i + 3];
        float norm = sqrtf(fmaf(z, z, fmaf(y, y, fmaf(x, x, r * r))));
        r /= norm, x /= norm, y /= norm, z /= norm;
        float R[3][3] = {
            {fmaf(-2.f, fmaf(z, z, y * y), 1.f), 2.f * fmaf(x, y, -(r * z)), 2.f * fmaf(x, z, r * y)},
            {2.f * fmaf(x, y, r * z), fmaf(-2.f, fmaf(z, z, x * x), 1.f), 2.f * fmaf(y, z, -(r * x))},
            {2.f * fmaf(x, z, -(r * y)), 2.f * fmaf(y, z, r * x), fmaf(-2.f, fmaf(y, y, x * x), 1.f)}};
        // The reference marks an instance visible iff sf > 0 and any size > 0; with all sizes required > 0 the
        // analytic inverse W = diag(1/S) R^T exists (a zero size is a degenerate transform OptiX cannot invert).
        bool visible = sf > 0.0f && S.x > 0.0f && S.y > 0.0f && S.z > 0.0f && std::isfinite(norm) && norm > 0.0f &&
                       std::isfinite(sc->mean[3 * i] + sc->mean[3 * i + 1] + sc->mean[3 * i + 2]) &&
                       std::isfinite(S.x + S.y + S.z);
        float Sv[3] = {S.x, S.y, S.z};
        float *o = rec + 16 * (size_t)i;
        float rad = fmaxf(S.x, fmaxf(S.y, S.z)) * 1.001f;
        o[0] = sc->mean[3 * i], o[1] = sc->mean[3 * i + 1], o[2] = sc->mean[3 * i + 2];
        o[3] = visible ? rad * rad : 0.0f;
        for (int rr = 0; rr < 3; ++rr)
            for (int c = 0; c < 3; ++c) o[4 + 4 * rr + c] = visible ? R[c][rr] / Sv[rr] : 0.0f;
        o[7] = sf;
        o[11] = sigma;
        uint32_t id = (uint32_t)i;
        std::memcpy(&o[15], &id, 4);
    }
}

}  // extern "C"

namespace {

// Morton (leaf) order of the Gaussians, as the CUDA LBVH build defines it: 63-bit code of the mean on a 2^21 grid
// over the bounds of the finite means, stable sort (ties keep id order).  Only used to take the product of the
// total transmittance in the same order as the device (float products do not commute bit for bit).
inline uint64_t expand21(uint32_t v) {
    uint64_t x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}
std::vector<uint32_t> morton_rank(const float *mean, int n) {
    float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
    for (int i = 0; i < n; ++i)
        for (int c = 0; c < 3; ++c) {
            float v = mean[3 * (size_t)i + c];
            if (std::isfinite(v)) lo[c] = fminf(lo[c], v), hi[c] = fmaxf(hi[c], v);
        }
    std::vector<uint64_t> key(n);
    for (int i = 0; i < n; ++i) {
        uint32_t q[3];
        for (int c = 0; c < 3; ++c) {
            float ext = hi[c] - lo[c];
            float v = mean[3 * (size_t)i + c];
            float u = ext > 0.0f ? (v - lo[c]) / ext : 0.0f;
            u = std::isfinite(u) ? fminf(fmaxf(u, 0.0f), 1.0f) : 0.0f;
            uint32_t qq = (uint32_t)(u * 2097152.0f);
            q[c] = qq < 2097151u ? qq : 2097151u;
        }
        key[i] = (expand21(q[0]) << 2) | (expand21(q[1]) << 1) | expand21(q[2]);
    }
    std::vector<uint32_t> order(n), rank(n);
    for (int i = 0; i < n; ++i) order[i] = (uint32_t)i;
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return key[a] < key[b]; });
    for (int i = 0; i < n; ++i) rank[order[i]] = (uint32_t)i;
    return rank;
}

struct Hit {
    float t;
    int32_t id;
    float alpha, gaussval;
    f3 x;  // local hit (scaled by sf)
};

struct RayResult {
    float T_full;
    int n_accepted;
    std::vector<Hit> hits;  // all accepted hits, sorted by (t, id)
};

// shaders.cu:9-75 for one (ray, Gaussian) pair given the packed record.  Returns true when accepted.
inline bool candidate(
    const float *rec, const float *normals, const OrcConfig &cfg, int step, f3 o, f3 d, float tmin, float tmax, Hit &h) {
    if (rec[3] == 0.0f) return false;
    f3 w0 = mk(rec[4], rec[5], rec[6]), w1 = mk(rec[8], rec[9], rec[10]), w2 = mk(rec[12], rec[13], rec[14]);
    f3 mu = mk(rec[0], rec[1], rec[2]);
    f3 p = o - mu;
    f3 ol = mk(dot3(w0, p), dot3(w1, p), dot3(w2, p));
    f3 dl = mk(dot3(w0, d), dot3(w1, d), dot3(w2, d));
    if (dot3(ol, dl) > 0.0f) return false;  // shaders.cu:36
    float nn = sqrtf(dot3(dl, dl));          // shaders.cu:41
    float inv = 1.0f / nn;                   // vec_math.h:337-340
    f3 dh = dl * inv;
    float s = dot3(-ol, dh);                 // shaders.cu:43
    float t = s / nn;                        // shaders.cu:44
    f3 xu = mk(fmaf(s, dh.x, ol.x), fmaf(s, dh.y, ol.y), fmaf(s, dh.z, ol.z));  // shaders.cu:45
    if (dot3(xu, xu) > 1.0f) return false;   // shaders.cu:48-51
    // OptiX presents only instances whose unit box is crossed by the segment [tmin,tmax]
    // (forward_pass.cu:24-43).  The closest-approach point lies inside the box, so the slab test is
    // only needed when t itself is outside the segment.
    if (t < tmin || t > tmax) {
        f3 idl = mk(1.0f / dl.x, 1.0f / dl.y, 1.0f / dl.z);
        float ax = (-1.0f - ol.x) * idl.x, bx = (1.0f - ol.x) * idl.x;
        float ay = (-1.0f - ol.y) * idl.y, by = (1.0f - ol.y) * idl.y;
        float az = (-1.0f - ol.z) * idl.z, bz = (1.0f - ol.z) * idl.z;
        float tnear = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
        float tfar = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
        if (!(tfar >= tmin && tnear <= tmax)) return false;
    }
    uint32_t id;
    std::memcpy(&id, &rec[15], 4);
    if (step != 0 && t < cfg.backfacing_max_dist) {  // shaders.cu:53-61 (object-space dh vs world normal, as written)
        f3 nrm = ld3(normals, id);
        if (sqrtf(dot3(nrm, nrm)) > cfg.backfacing_invalid_normal_threshold && dot3(nrm, dh) > 0.0f) return false;
    }
    float sf = rec[7], sigma = rec[11];
    f3 x = xu * sf;                                             // shaders.cu:64
    float d2 = dot3(x, x);
    float g = det_exp(-det_pow(d2, cfg.exp_power) / (2.0f * cfg.exp_power));  // kernel.cu:8-12
    float alpha = (A_MAX * g) * sigma;                          // kernel.cu:14-16
    h.t = t;
    h.id = (int32_t)id;
    h.alpha = alpha;
    h.gaussval = g;
    h.x = x;
    return true;
}

struct BoundSphere {
    f3 c;
    float r2;
};

void trace_ray(
    const float *recs, const std::vector<BoundSphere> &bs, const float *normals, int n, const OrcConfig &cfg, int step,
    f3 o, f3 d, float tmin, float tmax, RayResult &out, const std::vector<uint32_t> &rank) {
    out.hits.clear();
    out.T_full = 1.0f;
    out.n_accepted = 0;
    // a ray with a non-finite origin / direction is invalid for optixTraverse and finds nothing (the reference's
    // tangent frame, ggx_brdf.h:163, produces such rays when the shading normal is exactly -z)
    if (!std::isfinite(o.x + o.y + o.z + d.x + d.y + d.z) || !(dotp(d, d) > 0.0f)) return;
    for (int i = 0; i < n; ++i) {
        // conservative prefilter (pure accelerator): the hit point lies inside the proxy's bounding sphere
        f3 pc = bs[i].c - o;
        float b = dotp(pc, d);
        float dist2 = dotp(pc, pc) - b * b;
        if (dist2 > bs[i].r2) continue;
        Hit h;
        if (!candidate(recs + 16 * (size_t)i, normals, cfg, step, o, d, tmin, tmax, h)) continue;
        out.n_accepted++;
        out.hits.push_back(h);
    }
    // exact total transmittance (shaders.cu:69-71); the reference multiplies in OptiX traversal order (unspecified),
    // here in Morton order, which is the order the CUDA walk visits leaves in
    std::sort(out.hits.begin(), out.hits.end(), [&](const Hit &a, const Hit &b) { return rank[a.id] < rank[b.id]; });
    for (const Hit &h : out.hits) out.T_full = (float)((double)out.T_full * (1.0 - (double)h.alpha));
    std::sort(out.hits.begin(), out.hits.end(), [&](const Hit &a, const Hit &b) {
        return a.t < b.t || (a.t == b.t && rank[a.id] < rank[b.id]);
    });
}

struct StepState {
    f3 rgb, normal, f0, rem_rgb, rem_normal, rem_f0, ray_o, ray_d, throughput;
    float depth, roughness, T, T_full, rem_depth, rem_roughness;
    std::vector<Hit> comp;        // composited hits in composite order
    std::vector<float> comp_T;    // transmittance after each composited hit
    f3 used_o, used_d;
    bool executed;
};

}  // namespace

extern "C" {

// Full render of one view: forward for all steps, then (grads on) the backward, else accumulate + write.
// `total_num_calls` is the value AFTER the per-call increment (metadata.h:30), i.e. what the kernel reads.
void orc_render(
    const OrcScene *sc, const OrcConfig *cfgp, const OrcCamera *cam, int width, int height, uint32_t total_num_calls,
    int grads_enabled, const OrcTargets *tg, OrcOutputs *out, OrcGrads *gr, OrcHitLists *hl, const OrcOverride *ov,
    int pixel_begin, int pixel_end) {
    const OrcConfig cfg = *cfgp;
    const int n = sc->n;
    const int P = width * height;
    const int S = cfg.max_bounces + 1;
    if (pixel_end <= 0 || pixel_end > P) pixel_end = P;
    if (pixel_begin < 0) pixel_begin = 0;

    std::vector<float> own_recs;
    const float *recs = ov && ov->records ? ov->records : nullptr;
    if (!recs) {
        own_recs.resize((size_t)n * 16);
        orc_prepare(sc, &cfg, own_recs.data());
        recs = own_recs.data();
    }
    std::vector<BoundSphere> bs(n);
    for (int i = 0; i < n; ++i) {
        const float *r = recs + 16 * (size_t)i;
        bs[i].c = mk(r[0], r[1], r[2]);
        if (r[3] == 0.0f) {
            bs[i].r2 = -1.0f;
            continue;
        }
        // rows of W have norm 1/S_r; the unit box maps into a sphere of radius sqrt(3)*max S
        float smax = 0.0f;
        for (int rr = 0; rr < 3; ++rr) {
            const float *wr = r + 4 + 4 * rr;
            float nr = sqrtf(wr[0] * wr[0] + wr[1] * wr[1] + wr[2] * wr[2]);
            smax = fmaxf(smax, 1.0f / nr);
        }
        float rad = 1.7320508f * smax * 1.001f + 1e-6f;
        bs[i].r2 = rad * rad;
    }

    // leaf order: from the caller (the device's own order, e.g. after refits without a rebuild) or recomputed
    std::vector<uint32_t> rank;
    if (ov && ov->leaf_rank) rank.assign(ov->leaf_rank, ov->leaf_rank + n);
    else rank = morton_rank(sc->mean, n);

    std::vector<double> A_rgb, A_normal, A_f0, A_rough, A_opac, A_scale, A_mean, A_rot, A_weight;
    if (grads_enabled) {
        A_rgb.assign((size_t)n * 3, 0), A_normal.assign((size_t)n * 3, 0), A_f0.assign((size_t)n * 3, 0);
        A_rough.assign(n, 0), A_opac.assign(n, 0), A_scale.assign((size_t)n * 3, 0), A_mean.assign((size_t)n * 3, 0);
        A_rot.assign((size_t)n * 4, 0), A_weight.assign(n, 0);
    }
    const int B = std::min(cfg.num_bounces, cfg.max_bounces);  // shaders.cu:85
    const f3 cam_o = mk(cam->origin[0], cam->origin[1], cam->origin[2]);
    // rows of rotation_w2c = columns of rotation_c2w (camera.h:62-68)
    const f3 w2c0 = mk(cam->rot_c2w[0], cam->rot_c2w[3], cam->rot_c2w[6]);
    const f3 w2c1 = mk(cam->rot_c2w[1], cam->rot_c2w[4], cam->rot_c2w[7]);
    const f3 w2c2 = mk(cam->rot_c2w[2], cam->rot_c2w[5], cam->rot_c2w[8]);

#pragma omp parallel
    {
        RayResult rr;
        std::vector<StepState> st(S);
#pragma omp for schedule(dynamic, 16)
        for (int pix = pixel_begin; pix < pixel_end; ++pix) {
            const int ix = pix % width, iy = pix / width;
            uint32_t seed = tea4((uint32_t)pix, total_num_calls);  // shaders.cu:88
            // primary ray (camera.h:17-36)
            float vs_, vc_;
            egr_det_sincos(cam->fovy / 2, &vs_, &vc_);
            float view_size = vs_ / vc_;  // tan(fov/2)
            float aspect = (float)width / (float)height;
            float fx = (float)ix, fy = (float)iy;
            if (cfg.jitter_primary_rays) {
                float jx = rnd(seed) - 0.5f;
                float jy = rnd(seed) - 0.5f;
                fx += jx, fy += jy;
            }
            float y = view_size * (1.0f - 2.0f * (fy + 1.0f / 2.0f) / (float)height);
            float x = aspect * view_size * (2.0f * (fx + 1.0f / 2.0f) / (float)width - 1.0f);
            f3 ray_o = cam_o;
            f3 ray_d = norm3c(fma3(w2c1, y, w2c0 * x) - w2c2);

            for (int s = 0; s < S; ++s) {  // Pixel ctor, framebuffer.h:40-68
                StepState &q = st[s];
                q.rgb = q.normal = q.f0 = q.rem_rgb = q.rem_normal = q.rem_f0 = q.ray_o = q.ray_d = mk(0, 0, 0);
                q.throughput = mk(1, 1, 1);
                q.depth = q.roughness = q.rem_depth = q.rem_roughness = 0.0f;
                q.T = q.T_full = 1.0f;
                q.comp.clear(), q.comp_T.clear();
                q.used_o = q.used_d = mk(0, 0, 0);
                q.executed = false;
            }
            int num_hits[16] = {0};
            int total_effective_steps = 0;
            int last_accumulated = 0, sum_accepted = 0, sum_comp = 0;

            for (int step = 0; step < B + 1; ++step) {  // shaders.cu:104
                if (ov && ov->step_executed && !ov->step_executed[(size_t)step * P + pix]) break;
                if (ov && ov->ray_origin) {
                    ray_o = ld3(ov->ray_origin, (size_t)step * P + pix);
                    ray_d = ld3(ov->ray_direction, (size_t)step * P + pix);
                }
                total_effective_steps++;
                StepState &q = st[step];
                q.executed = true;
                q.used_o = ray_o, q.used_d = ray_d;
                // ---- forward_pass (forward_pass.cu:3-156)
                const float tmin0 = step != 0 ? 0.0f : cam->znear;
                const float tmax = cam->zfar;
                trace_ray(recs, bs, sc->normal, n, cfg, step, ray_o, ray_d, tmin0, tmax, rr, rank);
                q.T = 1.0f;
                q.T_full = rr.T_full;
                sum_accepted += rr.n_accepted;
                float w = tmin0;
                size_t pos = 0;
                int total_acc = 0;
                const size_t H = rr.hits.size();
                for (int it = 0; it < MAX_ITERATIONS && w < tmax; ++it) {
                    // the (up to) 16 nearest hits with t > w, strictly (forward_pass.cu:62)
                    while (pos < H && !(rr.hits[pos].t > w)) ++pos;
                    if (pos >= H) break;  // forward_pass.cu:80-83
                    size_t endb = std::min(H, pos + (size_t)BUFFER_SIZE);
                    bool stop = false;
                    for (size_t i = pos; i < endb; ++i) {
                        const Hit &h = rr.hits[i];
                        w = fmaxf(w, h.t);
                        if (h.t < tmax) {
                            num_hits[step]++;
                            total_acc++;
                            uint32_t id = (uint32_t)h.id;
                            f3 c_rgb = ld3(sc->rgb, id);
                            c_rgb = mk(fmaxf(0.0f, c_rgb.x), fmaxf(0.0f, c_rgb.y), fmaxf(0.0f, c_rgb.z));
                            f3 c_n = ld3(sc->normal, id);
                            f3 c_f0 = ld3(sc->f0, id);
                            c_f0 = mk(fminf(fmaxf(0.0f, c_f0.x), 1.0f), fminf(fmaxf(0.0f, c_f0.y), 1.0f),
                                      fminf(fmaxf(0.0f, c_f0.z), 1.0f));
                            float c_r = fminf(fmaxf(0.0f, sc->roughness[id]), 1.0f);
                            float nextT = q.T * (1.0f - h.alpha);  // forward_pass.cu:108-115
                            float wt = q.T - nextT;
                            q.rgb = fma3(c_rgb, wt, q.rgb);
                            q.normal = fma3(c_n, wt, q.normal);
                            q.f0 = fma3(c_f0, wt, q.f0);
                            q.roughness = fmaf(c_r, wt, q.roughness);
                            q.depth = fmaf(h.t, wt, q.depth);
                            q.T = nextT;
                            q.comp.push_back(h);
                            q.comp_T.push_back(q.T);
                            if (q.T < cfg.transmittance_threshold) {  // forward_pass.cu:131-134
                                w = tmax;
                                stop = true;
                                break;
                            }
                        }
                    }
                    if (stop) break;
                    // a batch with an empty slot drives the window to FLT_MAX through the fill value
                    // (forward_pass.cu:56,89), which ends the loop
                    if (endb - pos < (size_t)BUFFER_SIZE) break;
                    pos = endb;
                }
                last_accumulated = total_acc;
                sum_comp += total_acc;
                // remaining fill (forward_pass.cu:143-155)
                float rem_T = q.T - q.T_full;
                float nrm = fmaxf(1.0f - q.T, cfg.eps_forward_normalization);
                q.rem_rgb = div_s(q.rgb, nrm);
                q.rgb = fma3(q.rem_rgb, rem_T, q.rgb);
                q.rem_depth = q.depth / nrm;
                q.depth = fmaf(rem_T, q.rem_depth, q.depth);
                q.rem_normal = div_s(q.normal, nrm);
                q.normal = fma3(q.rem_normal, rem_T, q.normal);
                q.rem_f0 = div_s(q.f0, nrm);
                q.f0 = fma3(q.rem_f0, rem_T, q.f0);
                q.rem_roughness = q.roughness / nrm;
                q.roughness = fmaf(rem_T, q.rem_roughness, q.roughness);

                if (step > 0) q.rgb = q.rgb * st[step - 1].throughput;  // shaders.cu:112-114
                f3 un = q.normal;
                float er = fmaxf(q.roughness, cfg.eps_min_roughness);
                if (sqrtf(dot3(un, un)) < cfg.reflection_invalid_normal_threshold) break;  // shaders.cu:123-125
                f3 en = norm3c(un);
                f3 epos = fma3(ray_d, q.depth, ray_o);
                // sample_cook_torrance (ggx_brdf.h:152-168); argument order u1 then u2; canonical op order shared
                // with the CUDA kernel (explicit fmaf where nvcc would contract)
                float u1 = rnd(seed);
                float u2 = rnd(seed);
                f3 V = -ray_d;
                float a = er * er;
                float phi = (float)(2.0f * M_PI * (double)u1);
                float cosT = sqrtf((1.0f - u2) / fmaf(a * a - 1.0f, u2, 1.0f));
                float sinT = sqrtf(fmaf(-cosT, cosT, 1.0f));
                float sp, cp;
                egr_det_sincos(phi, &sp, &cp);
                f3 Hl = mk(sinT * cp, sinT * sp, cosT);
                f3 Tn = norm3c(cross3c(en.z < 0.999f ? mk(0, 0, 1) : mk(1, 0, 0), en));
                f3 Bn = cross3c(en, Tn);
                f3 Hh = fma3(en, Hl.z, fma3(Bn, Hl.y, Tn * Hl.x));
                f3 inc = -V;
                f3 next_d = fma3(Hh, -2.0f * dot3(Hh, inc), inc);  // reflect(-V,H), vec_math.h:385
                f3 next_o = fma3(next_d, cfg.eps_ray_surface_offset, epos);
                f3 ef0 = q.f0;
                if (step > 0) q.throughput = st[step - 1].throughput;
                // cook_torrance_weight (ggx_brdf.h:134-150)
                f3 wgt = mk(0, 0, 0);
                if (!(ef0.x == 0.0f && ef0.y == 0.0f && ef0.z == 0.0f)) {
                    f3 L = next_d;
                    f3 Hn = norm3c(V + L);
                    float NdotH = fmaxf(dot3(en, Hn), 0.0f);
                    float VdotH = fmaxf(dot3(V, Hn), 0.0f);
                    float NdotV = fmaxf(dot3(en, V), 0.0f);
                    float kk = (a * a) / 2.0f;
                    float NdotL = fmaxf(dot3(en, L), 0.0f);
                    float g1v = (float)((double)NdotV / ((double)fmaf(NdotV, 1.0f - kk, kk) + 1e-8));
                    float g1l = (float)((double)NdotL / ((double)fmaf(NdotL, 1.0f - kk, kk) + 1e-8));
                    float G = g1v * g1l;
                    float pw = 1.0f - VdotH;
                    float p5 = pw * pw;
                    p5 = (p5 * p5) * pw;
                    f3 F = fma3(mk(1, 1, 1) - ef0, p5, ef0);
                    // float3 / double resolves to operator/(float3, float): the denominator is rounded to
                    // float and the numerator multiplied by its reciprocal (vec_math.h:330-333)
                    float den = (float)((double)(NdotH * NdotV) + 1e-8);
                    wgt = div_s((F * G) * VdotH, den);
                }
                q.throughput = q.throughput * wgt;
                ray_o = next_o, ray_d = next_d;
                q.ray_o = ray_o, q.ray_d = ray_d;
            }
            f3 fin = mk(0, 0, 0);
            for (int step = 0; step < B + 1; ++step) fin = fin + st[step].rgb;  // shaders.cu:150-152

            if (out->num_accumulated) out->num_accumulated[pix] = last_accumulated;
            if (out->num_hits_per_pixel) out->num_hits_per_pixel[pix] = sum_accepted;
            if (out->num_composited_per_pixel) out->num_composited_per_pixel[pix] = sum_comp;
            if (out->num_steps_per_pixel) out->num_steps_per_pixel[pix] = total_effective_steps;
            for (int s = 0; s < S; ++s) {
                size_t o_ = (size_t)s * P + pix;
                if (out->used_ray_origin) st3(out->used_ray_origin, o_, st[s].used_o);
                if (out->used_ray_direction) st3(out->used_ray_direction, o_, st[s].used_d);
                if (out->step_executed) out->step_executed[o_] = st[s].executed ? 1 : 0;
                if (hl && hl->cap > 0) {
                    hl->counts[o_] = (int32_t)st[s].comp.size();
                    for (int k = 0; k < hl->cap; ++k) {
                        bool have = k < (int)st[s].comp.size();
                        hl->ids[o_ * hl->cap + k] = have ? st[s].comp[k].id : -1;
                        if (hl->ts) hl->ts[o_ * hl->cap + k] = have ? st[s].comp[k].t : 0.0f;
                    }
                }
            }

            if (grads_enabled) {
                // ---- backward (shaders.cu:155-162, backward_pass.cu:3-222)
                f3 t_diff = ld3(tg->diffuse, pix), t_spec = ld3(tg->specular, pix), t_n = ld3(tg->normal, pix),
                   t_f0 = ld3(tg->f0, pix);
                float t_depth = tg->depth[pix], t_rough = tg->roughness[pix];
                const float kexp = cfg.exp_power;
                for (int step = total_effective_steps - 1; step >= 0; --step) {
                    if (num_hits[step] <= 0) continue;
                    StepState &q = st[step];
                    f3 thr = step == 0 ? mk(1, 1, 1) : st[step - 1].throughput;
                    float rdw = powf(1.0f - st[std::max(step - 1, 0)].roughness, ROUGH_DOWNWEIGHT_POW);
                    f3 d_rgb = mk(0, 0, 0), d_n = mk(0, 0, 0), d_f0 = mk(0, 0, 0);
                    float d_depth = 0.0f, d_rough = 0.0f;
                    if (step == 0) {  // backward_pass.cu:89-99
                        d_rgb = 1.0f / 3.0f * sign3(st[0].rgb - t_diff) * cfg.loss_weight_diffuse;
                        d_depth = sign1(st[0].depth - t_depth) * cfg.loss_weight_depth;
                        d_n = 1.0f / 3.0f * sign3(st[0].normal - t_n) * cfg.loss_weight_normal;
                        d_f0 = 1.0f / 3.0f * sign3(st[0].f0 - t_f0) * cfg.loss_weight_f0;
                        d_rough = sign1(st[0].roughness - t_rough) * cfg.loss_weight_roughness;
                    } else {  // backward_pass.cu:100-108
                        f3 spec = mk(0, 0, 0);
                        for (int j = 1; j < B + 1; ++j) spec = spec + st[j].rgb;
                        d_rgb = 1.0f / 3.0f * sign3(spec - t_spec) * cfg.loss_weight_specular * rdw;
                        d_rgb = d_rgb * thr;
                    }
                    f3 p_rgb = mk(0, 0, 0), w_rgb = mk(0, 0, 0), p_n = mk(0, 0, 0), w_n = mk(0, 0, 0),
                       p_f0 = mk(0, 0, 0), w_f0 = mk(0, 0, 0);
                    float p_r = 0, w_r = 0, p_d = 0, w_d = 0;
                    for (int hi = (int)q.comp.size() - 1; hi >= 0; --hi) {  // farthest -> nearest
                        const Hit &h = q.comp[hi];
                        const uint32_t id = (uint32_t)h.id;
                        const float Tr = q.comp_T[hi];
                        const float alpha = h.alpha, gv = h.gaussval;
                        f3 c_rgb = ld3(sc->rgb, id);
                        c_rgb = mk(fmaxf(0.0f, c_rgb.x), fmaxf(0.0f, c_rgb.y), fmaxf(0.0f, c_rgb.z));
                        f3 c_n = mk(0, 0, 0), c_f0 = mk(0, 0, 0);
                        float c_r = 0.0f;
                        if (step == 0) {
                            c_n = ld3(sc->normal, id);
                            c_f0 = ld3(sc->f0, id);
                            c_f0 = mk(fminf(fmaxf(0.0f, c_f0.x), 1.0f), fminf(fmaxf(0.0f, c_f0.y), 1.0f),
                                      fminf(fmaxf(0.0f, c_f0.z), 1.0f));
                            c_r = fminf(fmaxf(0.0f, sc->roughness[id]), 1.0f);
                        }
                        const float *rec = recs + 16 * (size_t)id;
                        float sigma = rec[11], sf = rec[7];
                        f3 sc_e = mk(expf(sc->scale[3 * id]), expf(sc->scale[3 * id + 1]), expf(sc->scale[3 * id + 2]));
                        float qr = sc->rotation[4 * id], qx = sc->rotation[4 * id + 1], qy = sc->rotation[4 * id + 2],
                              qz = sc->rotation[4 * id + 3];
                        float qn = sqrtf(qr * qr + qx * qx + qy * qy + qz * qz);
                        float r = qr / qn, xq = qx / qn, yq = qy / qn, zq = qz / qn;
                        // colour gradients (backward_pass.cu:110-115); activations' backward is always open
                        // because it is evaluated on the activated value (activations.cu:42,54-56)
                        float weight = (float)((double)Tr / (1.0 - (double)alpha) * (double)alpha);
                        f3 g_c = d_rgb * weight;
                        f3 g_nn = d_n * weight;
                        f3 g_ff = d_f0 * weight;
                        float g_rr = d_rough * weight;
                        w_rgb = w_rgb + (c_rgb - p_rgb) * Tr, p_rgb = c_rgb;
                        if (step == 0) {
                            w_n = w_n + (c_n - p_n) * Tr, p_n = c_n;
                            w_f0 = w_f0 + (c_f0 - p_f0) * Tr, p_f0 = c_f0;
                            w_r += (c_r - p_r) * Tr, p_r = c_r;
                            w_d += (h.t - p_d) * Tr, p_d = h.t;
                        }
                        float dL_da = 0.0f;
                        float tmp1 = 1.0f / (1.0f - alpha);
                        dL_da += dotp(w_rgb * tmp1, d_rgb);
                        dL_da += dotp(w_n * tmp1, d_n);
                        dL_da += dotp(w_f0 * tmp1, d_f0);
                        dL_da += (w_r * tmp1) * d_rough;
                        dL_da += (w_d * tmp1) * d_depth;
                        float tmp2 = -((q.T - q.T_full) / (1.0f - alpha));
                        dL_da += tmp2 * dotp(q.rem_rgb, d_rgb);
                        dL_da += tmp2 * dotp(q.rem_normal, d_n);
                        dL_da += tmp2 * dotp(q.rem_f0, d_f0);
                        dL_da += tmp2 * (q.rem_roughness * d_rough);
                        dL_da += tmp2 * (q.rem_depth * d_depth);
                        float g_op = A_MAX * dL_da * gv;
                        g_op = g_op * sigma * (1.0f - sigma);  // activations.cu:8
                        float dL_dg = A_MAX * dL_da * sigma;
                        float sqn = dotp(h.x, h.x);
                        float dsq = gv * det_pow(sqn, kexp - 1.0f);  // backward_pass.cu:157
                        f3 dxl = -h.x * dsq * dL_dg;
                        // world_to_local columns dotted with dxl (backward_pass.cu:160-167)
                        f3 w0 = mk(rec[4], rec[5], rec[6]), w1 = mk(rec[8], rec[9], rec[10]), w2 = mk(rec[12], rec[13], rec[14]);
                        f3 dxw = mk(dotp(mk(w0.x, w1.x, w2.x), dxl), dotp(mk(w0.y, w1.y, w2.y), dxl),
                                    dotp(mk(w0.z, w1.z, w2.z), dxl)) *
                                 sf;
                        f3 l0 = -dxw.x * h.x, l1 = -dxw.y * h.x, l2 = -dxw.z * h.x;
                        f3 g_mean = -dxw;
                        // local_to_world rows: M[r][c] = R[r][c] * s_c * sf * global (bvh_wrapper.cu:25-30)
                        float Rm[3][3] = {
                            {1.f - 2.f * (yq * yq + zq * zq), 2.f * (xq * yq - r * zq), 2.f * (xq * zq + r * yq)},
                            {2.f * (xq * yq + r * zq), 1.f - 2.f * (xq * xq + zq * zq), 2.f * (yq * zq - r * xq)},
                            {2.f * (xq * zq - r * yq), 2.f * (yq * zq + r * xq), 1.f - 2.f * (xq * xq + yq * yq)}};
                        f3 Sg = sc_e * sf * cfg.global_scale_factor;
                        f3 den = sc_e * sf + mk(cfg.eps_scale_grad, cfg.eps_scale_grad, cfg.eps_scale_grad);
                        f3 rot0 = mk(Rm[0][0] * Sg.x, Rm[0][1] * Sg.y, Rm[0][2] * Sg.z) / den;
                        f3 rot1 = mk(Rm[1][0] * Sg.x, Rm[1][1] * Sg.y, Rm[1][2] * Sg.z) / den;
                        f3 rot2 = mk(Rm[2][0] * Sg.x, Rm[2][1] * Sg.y, Rm[2][2] * Sg.z) / den;
                        f3 g_scale = (l0 * rot0 + l1 * rot1 + l2 * rot2) * sc_e;  // backward_exp_act
                        f3 dr0 = l0 * sc_e, dr1 = l1 * sc_e, dr2 = l2 * sc_e;
                        float dLr = 2.f * xq * (dr2.y - dr1.z) + 2.f * yq * (dr0.z - dr2.x) + 2.f * zq * (dr1.x - dr0.y);
                        float dLx = -4.f * xq * (dr1.y + dr2.z) + 2.f * yq * (dr0.y + dr1.x) + 2.f * zq * (dr0.z + dr2.x) +
                                    2.f * r * (dr2.y - dr1.z);
                        float dLy = 2.f * xq * (dr0.y + dr1.x) - 4.f * yq * (dr0.x + dr2.z) + 2.f * zq * (dr1.z + dr2.y) +
                                    2.f * r * (dr0.z - dr2.x);
                        float dLz = 2.f * xq * (dr0.z + dr2.x) + 2.f * yq * (dr1.z + dr2.y) - 4.f * zq * (dr0.x + dr1.y) +
                                    2.f * r * (dr1.x - dr0.y);
                        // backward_normalize_act (activations.cu:71-73): dot(dL,x)*(-x)/|x|^3 + dL/|x|
                        float dq = dLr * qr + dLx * qx + dLy * qy + dLz * qz;
                        float i3 = 1.0f / powf(qn, 3.0f), i1 = 1.0f / qn;  // float4 / float = times reciprocal
                        float gq[4] = {(dq * -qr) * i3 + dLr * i1, (dq * -qx) * i3 + dLx * i1, (dq * -qy) * i3 + dLy * i1,
                                       (dq * -qz) * i3 + dLz * i1};
#pragma omp atomic
                        A_opac[id] += g_op;
                        float gs3[3] = {g_scale.x, g_scale.y, g_scale.z}, gm3[3] = {g_mean.x, g_mean.y, g_mean.z},
                              gc3[3] = {g_c.x, g_c.y, g_c.z}, gn3[3] = {g_nn.x, g_nn.y, g_nn.z},
                              gf3[3] = {g_ff.x, g_ff.y, g_ff.z};
                        for (int c = 0; c < 3; ++c) {
#pragma omp atomic
                            A_scale[3 * (size_t)id + c] += gs3[c];
#pragma omp atomic
                            A_mean[3 * (size_t)id + c] += gm3[c];
#pragma omp atomic
                            A_rgb[3 * (size_t)id + c] += gc3[c];
                        }
                        for (int c = 0; c < 4; ++c) {
#pragma omp atomic
                            A_rot[4 * (size_t)id + c] += gq[c];
                        }
                        if (step == 0) {
                            for (int c = 0; c < 3; ++c) {
#pragma omp atomic
                                A_normal[3 * (size_t)id + c] += gn3[c];
#pragma omp atomic
                                A_f0[3 * (size_t)id + c] += gf3[c];
                            }
#pragma omp atomic
                            A_rough[id] += g_rr;
                        }
#pragma omp atomic
                        A_weight[id] += weight;
                    }
                }
            } else {
                // ---- accumulate + write (shaders.cu:163-169, framebuffer.h:104-143), all S steps
                if (cfg.accumulate_samples) {
                    int cnt = *out->acc_count;
                    float dv = (float)(cnt + 1);
                    fin = mk(0, 0, 0);
                    for (int s = 0; s < S; ++s) {
                        size_t o_ = (size_t)pix + (size_t)P * s;
                        StepState &q = st[s];
                        f3 a3 = ld3(out->acc_rgb, o_) + q.rgb;
                        st3(out->acc_rgb, o_, a3), q.rgb = div_s(a3, dv);
                        out->acc_transmittance[o_] += q.T, q.T = out->acc_transmittance[o_] / dv;
                        out->acc_total_transmittance[o_] += q.T_full, q.T_full = out->acc_total_transmittance[o_] / dv;
                        out->acc_depth[o_] += q.depth, q.depth = out->acc_depth[o_] / dv;
                        a3 = ld3(out->acc_normal, o_) + q.normal;
                        st3(out->acc_normal, o_, a3), q.normal = div_s(a3, dv);
                        a3 = ld3(out->acc_f0, o_) + q.f0;
                        st3(out->acc_f0, o_, a3), q.f0 = div_s(a3, dv);
                        out->acc_roughness[o_] += q.roughness, q.roughness = out->acc_roughness[o_] / dv;
                        fin = fin + q.rgb;
                    }
                }
                for (int s = 0; s < S; ++s) {
                    size_t o_ = (size_t)pix + (size_t)P * s;
                    const StepState &q = st[s];
                    if (out->rgb) st3(out->rgb, o_, q.rgb);
                    if (out->depth) out->depth[o_] = q.depth;
                    if (out->normal) st3(out->normal, o_, q.normal);
                    if (out->f0) st3(out->f0, o_, q.f0);
                    if (out->roughness) out->roughness[o_] = q.roughness;
                    if (out->transmittance) out->transmittance[o_] = q.T;
                    if (out->total_transmittance) out->total_transmittance[o_] = q.T_full;
                    if (out->ray_origin) st3(out->ray_origin, o_, q.ray_o);
                    if (out->ray_direction) st3(out->ray_direction, o_, q.ray_d);
                }
                if (out->final_) st3(out->final_, pix, fin);
            }
            if (out->random_seeds) out->random_seeds[pix] = seed;  // shaders.cu:172
        }
    }
    if (grads_enabled) {
        for (size_t i = 0; i < (size_t)n; ++i) {
            for (int c = 0; c < 3; ++c) {
                gr->dL_drgb[3 * i + c] += (float)A_rgb[3 * i + c];
                gr->dL_dnormal[3 * i + c] += (float)A_normal[3 * i + c];
                gr->dL_df0[3 * i + c] += (float)A_f0[3 * i + c];
                gr->dL_dscale[3 * i + c] += (float)A_scale[3 * i + c];
                gr->dL_dmean[3 * i + c] += (float)A_mean[3 * i + c];
            }
            for (int c = 0; c < 4; ++c) gr->dL_drotation[4 * i + c] += (float)A_rot[4 * i + c];
            gr->dL_droughness[i] += (float)A_rough[i];
            gr->dL_dopacity[i] += (float)A_opac[i];
            gr->total_weight[i] += (float)A_weight[i];
        }
    }
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------
// Row f3: mean squared distance to the three nearest neighbours, restating simple-knn's definition
// (submodules/simple-knn/simple_knn.cu:141-199): for point i the three smallest squared distances to the OTHER
// points (index != i; duplicates count), kept sorted ascending by the swap cascade of `updateKBest` (:141-154), then
// (best[0] + best[1] + best[2]) / 3.0f (:198).  Brute force over all pairs (the reference's box pruning only skips
// points that cannot enter the best three).  The squared distance is the fused form nvcc contracts the reference's
// `d.x*d.x + d.y*d.y + d.z*d.z` into (SASS of the sm_100 build: FMUL y, FFMA x, FFMA z): fma(dz,dz,fma(dx,dx,dy*dy)) -- fmaf() is correctly rounded, -ffp-contract=off
// keeps everything else as written.
// ---------------------------------------------------------------------------------------------------------
extern "C" void orc_knn3_mean_dist2(const float *points, int n, float *out) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int i = 0; i < n; ++i) {
        const float px = points[3 * i], py = points[3 * i + 1], pz = points[3 * i + 2];
        float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
        for (int j = 0; j < n; ++j) {
            if (j == i) continue;
            const float dx = points[3 * j] - px, dy = points[3 * j + 1] - py, dz = points[3 * j + 2] - pz;
            float dist = fmaf(dz, dz, fmaf(dx, dx, dy * dy));
            for (int k = 0; k < 3; ++k) {  // updateKBest<3>
                if (best[k] > dist) {
                    const float t = best[k];
                    best[k] = dist;
                    dist = t;
                }
            }
        }
        out[i] = ((best[0] + best[1]) + best[2]) / 3.0f;
    }
}
