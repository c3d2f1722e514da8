/*
 * egr_b200.h — C ABI of the B200-native differentiable Gaussian ray tracer.
 *
 * Drop-in boundary for the hot path of graphdeco-inria/editable-gaussian-reflections: everything the
 * TorchScript class `raytracer.Raytracer` (reference: editable_gauss_refl/cuda/csrc/raytracer.cpp:24-218)
 * does on the device is reachable through the entry points below.  Plain pointers and sizes only: no
 * torch, no C++ types.  Every pointer is a DEVICE pointer owned by the caller unless stated otherwise; the
 * library allocates nothing (all storage / scratch is sized by the *_bytes() queries and passed in), never
 * synchronises the host, and enqueues all work on the stream handed in (`void*` = cudaStream_t).
 *
 * Every function returns 0 on success or a negative EGR_E* code; egr_last_error() returns the message of the
 * last failure on the calling thread (the Python shim raises RuntimeError from it, preserving the reference's
 * exception convention, csrc/utils/exception.h:48-78).
 *
 * The structs mirror the reference's raw-pointer device structs one to one:
 *   EgrCamera      <- struct Camera       csrc/core/camera.h:8-15
 *   EgrConfig      <- struct Config       csrc/core/config.h:5-26
 *   EgrFramebuffer <- struct Framebuffer  csrc/core/framebuffer.h:72-102
 *   EgrGaussians   <- struct Gaussians    csrc/core/gaussians.h:3-25
 *   EgrMetadata    <- struct Metadata     csrc/core/metadata.h:3-7
 *   EgrStats       <- struct Stats        csrc/core/stats.h:3-6
 *   EgrParams      <- struct Params       csrc/params.h:11-26
 */
#ifndef EGR_B200_H
#define EGR_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EGR_ABI_VERSION 2
#define EGR_MAX_STEPS 8          /* upper bound on MAX_BOUNCES+1 (reference: flags.h:5 fixes MAX_BOUNCES=2) */
#define EGR_GRAD_STRIDE 24       /* floats per Gaussian in the flat gradient record, see EgrGaussians */
#define EGR_RECORD_FLOATS 16     /* floats per packed geometry record */

#define EGR_OK 0
#define EGR_EINVAL (-1)          /* bad argument (null pointer, size mismatch, too many steps ...) */
#define EGR_ENOMEM (-2)          /* a caller-provided buffer is smaller than the matching *_bytes() query */
#define EGR_ECUDA (-3)           /* a CUDA call failed; message holds cudaGetErrorString */

typedef struct EgrCamera {       /* csrc/core/camera.h:8-15 */
    const float *origin;                 /* [3]   */
    const float *vertical_fov_radians;   /* [1]   */
    const float *rotation_c2w;           /* [3,3] row-major */
    const float *rotation_w2c;           /* [3,3] row-major (= transpose) */
    const float *znear;                  /* [1]   */
    const float *zfar;                   /* [1]   */
} EgrCamera;

typedef struct EgrConfig {       /* csrc/core/config.h:5-26; one-element device scalars, read on the device */
    const float *exp_power;
    const float *alpha_threshold;
    const float *transmittance_threshold;
    const uint8_t *accumulate_samples;   /* bool */
    const uint8_t *jitter_primary_rays;  /* bool */
    const int32_t *num_bounces;
    const float *global_scale_factor;
    const float *loss_weight_diffuse;
    const float *loss_weight_specular;
    const float *loss_weight_depth;
    const float *loss_weight_normal;
    const float *loss_weight_f0;
    const float *loss_weight_roughness;
    const float *eps_forward_normalization;
    const float *eps_scale_grad;
    const float *eps_ray_surface_offset;
    const float *eps_min_roughness;
    const float *reflection_invalid_normal_threshold;
    const float *backfacing_invalid_normal_threshold;
    const float *backfacing_max_dist;
} EgrConfig;

typedef struct EgrFramebuffer {  /* csrc/core/framebuffer.h:72-102; planes are [S,H,W,C], step-major */
    float *output_rgb, *output_depth, *output_normal, *output_f0, *output_roughness, *output_transmittance,
        *output_total_transmittance, *output_ray_origin, *output_ray_direction;
    float *output_final, *output_denoised;          /* [1,H,W,3] */
    float *accumulated_rgb, *accumulated_transmittance, *accumulated_total_transmittance, *accumulated_depth,
        *accumulated_normal, *accumulated_f0, *accumulated_roughness;
    int32_t *accumulated_sample_count;              /* [1] */
    const float *target_diffuse, *target_specular, *target_depth, *target_normal, *target_f0,
        *target_roughness;                          /* [H,W,C] */
} EgrFramebuffer;

typedef struct EgrGaussians {    /* csrc/core/gaussians.h:3-25; raw (pre-activation) parameters */
    uint32_t count;
    const float *rgb, *normal, *f0;      /* [N,3] */
    const float *roughness, *opacity;    /* [N,1] */
    const float *scale, *mean;           /* [N,3] */
    const float *rotation;               /* [N,4] (r,x,y,z) */
    /* The reference's eight dL_d* arrays + total_weight are ONE flat [N, EGR_GRAD_STRIDE] fp32 record here so
     * that a hit is flushed with 128-bit vector reductions and a multi-GPU job all-reduces one buffer:
     *   0-2 dL_dmean   3 dL_dopacity | 4-6 dL_dscale  7 dL_droughness | 8-11 dL_drotation |
     *   12-14 dL_drgb 15 total_weight | 16-18 dL_dnormal 19 pad | 20-22 dL_df0 23 pad
     * (the Python binding exposes the reference's tensors as strided views of it).  16-byte aligned. */
    float *grad_records;
} EgrGaussians;

typedef struct EgrMetadata {     /* csrc/core/metadata.h:3-7 */
    uint8_t *grads_enabled;      /* [1] bool, written by egr_raytrace */
    uint32_t *total_num_calls;   /* [1], incremented by egr_raytrace before tracing (metadata.h:28-31) */
    uint32_t *random_seeds;      /* [H,W,1] */
} EgrMetadata;

typedef struct EgrStats {        /* csrc/core/stats.h:3-6 + three extra planes used by the roofline formula */
    int32_t *num_accumulated_per_pixel;  /* [H,W] composited hits of the last executed step */
    int32_t *num_traversed_per_pixel;    /* [H,W] candidates tested (implementation-defined, as in OptiX) */
    int32_t *num_hits_per_pixel;         /* [H,W] accepted hits, summed over steps */
    int32_t *num_composited_per_pixel;   /* [H,W] composited hits, summed over steps */
    int32_t *num_steps_per_pixel;        /* [H,W] executed ray-steps */
    uint32_t *hitlog_status;             /* [4]: [0] hit-log chunks used by the last grad-mode call, [1] 1 when that call
                                          * ran out of hit-log space (the affected 32-ray groups then get NO backward:
                                          * their gradients are missing), [2] sticky copy of [1] (set by the library,
                                          * cleared only by the caller; Raytracer.hitlog_overflowed() reads it), [3] unused */
} EgrStats;

typedef struct EgrParams {       /* csrc/params.h:11-26 */
    uint32_t image_width, image_height;
    uint32_t num_steps;          /* S = MAX_BOUNCES+1 the planes were allocated with (<= EGR_MAX_STEPS) */
    uint32_t tile_rows;          /* 0 = the whole image; else (first) | (end << 16) in units of 4-pixel tile rows: only the
                                  * pixels of rows [4*first, 4*end) are traced (their bounce rays included).  Used to
                                  * split ONE view over several GPUs by horizontal bands: per-pixel seeds do not depend
                                  * on the band, so the bands of all ranks tile the single-GPU image bit for bit and the
                                  * summed gradient records equal the single-GPU gradients (float-atomic order aside).
                                  * Outside the band the outputs hold the values of a pixel without rays. */
    EgrCamera camera;
    EgrConfig config;
    EgrFramebuffer framebuffer;
    EgrGaussians gaussians;
    EgrMetadata metadata;
    EgrStats stats;
    void *bvh_storage;           /* built by egr_bvh_build, sized by egr_bvh_storage_bytes(count) */
    size_t bvh_storage_bytes;
    void *pixel_state;           /* inter-step per-pixel state, sized by egr_pixel_state_bytes */
    size_t pixel_state_bytes;
    void *hitlog;                /* composited-hit log read by the backward; may be NULL when grads are off */
    size_t hitlog_bytes;
    /* debug planes, may be NULL: the ray traced at each step and whether the step ran ([S,H,W,3] / [S,H,W]) */
    float *used_ray_origin, *used_ray_direction;
    int32_t *step_executed;
    /* test inputs, may be NULL ([S,H,W,3] / [S,H,W]): when set, the ray of every step >= 1 is read from these planes
     * instead of the ray spawned by the previous step, and step s+1 of a pixel runs iff override_step_executed[s+1] != 0.
     * Lets a test feed the bounce rays logged by another implementation (the reference), so that ulp-level
     * differences in a reflected ray are not amplified by the next step. */
    const float *override_ray_origin, *override_ray_direction;
    const int32_t *override_step_executed;
    /* 0 (default): exact mode -- every hit along a ray is enumerated, total transmittance and hit lists are the
     * reference's.  > 0 (opt-in, no counterpart in the reference): early-stop mode -- hits are collected nearest first
     * and a ray stops collecting once its composite has terminated AND the running total transmittance is below this
     * threshold; output_total_transmittance is then an upper bound within the threshold of the exact value, everything
     * composited is unchanged (ties at exactly equal distance may be ordered differently).  For opaque scenes. */
    float early_stop_transmittance;
    uint32_t reserved1;
} EgrParams;

/* ---- library ------------------------------------------------------------------------------------------ */
int egr_abi_version(void);
const char *egr_last_error(void);

/* Compile-time constants the reference exports as static methods (csrc/raytracer.cpp:145-148, flags.h). */
double egr_max_alpha(void);
int egr_roughness_downweight_grad(void);
double egr_roughness_downweight_grad_power(void);
int egr_buffer_size(void);      /* flags.h:16 */
int egr_max_iterations(void);   /* flags.h:17 */

/* ---- BVH (replaces csrc/optix/bvh_wrapper.{h,cu}) ------------------------------------------------------ */
size_t egr_bvh_storage_bytes(uint32_t num_gaussians);
size_t egr_bvh_scratch_bytes(uint32_t num_gaussians);
/* Full build: Morton codes of the means, radix sort, Karras topology, preorder layout, then a refit.
 * Replaces BVHWrapper::rebuild / build_tlas (bvh_wrapper.h:24-30,118-157). */
int egr_bvh_build(const EgrGaussians *g, const EgrConfig *cfg, void *storage, size_t storage_bytes, void *scratch,
                  size_t scratch_bytes, void *stream);
/* Refit: re-pack the geometry records from the current parameters and refresh all boxes bottom-up.
 * Replaces BVHWrapper::update = populateBVH + optixAccelBuild(UPDATE) (bvh_wrapper.h:32-59,
 * bvh_wrapper.cu:33-70) without its three host syncs. */
int egr_bvh_refit(const EgrGaussians *g, const EgrConfig *cfg, void *storage, size_t storage_bytes, void *stream);

/* No counterpart in the reference (its rebuild calls cudaFree + cudaMalloc three times, bvh_wrapper.h:24-30).
 * Same result as egr_bvh_build, for hosts that must not allocate device memory when a rebuild is due: e.g. an
 * application that has handed the whole GPU to its own memory pool (a raw allocation then fails, or forces the
 * pool to flush and synchronise), or an interactive edit -> re-render loop where the driver-wide lock and implicit
 * synchronisation of an allocation would show up as a hitch.  The build scratch is a 6 MiB block inside the
 * library image instead of a caller buffer, which caps the size at EGR_NOALLOC_MAX_GAUSSIANS; larger counts are
 * rejected with EGR_EINVAL and a message.  The caller still passes the parameters and the BVH storage.
 *  - The first call on each GPU looks up the block's device address (with that GPU current); make that first call
 *    outside any stream capture.
 *  - All calls on a GPU share the one block (it simply keeps whatever the previous call left; nothing in it needs
 *    clearing).  Two builds that overlap on the same GPU (different streams or threads) would corrupt each other:
 *    serialise them.
 *  - Allocates nothing, directly or through a library. */
#define EGR_NOALLOC_MAX_GAUSSIANS 65536u
int egr_bvh_build_noalloc(const EgrGaussians *g, const EgrConfig *cfg, void *storage, size_t storage_bytes,
                          void *stream);

/* ---- tracer (replaces Raytracer::raytrace, csrc/raytracer.cpp:81-94, and the OptiX launch of
 *      __raygen__rg / __intersection__gaussian, csrc/shaders.cu:9-173) -------------------------------- */
size_t egr_pixel_state_bytes(uint32_t width, uint32_t height, uint32_t num_steps);
size_t egr_hitlog_bytes_for_hits(uint64_t num_hits_capacity);
/* One raytrace() call: metadata update, stats reset, forward for all steps and then either the fused-loss
 * backward (grads_enabled != 0; nothing is written to the framebuffer, shaders.cu:155-169) or accumulation +
 * output writes, and the accumulated_sample_count bump. */
int egr_raytrace(const EgrParams *p, int grads_enabled, void *stream);
/* The same call split into its launches so that a caller can bracket each kernel with events (bench.py times the
 * dominant kernel this way): BEGIN = appearance pack + metadata/stats reset, STEP(s) = forward of ray-step s,
 * FINISH = backward (grads) or resolve/write (no grads) + sample-count bump.  egr_raytrace == EGR_PHASE_ALL. */
#define EGR_PHASE_BEGIN 0x1u
#define EGR_PHASE_STEP(s) (0x2u << (s))
#define EGR_PHASE_FINISH 0x10000u
#define EGR_PHASE_ALL 0xFFFFFFFFu
int egr_raytrace_phases(const EgrParams *p, int grads_enabled, uint32_t phase_mask, void *stream);

/* CameraDataHolder::set_pose (csrc/core/camera.h:43-77) for a pose that lives in HOST memory: origin[3] and the
 * row-major c2w rotation[9] are read during this call and travel as kernel arguments of one tiny launch that writes
 * origin, rotation_c2w and rotation_w2c (= transpose).  No staging buffer, no host synchronisation (the reference's
 * `tensor.copy_` of a pageable host tensor blocks the host until every earlier launch on the stream has finished).
 * Either pointer may be NULL to leave that part unchanged. */
int egr_camera_set_pose(const EgrCamera *cam, const float *host_origin, const float *host_rotation_c2w, void *stream);

/* GaussianRaytracer._import_param_gradients (editable_gauss_refl/renderer/gaussian_raytracer.py:53-62: eight
 * `param.grad.add_(holder.grad)`) as ONE launch: dst[k] += column group k of the flat gradient record, each dst a
 * contiguous [N, ch] fp32 array of the caller (the model's .grad tensors); a NULL dst is skipped.  Reads the record
 * once (96 B/Gaussian) instead of eight strided passes over it. */
typedef struct EgrGradTargets {
    float *mean, *opacity, *scale, *roughness, *rotation, *rgb, *normal, *f0; /* [N,3] [N,1] [N,3] [N,1] [N,4] [N,3] [N,3] [N,3] */
} EgrGradTargets;
int egr_gradients_accumulate(const float *grad_records, uint32_t count, const EgrGradTargets *dst, void *stream);
/* GaussianRaytracer._export_param_values (renderer/gaussian_raytracer.py:41-51: eight `holder.copy_(model tensor)`)
 * as ONE launch: dst[k] = src[k] for the eight contiguous [N, ch] fp32 arrays (a NULL pair is skipped). */
int egr_params_export(const EgrGradTargets *src, const EgrGradTargets *dst, uint32_t count, void *stream);

/* FramebufferDataHolder::reset_accumulators (csrc/core/framebuffer.h:249-258): eight stream-ordered memsets. */
int egr_reset_accumulators(const EgrParams *p, void *stream);
/* DenoiserWrapper::run (csrc/optix/denoiser_wrapper.h:76-104): the OptiX AI denoiser is a proprietary
 * driver-resident network, unavailable here; this is the documented pass-through output_final -> output_denoised. */
int egr_denoise_passthrough(const EgrParams *p, void *stream);
/* Headless stand-in for DenoiserWrapper::run, used by Raytracer.denoise(): an edge-avoiding a-trous wavelet filter
 * (Dammertz et al. 2010) of output_final into output_denoised, guided by the primary-step normal (the guide layer the
 * reference gives OptiX, denoiser_wrapper.h:88-96) and the primary-step depth.  EGR_DENOISE_PASSES passes of a 5x5
 * B3-spline kernel with hole sizes 1, 2, 4 ...; tap weight = kernel * exp(-(|dc|^2/sigma_c^2 + |dn|^2/sigma_n^2 +
 * (dd/(|d|+eps))^2/sigma_d^2)), sigma_c halved after every pass.  It is NOT the OptiX network (which is unavailable
 * on this box); it keeps geometric edges and removes the Monte-Carlo noise of the sampled reflection directions.
 * Scratch: six step-0 planes of pixel_state that are dead between two egr_raytrace calls (call it after the trace
 * whose output it filters, and before egr_hitlog_export's of an earlier trace are no longer needed). */
#define EGR_DENOISE_PASSES 3
#define EGR_DENOISE_SIGMA_COLOR 0.6f
#define EGR_DENOISE_SIGMA_NORMAL 0.3f
#define EGR_DENOISE_SIGMA_DEPTH 0.05f
#define EGR_DENOISE_DEPTH_EPS 1.0e-3f
int egr_denoise(const EgrParams *p, void *stream);

/* ---- next row f3: `distCUDA2` of the reference's simple-knn submodule
 *      (submodules/simple-knn/simple_knn.cu:46-222, spatial.cu; Python call sites scene/gaussian_model.py:197-200,
 *      246-249: initial Gaussian scales from a point cloud) --------------------------------------------------
 * out[i] = mean of the three smallest squared distances from points[i] to OTHER points (duplicates count, at 0;
 * fewer than four points leave FLT_MAX terms in the sum, as in the reference).  points: [n,3] fp32 device, out: [n]
 * fp32 device; scratch: caller-owned, egr_knn3_scratch_bytes(n) bytes, 256-byte aligned.  Exact search over the
 * tracer's Morton order + LBVH; d2 = fma(dz,dz,fma(dx,dx,dy*dy)), result = ((b0 + b1) + b2) / 3. */
size_t egr_knn3_scratch_bytes(uint32_t n);
int egr_knn3_mean_dist2(const float *points, uint32_t n, float *out, void *scratch, size_t scratch_bytes, void *stream);

/* ---- next row f4: pruning as one stream-ordered pipeline (train.py:238-245 builds the mask
 *      `total_weight / pruning_interval < min_weight`; scene/gaussian_model.py:479-532 then compacts the 8 parameter
 *      tensors, their two Adam moments and six side arrays with ~30 separate `tensor[mask]` calls, each of which
 *      synchronises the host to learn the new size) -----------------------------------------------------------
 * egr_prune_keep_mask: keep[i] = !(total_weight[i] / interval < min_weight), the complement of the reference's
 *   prune mask (NaN weights are kept, as `<` is false for them).
 * egr_compact_rows: stable compaction (== `src[keep]` for every array) of `num_arrays` row-major arrays that share
 *   the row count n; array a has rows of row_bytes[a] bytes (a multiple of 4, <= 64; base pointers 4-byte aligned).
 *   dst[a] must not alias src[a] and needs room for n rows; *count_out (device, uint32) receives the number of
 *   kept rows; the host reads it once to shrink its views.  src/dst/row_bytes are HOST arrays of num_arrays
 *   (<= EGR_COMPACT_MAX_ARRAYS) entries, read during the call.  scratch: egr_compact_scratch_bytes(n), 256-aligned. */
/* egr_prune_near_cameras: keep[i] = 0 for every point closer to ANY camera centre than that camera's znear
 *   (Scene.select_points_to_prune_near_cameras, scene/scene.py:88-105: one `(points - T).norm(dim=1) < znear`
 *   pass per training camera, OR-ed); other entries of keep are left as they are.  centers [num_cameras,3] and
 *   znear [num_cameras] are device fp32 arrays. */
int egr_prune_near_cameras(const float *mean, uint32_t n, const float *centers, const float *znear,
                           uint32_t num_cameras, uint8_t *keep, void *stream);
#define EGR_COMPACT_MAX_ARRAYS 40
int egr_prune_keep_mask(const float *total_weight, uint32_t n, float interval, float min_weight, uint8_t *keep,
                        void *stream);
size_t egr_compact_scratch_bytes(uint32_t n);
int egr_compact_rows(const uint8_t *keep, uint32_t n, uint32_t num_arrays, const void *const *src, void *const *dst,
                     const uint32_t *row_bytes, void *scratch, size_t scratch_bytes, uint32_t *count_out, void *stream);

/* ---- next row f2: the edit path on the device (scene/editable_gaussian_model.py:30-50 selection
 *      masks, :201-279 translate / scale / rotate of the selected Gaussians, :324-328 delete; driven by
 *      gaussian_viewer.py:283-344).  The reference runs N-sized torch / kornia op chains with boolean-mask
 *      indexing (a host sync each); here one launch per operation, in place, no sync. ---------------------------
 * egr_select_box:      mask[i] = all(mean[i] - min >= 0) && all(mean[i] - max <= 0)              (:45-48)
 * egr_select_cylinder: elliptic cylinder inscribed in the box: ((x-cx)/hx)^2 + ((y-cy)/hy)^2 <= 1 and
 *                      zmin <= z <= zmax, with c = 0.5*(min+max), h = 0.5*(max-min) in fp32       (:30-44)
 *   `min`/`max` are HOST float[3], read during the call.  mask: device uint8[n] (1 = selected).
 * egr_edit_apply: for every selected Gaussian, in place (each array may be NULL to skip it):
 *   mean     <- R*(((mean + t) - oc)*s + oc - oc) + oc with oc = center + t   (:208-236, same op order; the 3x3
 *               product is an explicit fma chain where the reference calls a GEMM)
 *   scale    <- log(exp(scale) * s)                                            (:240-254, raw log-scales)
 *   rotation <- unit(q_edit (x) rotation/|rotation|)  (:256-279 composes rotation MATRICES through kornia and
 *               converts back: the same rotation, as a unit quaternion that may differ in sign)
 *   opacity  <- opacity*0 - 1e8 when `kill` != 0                              (:324-328)
 *   `edit` is a HOST struct read during the call. */
typedef struct EgrEdit {
    float translate[3];
    float center[3];      /* centre of the object's bounding box, 0.5*(min+max) */
    float scale;          /* uniform */
    float rotation[9];    /* row-major matrix of the edit rotation (identity for none) */
    float quaternion[4];  /* the same rotation as a unit quaternion (r,x,y,z) */
    int kill;
} EgrEdit;
int egr_select_box(const float *mean, uint32_t n, const float *min3, const float *max3, uint8_t *mask, void *stream);
int egr_select_cylinder(const float *mean, uint32_t n, const float *min3, const float *max3, uint8_t *mask, void *stream);
int egr_edit_apply(const uint8_t *mask, uint32_t n, float *mean, float *scale, float *rotation, float *opacity,
                   const EgrEdit *edit, void *stream);
/* egr_duplicate_rows: EditableGaussianModel.duplicate_object (scene/editable_gaussian_model.py:284-322: nine
 *   `tensor[selection].clone()` + nine `torch.cat`, i.e. ~27 launches and a host sync per indexing) as one scan + ONE
 *   launch: for every array a, dst[a] = cat(src[a], src[a][mask]); the appended copy of array `add_array` (the means;
 *   -1 for none) additionally gets the HOST row `add_row` added (offset + edit translation, :290).  Same argument
 *   conventions as egr_compact_rows; dst[a] must have room for n + (number of selected rows) rows -- the caller
 *   learns that number from the mask (one host read, as the reference's indexing needs) or from an upper bound;
 *   *count_out (device) receives it as well.  scratch: egr_compact_scratch_bytes(n), 256-aligned. */
int egr_duplicate_rows(const uint8_t *mask, uint32_t n, uint32_t num_arrays, const void *const *src, void *const *dst,
                       const uint32_t *row_bytes, int add_array, const float *add_row, void *scratch,
                       size_t scratch_bytes, uint32_t *count_out, void *stream);

/* ---- next row f1: the host step around the tracer as ONE launch.  Reference, per iteration:
 *      _import_param_gradients (renderer/gaussian_raytracer.py:53-62, 8 add_), the scale decay
 *      `_scaling = log(exp(_scaling) * scale_decay)` (train.py:224-226), torch.optim.Adam.step over 8 parameter
 *      groups with eps 1e-15 (scene/gaussian_model.py:333-338; train.py:247), zero_grad of the model and of the
 *      tracer (train.py:248-249) and three clamps (train.py:251-254): ~30 launches, ~1 GB of elementwise traffic.
 * egr_adam_step reads the gradient straight from the flat record (96 B), updates the 8 raw parameter arrays and
 * their two moments in place (Adam without weight decay / amsgrad, bias-corrected as torch does:
 *   m = m + (1-b1)(g-m); v = b2 v + (1-b2) g g; p -= lr/(1-b1^t) * m / (sqrt(v)/sqrt(1-b2^t) + eps)),
 * applies the scale decay to `scale` BEFORE its update when scale_decay < 1, clamps rgb >= 0, roughness and f0 to
 * [0,1] AFTER the update when `clamp` != 0, and zeroes the record (total_weight column included only when
 * zero_total_weight != 0; the reference zeroes it every pruning interval, train.py:243).  Arrays of both structs
 * follow EgrGradTargets' layout; `opt` is a HOST struct read during the call; `step` is the 1-based step count. */
typedef struct EgrAdam {   /* hyper-parameters are doubles, as the Python floats torch derives its fp32 constants from */
    double lr[8];         /* order: mean, opacity, scale, roughness, rotation, rgb, normal, f0 */
    double beta1, beta2, eps;
    double scale_decay;   /* 1 = none */
    uint32_t step;
    int clamp;
    int zero_grads;
    int zero_total_weight;
} EgrAdam;
int egr_adam_step(float *grad_records, uint32_t count, const EgrGradTargets *params, const EgrGradTargets *exp_avg,
                  const EgrGradTargets *exp_avg_sq, const EgrAdam *opt, void *stream);

/* Debug/test access: copies of internal arrays (device to device, on `stream`). */
int egr_bvh_export(const void *storage, size_t storage_bytes, uint32_t num_gaussians, uint32_t *sorted_ids,
                   float *records, float *nodes, uint32_t *parents, void *stream);

/* The 4-wide tree the bounce steps walk (every second level of the binary tree collapsed; built by egr_bvh_build,
 * boxes refreshed by every egr_bvh_refit): wide [cap,32] floats = 4 x (child bmin.xyz, link bits) (child bmax.xyz, -),
 * wide_slot [2n-1] = for every binary node the slot (4 * wide node + child) that mirrors its box (0xFFFFFFFF: none),
 * *num_wide = nodes in use; cap = egr_bvh_wide_capacity(num_gaussians).  A link is a wide-node index, or
 * 0x80000000 | (count-1) << 27 | first leaf for a subtree of <= 16 leaves, or 0xFFFFFFFF (empty slot). */
uint32_t egr_bvh_wide_capacity(uint32_t num_gaussians);
int egr_bvh_export_wide(const void *storage, size_t storage_bytes, uint32_t num_gaussians, float *wide,
                        uint32_t *wide_slot, uint32_t *num_wide, void *stream);

/* Decode the hit log of the last grads-enabled egr_raytrace: per pixel, the ORIGINAL Gaussian ids (and distances)
 * of the hits composited at `step`, in composite order; ids/ts are [H*W, cap], counts [H*W] (not capped).
 * Plays the role of walking the reference's backward per-pixel linked list (per_pixel_linked_list.h:19-75). */
int egr_hitlog_export(const EgrParams *p, int step, int cap, int32_t *ids, float *ts, int32_t *counts, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* EGR_B200_H */
