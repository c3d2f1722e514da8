// Backward pass: the in-kernel L1 loss of the reference and its per-Gaussian gradients.  One warp per 8x4 pixel
// tile (same mapping as the forward) re-walks each step's hit log farthest -> nearest, 512 contiguous bytes per
// slot, recomputes the local hit from the packed geometry record, and flushes the 22 per-Gaussian gradient floats
// into the flat [N,24] record with 128-bit vector reductions (REDG.F32x4) after a warp-level segmented sum over
// the lanes that carry the same Gaussian.
//
// Reference: editable_gauss_refl/cuda/csrc/backward_pass.cu:3-222, shaders.cu:155-162,
// core/framebuffer.h:145-152, utils/activations.cu, utils/helpers.cu:36-58, utils/misc.cu:5-24,56-63.
#include "egr_trace.cuh"

namespace egr {

__device__ __forceinline__ void red_add_v4(float *addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
}

// Warp-aggregated flush.  `peers` = lanes holding the same Gaussian (from __match_any_sync).  Tree reduction over
// the rank inside each group: in every round the even-ranked survivors absorb their next (odd-ranked) group
// member, which then drops out; the group's first lane ends with the sum and issues the vector reductions.
template <int NV>
__device__ __forceinline__ void flush_grads(float *grad_records, uint32_t id, bool has, float (&g)[24]) {
    const uint32_t lane = threadIdx.x & 31;
#ifdef EGR_NO_AGGREGATION
    if (has) {
        float *dst = grad_records + (size_t)id * EGR_GRAD_STRIDE;
#pragma unroll
        for (int k = 0; k < NV; k += 4) red_add_v4(dst + k, g[k], g[k + 1], g[k + 2], g[k + 3]);
    }
    return;
#endif
    unsigned peers = __match_any_sync(0xffffffffu, has ? id : (0x80000000u | lane));
    // The tree below costs ~3 shuffle rounds over all 24 values; it only pays when many lanes of the warp hit the
    // same Gaussian (large foreground Gaussians, where same-address reductions would serialise in L2).  With
    // mostly distinct Gaussians the 128-bit reductions go out directly.
    if (!__any_sync(0xffffffffu, __popc(peers) >= 8)) {
        if (has) {
            float *dst = grad_records + (size_t)id * EGR_GRAD_STRIDE;
#pragma unroll
            for (int k = 0; k < NV; k += 4) red_add_v4(dst + k, g[k], g[k + 1], g[k + 2], g[k + 3]);
        }
        return;
    }
    const unsigned lt = (1u << lane) - 1u;
    const unsigned gt = ~lt & ~(1u << lane);
    const bool leader = has && (peers & lt) == 0;
    bool dropped = false;
    while (true) {
        const unsigned after = peers & gt;
        const bool odd = !dropped && (__popc(peers & lt) & 1);
        const int nxt = after ? __ffs(after) - 1 : -1;
        const bool take = !dropped && !odd && nxt >= 0;
        dropped |= odd;
        if (!__any_sync(0xffffffffu, take)) break;
        const int src = nxt >= 0 ? nxt : (int)lane;
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            const float o = __shfl_sync(0xffffffffu, g[k], src);
            g[k] += take ? o : 0.0f;
        }
        peers &= ~__ballot_sync(0xffffffffu, odd);
    }
    if (leader) {
        float *dst = grad_records + (size_t)id * EGR_GRAD_STRIDE;
#pragma unroll
        for (int k = 0; k < NV; k += 4) red_add_v4(dst + k, g[k], g[k + 1], g[k + 2], g[k + 3]);
    }
}

// STEP0: the primary step's loss terms (diffuse, depth, normal, f0, roughness); bounce steps only carry the specular
// radiance term, so their variant keeps far fewer accumulators live.
template <bool STEP0>
__global__ void __launch_bounds__(128, 4) k_backward(const TraceCtx c, const int step_arg) {
    const int step = STEP0 ? 0 : step_arg;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31;
    if (warp >= c.num_warps) return;
    const EgrConfig &cfg = c.p.config;
    const int B = min(*cfg.num_bounces, (int)c.S - 1);
    if (step > B) return;
    if (step > 0 && warp * 32 >= c.ray_count[step]) return;
    uint32_t pix = 0;
    const bool valid = lane_pixel(c, step, warp, lane, pix);  // same lane <-> ray mapping as the forward of `step`
    // What a launch forked under a later forward may read: the backward of step 0 runs on the side stream while the
    // forward of steps >= 1 writes steps_executed[] and the planes of those steps (egr_trace.cu, egr_raytrace_phases),
    // so at step 0 it reads neither: every valid pixel executed step 0, and the primary loss needs no bounce radiance.
    const int nexec = STEP0 ? (valid ? 1 : 0) : (valid ? (int)(c.steps_executed[pix] & 0xFFu) : 0);  // total_effective_steps
    const float exp_power = *cfg.exp_power;
    const float eps_scale_grad = *cfg.eps_scale_grad;
    const float gsf = *cfg.global_scale_factor;
    const EgrFramebuffer &fb = c.p.framebuffer;
    const EgrGaussians &gs = c.p.gaussians;

    // fetch_targets (framebuffer.h:145-152)
    f3 t_diff = mk3(0, 0, 0), t_spec = mk3(0, 0, 0), t_n = mk3(0, 0, 0), t_f0 = mk3(0, 0, 0);
    float t_depth = 0.0f, t_rough = 0.0f;
    f3 spec = mk3(0, 0, 0);
    if (valid) {  // each step reads only the targets its loss uses (backward_pass.cu:81-108)
        if (STEP0) {
            t_diff = ld3(fb.target_diffuse, pix);
            t_n = ld3(fb.target_normal, pix), t_f0 = ld3(fb.target_f0, pix);
            t_depth = fb.target_depth[pix], t_rough = fb.target_roughness[pix];
        } else {
            t_spec = ld3(fb.target_specular, pix);
        }
        if (!STEP0)
            for (int j = 1; j < B + 1; ++j)  // backward_pass.cu:101-104 (non-executed steps hold zeros)
                if (j < nexec) spec = spec + plane3(c, SP_RGB, j, pix);
    }

    {  // one launch per step, last step first (shaders.cu:157)
        const uint32_t *wt_ = c.warp_tail + ((size_t)warp * c.S + step) * 2;
        uint32_t chunk = wt_[0];
        const int nb = (int)wt_[1];
        if (nb == 0 || chunk == 0xFFFFFFFFu) return;
        const int nh = (valid && step < nexec) ? __float_as_int(plane(c, SP_NHITS, step, pix)) : 0;
        const bool on = nh > 0;
        // per-pixel upstream gradients (backward_pass.cu:81-108)
        f3 d_rgb = mk3(0, 0, 0), d_n = mk3(0, 0, 0), d_f0 = mk3(0, 0, 0);
        float d_depth = 0.0f, d_rough = 0.0f;
        f3 o = mk3(0, 0, 0), d = mk3(0, 0, 1);
        f3 rem_rgb = mk3(0, 0, 0), rem_n = mk3(0, 0, 0), rem_f0 = mk3(0, 0, 0);
        float rem_depth = 0.0f, rem_rough = 0.0f, T_end = 1.0f, T_full = 1.0f;
        if (on) {
            if (STEP0) {
                d_rgb = 1.0f / 3.0f * sign3(plane3(c, SP_RGB, 0, pix) - t_diff) * *cfg.loss_weight_diffuse;
                d_depth = sign1(plane(c, SP_DEPTH, 0, pix) - t_depth) * *cfg.loss_weight_depth;
                d_n = 1.0f / 3.0f * sign3(plane3(c, SP_NORMAL, 0, pix) - t_n) * *cfg.loss_weight_normal;
                d_f0 = 1.0f / 3.0f * sign3(plane3(c, SP_F0, 0, pix) - t_f0) * *cfg.loss_weight_f0;
                d_rough = sign1(plane(c, SP_ROUGH, 0, pix) - t_rough) * *cfg.loss_weight_roughness;
            } else {
                const float rprev = plane(c, SP_ROUGH, max(step - 1, 0), pix);
                const float rdw = powf(1.0f - rprev, ROUGH_DOWNWEIGHT_POW);  // backward_pass.cu:11-16
                d_rgb = 1.0f / 3.0f * sign3(spec - t_spec) * *cfg.loss_weight_specular * rdw;
                d_rgb = d_rgb * plane3(c, SP_THR, step - 1, pix);
            }
            o = plane3(c, SP_RAY_O, step, pix), d = plane3(c, SP_RAY_D, step, pix);
            rem_rgb = plane3(c, SP_REM_RGB, step, pix);
            if (STEP0) {
                rem_n = plane3(c, SP_REM_NORMAL, step, pix), rem_f0 = plane3(c, SP_REM_F0, step, pix);
                rem_depth = plane(c, SP_REM_DEPTH, step, pix), rem_rough = plane(c, SP_REM_ROUGH, step, pix);
            }
            T_end = plane(c, SP_T, step, pix), T_full = plane(c, SP_TFULL, step, pix);
        }
        const float tmin = step != 0 ? 0.0f : *c.p.camera.znear, tmax = *c.p.camera.zfar;
        f3 p_rgb = mk3(0, 0, 0), w_rgb = mk3(0, 0, 0), p_n = mk3(0, 0, 0), w_n = mk3(0, 0, 0), p_f0 = mk3(0, 0, 0),
           w_f0 = mk3(0, 0, 0);
        float p_r = 0.0f, w_r = 0.0f, p_d = 0.0f, w_d = 0.0f;

        for (int b = nb - 1; b >= 0; --b) {
            EGR_CHECK(chunk < c.hitlog_chunks);
            const char *base = (const char *)c.p.hitlog + (size_t)chunk * HITLOG_CHUNK_BYTES;
            const uint2 *entries = (const uint2 *)(base + 128);
            const uint32_t prev = *(const uint32_t *)base;
            // the log entry of the next slot is fetched one iteration ahead (coalesced, streamed once from HBM)
            uint2 e_next = (b * KBUF + KBUF - 1) < nh ? __ldcs(&entries[(KBUF - 1) * 32 + lane]) : make_uint2(0, 0);
            for (int slot = KBUF - 1; slot >= 0; --slot) {
                const bool has = (b * KBUF + slot) < nh;
                const uint2 e = e_next;
                if (slot > 0 && (b * KBUF + slot - 1) < nh) e_next = __ldcs(&entries[(slot - 1) * 32 + lane]);
                if (!__any_sync(0xffffffffu, has)) continue;
                float g[24];
#pragma unroll
                for (int k = 0; k < 24; ++k) g[k] = 0.0f;
                uint32_t id = 0;
                if (has) {
                    const uint32_t j = e.x;
                    EGR_CHECK(j < c.bvh.n);
                    const float Tr = __uint_as_float(e.y);
                    float4 r0, r1, r2, r3;
                    ldg2(&c.bvh.records[4 * (size_t)j], r0, r1);
                    ldg2(&c.bvh.records[4 * (size_t)j + 2], r2, r3);
                    float4 q4, s4;  // raw quaternion; exp(scale), 1/|q| as of the last refit
                    ldg2(&c.bvh.rawgeo[2 * (size_t)j], q4, s4);
                    HitEval h;
                    eval_hit(r0, r1, r2, r3, o, d, tmin, tmax, exp_power, h);  // same bits as the forward's evaluation
                    const float dist = h.t, alpha = h.alpha;
                    const float sf = r1.w, sigma = r2.w;
                    id = __float_as_uint(r3.w);
                    EGR_CHECK(id < gs.count);
                    const float gv = h.gaussval;
                    const float4 a0 = __ldg(&c.bvh.appearance[3 * (size_t)j + 0]);
                    const f3 c_rgb = mk3(a0.x, a0.y, a0.z);
                    // colour gradients (backward_pass.cu:110-115): the activations' backward is evaluated on the
                    // ACTIVATED value and therefore always passes the gradient (activations.cu:42,54-56)
                    const float weight = (float)((double)Tr / (1.0 - (double)alpha) * (double)alpha);
                    const f3 g_c = d_rgb * weight;
                    w_rgb = w_rgb + (c_rgb - p_rgb) * Tr, p_rgb = c_rgb;  // backward_pass.cu:117-132
                    if (STEP0) {
                        const float4 a1 = __ldg(&c.bvh.appearance[3 * (size_t)j + 1]);
                        const float4 a2 = __ldg(&c.bvh.appearance[3 * (size_t)j + 2]);
                        const f3 c_n = mk3(a1.x, a1.y, a1.z), c_f0 = mk3(a2.x, a2.y, a2.z);
                        const float c_r = a0.w;
                        w_n = w_n + (c_n - p_n) * Tr, p_n = c_n;
                        w_f0 = w_f0 + (c_f0 - p_f0) * Tr, p_f0 = c_f0;
                        w_r += (c_r - p_r) * Tr, p_r = c_r;
                        w_d += (dist - p_d) * Tr, p_d = dist;
                        const f3 g_nn = d_n * weight, g_ff = d_f0 * weight;
                        g[16] = g_nn.x, g[17] = g_nn.y, g[18] = g_nn.z;
                        g[20] = g_ff.x, g[21] = g_ff.y, g[22] = g_ff.z;
                        g[7] = d_rough * weight;
                    }
                    // alpha gradient (backward_pass.cu:134-148)
                    float dL_da = 0.0f;
                    const float tmp1 = 1.0f / (1.0f - alpha);
                    dL_da += dotp(w_rgb * tmp1, d_rgb);
                    if (STEP0) {  // the bounce steps' upstream gradients of these channels are zero
                        dL_da += dotp(w_n * tmp1, d_n);
                        dL_da += dotp(w_f0 * tmp1, d_f0);
                        dL_da += (w_r * tmp1) * d_rough;
                        dL_da += (w_d * tmp1) * d_depth;
                    }
                    const float tmp2 = -((T_end - T_full) / (1.0f - alpha));
                    dL_da += tmp2 * dotp(rem_rgb, d_rgb);
                    if (STEP0) {
                        dL_da += tmp2 * dotp(rem_n, d_n);
                        dL_da += tmp2 * dotp(rem_f0, d_f0);
                        dL_da += tmp2 * (rem_rough * d_rough);
                        dL_da += tmp2 * (rem_depth * d_depth);
                    }
                    // opacity (backward_pass.cu:150-152), transform (154-167)
                    float g_op = A_MAX * dL_da * gv;
                    g_op = g_op * sigma * (1.0f - sigma);
                    const float dL_dg = A_MAX * dL_da * sigma;
                    const float sqn = dotp(h.x, h.x);
                    const float dsq = gv * det_pow(sqn, exp_power - 1.0f);
                    const f3 dxl = -h.x * dsq * dL_dg;
                    const f3 dxw = mk3(dotp(mk3(r1.x, r2.x, r3.x), dxl), dotp(mk3(r1.y, r2.y, r3.y), dxl),
                                       dotp(mk3(r1.z, r2.z, r3.z), dxl)) *
                                   sf;
                    const f3 l0 = -dxw.x * h.x, l1 = -dxw.y * h.x, l2 = -dxw.z * h.x;
                    // scale / rotation (backward_pass.cu:169-207) from the raw parameters
                    const f3 sc_e = mk3(s4.x, s4.y, s4.z);
                    const float iqn = s4.w;
                    const float r = q4.x * iqn, xq = q4.y * iqn, yq = q4.z * iqn, zq = q4.w * iqn;
                    const float R00 = 1.f - 2.f * (yq * yq + zq * zq), R01 = 2.f * (xq * yq - r * zq),
                                R02 = 2.f * (xq * zq + r * yq);
                    const float R10 = 2.f * (xq * yq + r * zq), R11 = 1.f - 2.f * (xq * xq + zq * zq),
                                R12 = 2.f * (yq * zq - r * xq);
                    const float R20 = 2.f * (xq * zq - r * yq), R21 = 2.f * (yq * zq + r * xq),
                                R22 = 1.f - 2.f * (xq * xq + yq * yq);
                    const f3 Sg = sc_e * sf * gsf;
                    const f3 den = sc_e * sf + mk3(eps_scale_grad, eps_scale_grad, eps_scale_grad);
                    const f3 sod = Sg / den;  // one division per axis instead of one per matrix entry
                    const f3 rot0 = mk3(R00, R01, R02) * sod;
                    const f3 rot1 = mk3(R10, R11, R12) * sod;
                    const f3 rot2 = mk3(R20, R21, R22) * sod;
                    const f3 g_scale = (l0 * rot0 + l1 * rot1 + l2 * rot2) * sc_e;
                    const f3 dr0 = l0 * sc_e, dr1 = l1 * sc_e, dr2 = l2 * sc_e;
                    const float dLr =
                        2.f * xq * (dr2.y - dr1.z) + 2.f * yq * (dr0.z - dr2.x) + 2.f * zq * (dr1.x - dr0.y);
                    const float dLx = -4.f * xq * (dr1.y + dr2.z) + 2.f * yq * (dr0.y + dr1.x) +
                                      2.f * zq * (dr0.z + dr2.x) + 2.f * r * (dr2.y - dr1.z);
                    const float dLy = 2.f * xq * (dr0.y + dr1.x) - 4.f * yq * (dr0.x + dr2.z) +
                                      2.f * zq * (dr1.z + dr2.y) + 2.f * r * (dr0.z - dr2.x);
                    const float dLz = 2.f * xq * (dr0.z + dr2.x) + 2.f * yq * (dr1.z + dr2.y) -
                                      4.f * zq * (dr0.x + dr1.y) + 2.f * r * (dr1.x - dr0.y);
                    const float dq = dLr * q4.x + dLx * q4.y + dLy * q4.z + dLz * q4.w;  // activations.cu:71-73
                    const float i1 = iqn, i3 = (iqn * iqn) * iqn;
                    g[0] = -dxw.x, g[1] = -dxw.y, g[2] = -dxw.z;  // mean (backward_pass.cu:175)
                    g[3] = g_op;
                    g[4] = g_scale.x, g[5] = g_scale.y, g[6] = g_scale.z;
                    g[8] = (dq * -q4.x) * i3 + dLr * i1, g[9] = (dq * -q4.y) * i3 + dLx * i1;
                    g[10] = (dq * -q4.z) * i3 + dLy * i1, g[11] = (dq * -q4.w) * i3 + dLz * i1;
                    g[12] = g_c.x, g[13] = g_c.y, g[14] = g_c.z;
                    g[15] = weight;  // total_weight (backward_pass.cu:220)
                }
                if (STEP0) flush_grads<24>(gs.grad_records, id, has, g);
                else flush_grads<16>(gs.grad_records, id, has, g);
            }
            chunk = prev;
        }
    }
}

// Debug/test: decode one step's hit log into per-pixel ordered lists of ORIGINAL Gaussian ids (composite order).
__global__ void __launch_bounds__(128)
k_hitlog_export(const TraceCtx c, const int step, const int cap, int32_t *ids, float *ts, int32_t *counts) {
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31;
    if (warp >= c.num_warps) return;
    if (step > 0 && warp * 32 >= c.ray_count[step]) return;
    uint32_t pix = 0;
    const bool valid = lane_pixel(c, step, warp, lane, pix);
    const int nexec = valid ? (int)(c.steps_executed[pix] & 0xFFu) : 0;
    const uint32_t *wt_ = c.warp_tail + ((size_t)warp * c.S + step) * 2;
    uint32_t chunk = wt_[0];
    const int nb = (int)wt_[1];
    const int nh = (valid && step < nexec && nb > 0) ? __float_as_int(plane(c, SP_NHITS, step, pix)) : 0;
    if (valid) counts[pix] = nh;
    if (nb == 0 || chunk == 0xFFFFFFFFu) return;
    const f3 o = nh > 0 ? plane3(c, SP_RAY_O, step, pix) : mk3(0, 0, 0), d = nh > 0 ? plane3(c, SP_RAY_D, step, pix) : mk3(0, 0, 1);
    const float tmin = step != 0 ? 0.0f : *c.p.camera.znear, tmax = *c.p.camera.zfar, exp_power = *c.p.config.exp_power;
    for (int b = nb - 1; b >= 0; --b) {
        const char *base = (const char *)c.p.hitlog + (size_t)chunk * HITLOG_CHUNK_BYTES;
        const uint2 *entries = (const uint2 *)(base + 128);
        for (int slot = 0; slot < KBUF; ++slot) {
            const int idx = b * KBUF + slot;
            if (idx < nh && idx < cap) {
                const uint32_t j = entries[slot * 32 + lane].x;
                const float4 *rp = &c.bvh.records[4 * (size_t)j];
                ids[(size_t)pix * cap + idx] = (int32_t)__float_as_uint(rp[3].w);
                if (ts) {  // the distance is not logged: recomputed exactly as the forward evaluated it
                    HitEval h;
                    eval_hit(rp[0], rp[1], rp[2], rp[3], o, d, tmin, tmax, exp_power, h);
                    ts[(size_t)pix * cap + idx] = h.t;
                }
            }
        }
        chunk = *(const uint32_t *)base;
    }
}

int launch_hitlog_export(const TraceCtx &c, int step, int cap, int32_t *ids, float *ts, int32_t *counts, cudaStream_t st) {
    const uint32_t blocks = (c.num_warps + 3) / 4;
    k_hitlog_export<<<blocks, 128, 0, st>>>(c, step, cap, ids, ts, counts);
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}

// Backward of the steps in `step_mask` (bit s = step s), last step first (shaders.cu:157).
int launch_backward_impl(const TraceCtx &c, uint32_t step_mask, cudaStream_t st) {
    const uint32_t blocks = (c.num_warps + 3) / 4;
    for (int step = (int)c.S - 1; step >= 0; --step)
        if (step_mask >> step & 1u) {
            if (step == 0) k_backward<true><<<blocks, 128, 0, st>>>(c, 0);
            else k_backward<false><<<blocks, 128, 0, st>>>(c, step);
        }
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}

}  // namespace egr
