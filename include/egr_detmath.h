/*
 * egr_detmath.h — the transcendental functions of the tracer as IEEE-only sequences.
 *
 * exp / log / sin / cos are evaluated with fixed sequences of fmaf, multiplications, additions, round-to-nearest
 * conversions and bit manipulation, so that the CUDA kernels (sm_100a) and a CPU build produce the SAME BITS.
 * That is what lets hit-ID lists, geometry records, rays and framebuffers be compared bit for bit between the
 * device and the CPU oracle instead of "within libm ulps".  Each function is within 2 ulp of the correctly rounded
 * result on the ranges the tracer uses (CUDA's own expf/logf/sinf/cosf are specified to 1-2 ulp as well), so
 * replacing the reference's libm calls (utils/kernel.cu:3-12, utils/activations.cu:5,59, core/camera.h:19,
 * utils/ggx_brdf.h:159) by these stays far inside the 1e-4 output tolerance.
 *
 * Plain C99 / C++ / CUDA; no dependencies.  Compile the CPU side with -ffp-contract=off.
 */
#ifndef EGR_DETMATH_H
#define EGR_DETMATH_H

#include <math.h>
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define EGR_HD __host__ __device__ __forceinline__
#else
#define EGR_HD static inline
#endif

EGR_HD float egr_bits_to_float(uint32_t u) {
    float f;
    memcpy(&f, &u, 4);
    return f;
}
EGR_HD uint32_t egr_float_to_bits(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    return u;
}

/* exp(u) for u in (-87, 88): Cody-Waite reduction by ln2 (hi/lo) and a degree-7 Horner polynomial. */
EGR_HD float egr_det_exp(float u) {
    if (!(u > -87.0f)) return 0.0f;
    if (u > 88.0f) u = 88.0f;
    float n = rintf(u * 1.44269504088896341f);
    float r = fmaf(-n, 0.693145751953125f, u);
    r = fmaf(-n, 1.42860682030941723212e-6f, r);
    float p = 1.0f / 5040.0f;
    p = fmaf(p, r, 1.0f / 720.0f);
    p = fmaf(p, r, 1.0f / 120.0f);
    p = fmaf(p, r, 1.0f / 24.0f);
    p = fmaf(p, r, 1.0f / 6.0f);
    p = fmaf(p, r, 0.5f);
    p = fmaf(p, r, 1.0f);
    p = fmaf(p, r, 1.0f);
    int e = (int)n;
    return p * egr_bits_to_float((uint32_t)(e + 127) << 23);
}

/* log(x) for positive normal x: x = m 2^e with m in [sqrt(1/2), sqrt(2)), log m = 2 atanh((m-1)/(m+1)). */
EGR_HD float egr_det_log(float x) {
    uint32_t b = egr_float_to_bits(x);
    int e = (int)((b >> 23) & 0xffu) - 127;
    float m = egr_bits_to_float((b & 0x007fffffu) | 0x3f800000u);
    if (m > 1.41421356237f) {
        m = m * 0.5f;
        e += 1;
    }
    float s = (m - 1.0f) / (m + 1.0f);
    float z = s * s;
    float p = 2.0f / 11.0f;
    p = fmaf(p, z, 2.0f / 9.0f);
    p = fmaf(p, z, 2.0f / 7.0f);
    p = fmaf(p, z, 2.0f / 5.0f);
    p = fmaf(p, z, 2.0f / 3.0f);
    p = fmaf(p, z, 2.0f);
    float fe = (float)e;
    return fmaf(fe, 0.693145751953125f, fmaf(fe, 1.42860682030941723212e-6f, s * p));
}

/* x^y for x > 0 through exp(y log x); relative error ~ |y log x| * 1e-7. */
EGR_HD float egr_det_pow(float x, float y) { return egr_det_exp(y * egr_det_log(x)); }

/* sin and cos of x, |x| <= 16: reduction by pi/2 in three parts, Cephes single-precision kernels. */
EGR_HD void egr_det_sincos(float x, float *sn, float *cs) {
    float k = rintf(x * 0.636619772367581343f);
    float r = fmaf(-k, 1.5703125f, x);
    r = fmaf(-k, 4.837512969970703125e-4f, r);
    r = fmaf(-k, 7.54978995489188216e-8f, r);
    float z = r * r;
    float ps = -1.9515295891e-4f;
    ps = fmaf(ps, z, 8.3321608736e-3f);
    ps = fmaf(ps, z, -1.6666654611e-1f);
    float s = fmaf(r * z, ps, r);
    float pc = 2.443315711809948e-5f;
    pc = fmaf(pc, z, -1.388731625493765e-3f);
    pc = fmaf(pc, z, 4.166664568298827e-2f);
    float c = fmaf(z * z, pc, fmaf(-0.5f, z, 1.0f));
    int q = ((int)k) & 3;
    float ss = (q & 1) ? c : s;
    float cc = (q & 1) ? s : c;
    if (q == 1 || q == 2) cc = -cc;
    if (q == 2 || q == 3) ss = -ss;
    *sn = ss;
    *cs = cc;
}

#endif /* EGR_DETMATH_H */
