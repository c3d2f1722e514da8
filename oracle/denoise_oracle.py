"""ORACLE -- TEST INFRASTRUCTURE ONLY.  CPU restatement (float32 numpy) of ``egr_denoise``
(editable-gaussian-reflections_b200/csrc/egr_denoise.cu): the headless stand-in for the reference's
``DenoiserWrapper::run`` (csrc/optix/denoiser_wrapper.h:76-104).  Parity unpinned against the reference by
construction: the reference calls the proprietary OptiX network, which is unavailable on this box; this oracle pins the
CUDA filter to its own published definition (Dammertz et al. 2010, edge-avoiding a-trous wavelets)."""
import numpy as np

PASSES, SIGMA_COLOR, SIGMA_NORMAL, SIGMA_DEPTH, DEPTH_EPS = 3, 0.6, 0.3, 0.05, 1.0e-3  # include/egr_b200.h EGR_DENOISE_*
KERN = np.array([1 / 16, 1 / 4, 3 / 8, 1 / 4, 1 / 16], np.float32)


def denoise(final, normal0, depth0):
    """final [H,W,3], normal0 [H,W,3], depth0 [H,W] (or [H,W,1]) -> denoised [H,W,3], all float32."""
    f32 = np.float32
    c = np.asarray(final, f32).copy()
    n = np.asarray(normal0, f32)
    d = np.asarray(depth0, f32).reshape(c.shape[:2])
    H, W = d.shape
    sigma_c = f32(SIGMA_COLOR)
    inv_sn2 = f32(1.0) / (f32(SIGMA_NORMAL) * f32(SIGMA_NORMAL))
    inv_sd2 = f32(1.0) / (f32(SIGMA_DEPTH) * f32(SIGMA_DEPTH))
    for p in range(PASSES):
        hole = 1 << p
        inv_sc2 = f32(1.0) / (sigma_c * sigma_c)
        acc = np.zeros_like(c)
        wsum = np.zeros((H, W), f32)
        for j in range(-2, 3):
            for i in range(-2, 3):
                dy, dx = j * hole, i * hole
                y0, y1 = max(0, -dy), min(H, H - dy)
                x0, x1 = max(0, -dx), min(W, W - dx)
                if y0 >= y1 or x0 >= x1:
                    continue
                P = (slice(y0, y1), slice(x0, x1))              # centre pixels whose tap (dy, dx) is inside
                Q = (slice(y0 + dy, y1 + dy), slice(x0 + dx, x1 + dx))
                dc = c[Q] - c[P]
                dn = n[Q] - n[P]
                dd = (d[Q] - d[P]) / (np.abs(d[P]) + f32(DEPTH_EPS))
                e = ((dc * dc).sum(-1, dtype=f32) * inv_sc2 + (dn * dn).sum(-1, dtype=f32) * inv_sn2 + dd * dd * inv_sd2)
                w = (KERN[i + 2] * KERN[j + 2]) * np.exp(-e, dtype=f32)
                acc[P] += c[Q] * w[..., None]
                wsum[P] += w
        c = acc * (f32(1.0) / wsum)[..., None]
        sigma_c = sigma_c * f32(0.5)
    return c
