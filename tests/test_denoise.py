"""denoise(): the headless stand-in for the reference's OptiX AI denoiser (csrc/optix/denoiser_wrapper.h:76-104).
The OptiX network itself is unavailable on this box (proprietary, inside the driver), so there is no reference output
to compare with; the CUDA filter is pinned to its own definition (oracle/denoise_oracle.py) and to the properties a
guided denoiser must have: constants stay constant, guide edges stay sharp, Monte-Carlo noise goes down."""
import numpy as np
import pytest

from egr_b200 import scenes
from oracle import denoise_oracle


def test_oracle_keeps_constants_and_guide_edges():
    H, W = 24, 40
    g = np.random.default_rng(0)
    nrm = np.zeros((H, W, 3), np.float32)
    nrm[..., 2] = 1.0
    depth = np.full((H, W), 2.0, np.float32)
    const = np.full((H, W, 3), 0.37, np.float32)
    assert np.abs(denoise_oracle.denoise(const, nrm, depth) - const).max() < 1e-6
    # two planes meeting at x = W/2: different colour, normal and depth; noise inside each plane
    img = np.zeros((H, W, 3), np.float32)
    img[:, : W // 2] = (0.8, 0.2, 0.1)
    img[:, W // 2:] = (0.1, 0.3, 0.9)
    nrm[:, W // 2:] = (1.0, 0.0, 0.0)
    depth[:, W // 2:] = 3.0
    noisy = img + g.normal(0, 0.05, img.shape).astype(np.float32)
    out = denoise_oracle.denoise(noisy, nrm, depth)
    assert np.abs(out - img).mean() < 0.4 * np.abs(noisy - img).mean()      # noise goes down
    assert np.abs(out[:, W // 2 - 1] - img[:, W // 2 - 1]).max() < 0.1      # ... and nothing bleeds over the edge
    assert np.abs(out[:, W // 2] - img[:, W // 2]).max() < 0.1


@pytest.mark.gpu
def test_cuda_denoise_matches_oracle_and_reduces_noise():
    import torch

    import helpers as H

    W, Hh = 96, 64
    sc = scenes.random_scene(4000, seed=6)
    sc["roughness"][:] = 0.35  # wide GGX lobes: visibly noisy reflections at one sample per pixel
    cam = scenes.random_camera(W, Hh)
    rt = H.make_gpu_raytracer(sc, cam, dict(num_bounces=2, jitter_primary_rays=True))
    fb = rt.get_framebuffer()
    with torch.no_grad():
        rt.raytrace()
        rt.denoise()
        one = fb.output_final[0].cpu().numpy().copy()
        den = fb.output_denoised[0].cpu().numpy().copy()
        ref = denoise_oracle.denoise(one, fb.output_normal[0].cpu().numpy(), fb.output_depth[0].cpu().numpy())
        assert np.abs(den - ref).max() <= 1e-5 * max(1.0, float(np.abs(ref).max())), float(np.abs(den - ref).max())
        rt.denoise(passthrough=True)
        assert torch.equal(fb.output_denoised, fb.output_final)
        # converged image: 128 accumulated samples (render.py's evaluation protocol)
        rt.get_config().accumulate_samples.fill_(True)
        rt.reset_accumulators()
        for _ in range(128):
            rt.raytrace()
        conv = fb.output_final[0].cpu().numpy().copy()
    e_noisy, e_den = float(np.abs(one - conv).mean()), float(np.abs(den - conv).mean())
    assert e_noisy > 0 and e_den < 0.9 * e_noisy, (e_noisy, e_den)
    # the next raytrace() is unaffected by the scratch planes the filter used
    rt.get_config().accumulate_samples.fill_(False)
    rt.get_metadata().total_num_calls.zero_()
    with torch.no_grad():
        rt.raytrace()
    assert np.array_equal(fb.output_final[0].cpu().numpy(), one)
