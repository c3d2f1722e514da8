"""Shared test plumbing: drive the CUDA tracer through its Python facade (which calls the C ABI) and the CPU
oracle on identical inputs."""
from __future__ import annotations

import numpy as np

import egr_b200
from egr_b200 import scenes
from oracle import oracle as orc

OUT_KEYS = list(orc.OUT_PLANES) + ["final"]
GRAD_MAP = dict(dL_drgb="rgb", dL_dnormal="normal", dL_df0="f0", dL_droughness="roughness", dL_dopacity="opacity",
                dL_dscale="scale", dL_dmean="mean", dL_drotation="rotation")


def oracle_config(rt_cfg: dict, max_bounces=2):
    return orc.make_config(max_bounces=max_bounces, **rt_cfg)


def make_gpu_raytracer(scene, cam, cfg: dict, max_bounces=2, debug_rays=True, hit_capacity=4_000_000):
    import torch

    n = scene["mean"].shape[0]
    rt = egr_b200.make_raytracer(cam["width"], cam["height"], n, 1000, hit_capacity, max_bounces=max_bounces,
                                 debug_rays=debug_rays)
    g = rt.get_gaussians()
    for k in scenes.PARAM_KEYS:
        getattr(g, k).copy_(torch.from_numpy(np.ascontiguousarray(scene[k])))
    c = rt.get_config()
    for k, v in cfg.items():
        getattr(c, k).fill_(v)
    camera = rt.get_camera()
    camera.znear.fill_(cam.get("znear", 0.01))
    camera.zfar.fill_(cam.get("zfar", 999.9))
    camera.vertical_fov_radians.fill_(cam["fovy"])
    camera.set_pose(torch.from_numpy(np.asarray(cam["origin"], np.float32)), torch.from_numpy(cam["c2w"]))
    rt.rebuild_bvh()
    return rt


def set_targets(rt, targets):
    import torch

    fb = rt.get_framebuffer()
    for k, v in targets.items():
        getattr(fb, "target_" + k).copy_(torch.from_numpy(np.ascontiguousarray(v)))


def gpu_outputs(rt):
    fb = rt.get_framebuffer()
    res = {k: getattr(fb, "output_" + k).cpu().numpy() for k in OUT_KEYS}
    st = rt.get_stats()
    for k in ("num_accumulated_per_pixel", "num_hits_per_pixel", "num_composited_per_pixel", "num_steps_per_pixel",
              "num_traversed_per_pixel", "hitlog_status"):
        res[k] = getattr(st, k).cpu().numpy()
    res["random_seeds"] = rt.get_metadata().random_seeds.cpu().numpy().astype(np.uint32)
    if fb.debug_ray_origin is not None:
        res["used_ray_origin"] = fb.debug_ray_origin.cpu().numpy()
        res["used_ray_direction"] = fb.debug_ray_direction.cpu().numpy()
        res["step_executed"] = fb.debug_step_executed.cpu().numpy()
    return res


def gpu_grads(rt):
    g = rt.get_gaussians()
    res = {k: getattr(g, k).cpu().numpy().copy() for k in GRAD_MAP}
    res["total_weight"] = g.total_weight.cpu().numpy().copy()
    return res


def oracle_camera(cam):
    return orc.make_camera(cam["origin"], cam["c2w"], cam["fovy"], cam.get("znear", 0.01), cam.get("zfar", 999.9))


def rel_err(a, b):
    """max |a-b| relative to max |b| (gradients are compared per tensor against the oracle's scale)."""
    return float(np.abs(a - b).max() / max(float(np.abs(b).max()), 1e-30))
