"""Renderer facade: the caller side of the boundary, kept call-for-call identical to the reference's
``editable_gauss_refl/renderer/gaussian_raytracer.py:10-151`` and ``gaussian_renderer.py:21-92`` so that
``train.py`` / ``render.py`` / ``measure_fps.py`` / the viewer's edit->re-render sequence run unchanged against
this binding.  ``pc`` is duck-typed (the attributes of ``GaussianModel`` the reference touches:
``cfg, get_scaling, _get_scaling, _get_rotation, get_xyz, _opacity, get_diffuse, get_normal, get_roughness,
get_f0`` and the raw ``_xyz ... _f0`` parameters with ``.grad``); ``camera`` needs ``R, FoVy, camera_center`` and
the ``*_image`` targets.
"""
from __future__ import annotations

import os
from types import SimpleNamespace

import numpy as np
import torch


class GaussianRaytracer:
    def __init__(self, pc, image_width: int, image_height: int, **raytracer_kwargs):
        from . import make_raytracer

        self.image_width = image_width
        self.image_height = image_height
        self.cuda_module = make_raytracer(image_width, image_height, pc.get_scaling.shape[0], **raytracer_kwargs)
        config = self.cuda_module.get_config()
        config.loss_weight_diffuse.fill_(pc.cfg.loss_weight_diffuse)
        config.loss_weight_specular.fill_(pc.cfg.loss_weight_specular)
        config.loss_weight_normal.fill_(pc.cfg.loss_weight_normal)
        config.loss_weight_depth.fill_(pc.cfg.loss_weight_depth)
        config.loss_weight_f0.fill_(pc.cfg.loss_weight_f0)
        config.loss_weight_roughness.fill_(pc.cfg.loss_weight_roughness)
        config.transmittance_threshold.fill_(pc.cfg.transmittance_threshold)
        config.alpha_threshold.fill_(pc.cfg.alpha_threshold)
        config.exp_power.fill_(pc.cfg.exp_power)
        self.pc = pc
        self._upload_stream = None
        # False when the optimiser reads the tracer's gradient record directly (optim.FusedAdam / trainer.Trainer):
        # the eight `.grad.add_` of _import_param_gradients are then skipped
        self.import_gradients = True
        self._export_param_values()
        self.cuda_module.rebuild_bvh()

    @torch.no_grad()
    def rebuild_bvh(self):
        # train.py calls this every pruning interval, right after the prune's own host reads: the cheapest place to
        # notice that a backward since the last rebuild ran out of hit-log space (its gradients were dropped)
        if self.cuda_module.hitlog_overflowed(clear=True):
            import warnings

            warnings.warn("egr_b200: the backward hit log overflowed since the last rebuild_bvh(); the gradients of the "
                          "affected pixel tiles were dropped.  Construct the tracer with a larger ppll_backward_size.",
                          RuntimeWarning, stacklevel=2)
        self.cuda_module.resize(self.pc._xyz.shape[0])
        self._export_param_values()
        self.cuda_module.rebuild_bvh()

    @torch.no_grad()
    def _export_param_values(self):
        g = self.cuda_module.get_gaussians()
        pc = self.pc
        srcs = dict(scale=pc._get_scaling, rotation=pc._get_rotation, mean=pc.get_xyz, opacity=pc._opacity,
                    rgb=pc.get_diffuse, normal=pc.get_normal, roughness=pc.get_roughness, f0=pc.get_f0)
        if g.mean.shape[0] == srcs["mean"].shape[0] and self.cuda_module.export_parameters_from(srcs):
            return  # the eight copies below as one launch
        g.scale.copy_(self.pc._get_scaling)
        g.rotation.copy_(self.pc._get_rotation)
        g.mean.copy_(self.pc.get_xyz)
        g.opacity.copy_(self.pc._opacity)
        g.rgb.copy_(self.pc.get_diffuse)
        g.normal.copy_(self.pc.get_normal)
        g.roughness.copy_(self.pc.get_roughness)
        g.f0.copy_(self.pc.get_f0)

    @torch.no_grad()
    def _import_param_gradients(self):
        g = self.cuda_module.get_gaussians()
        pairs = dict(mean=self.pc._xyz, opacity=self.pc._opacity, scale=self.pc._scaling, rotation=self.pc._rotation,
                     rgb=self.pc._diffuse, normal=self.pc._normal, roughness=self.pc._roughness, f0=self.pc._f0)
        n = g.grad_records.shape[0]
        fused = all(p.grad is not None and p.grad.is_cuda and p.grad.is_contiguous() and p.grad.dtype == torch.float32
                    and p.grad.numel() == n * getattr(g, k).shape[1] for k, p in pairs.items())
        if fused:  # the eight add_ of the reference (gaussian_raytracer.py:53-62) as one launch over the flat record
            self.cuda_module.accumulate_gradients_into({k: p.grad for k, p in pairs.items()})
            return
        self.pc._xyz.grad.add_(g.mean.grad)
        self.pc._opacity.grad.add_(g.opacity.grad)
        self.pc._scaling.grad.add_(g.scale.grad)
        self.pc._rotation.grad.add_(g.rotation.grad)
        self.pc._diffuse.grad.add_(g.rgb.grad)
        self.pc._normal.grad.add_(g.normal.grad)
        self.pc._roughness.grad.add_(g.roughness.grad)
        self.pc._f0.grad.add_(g.f0.grad)

    def zero_grad(self):
        g = self.cuda_module.get_gaussians()
        for name in ("rgb", "opacity", "scale", "rotation", "mean", "normal", "roughness", "f0"):
            getattr(g, name).grad.zero_()

    def __call__(self, viewpoint_camera, target=None, target_diffuse=None, target_specular=None, target_depth=None,
                 target_normal=None, target_roughness=None, target_f0=None, force_update_bvh=False, denoise=False,
                 znear=0.01, zfar=999.9):
        with torch.no_grad():
            R = viewpoint_camera.R
            # a numpy pose stays on the host (the reference uploads it with a blocking `.cuda()`); set_pose passes
            # host parts by value
            R = torch.from_numpy(np.ascontiguousarray(R, dtype=np.float32)) if isinstance(R, np.ndarray) else R.float()
            R_c2w_blender = -R
            R_c2w_blender[:, 0] = -R_c2w_blender[:, 0]
            camera = self.cuda_module.get_camera()
            camera.znear.fill_(float(os.getenv("ZNEAR", znear)))
            camera.zfar.fill_(float(os.getenv("ZFAR", zfar)))
            camera.vertical_fov_radians.fill_(float(viewpoint_camera.FoVy))
            camera.set_pose(viewpoint_camera.camera_center.contiguous(), R_c2w_blender.contiguous())
            self._export_param_values()
            fb = self.cuda_module.get_framebuffer()
            tgts = (("target_diffuse", target_diffuse), ("target_specular", target_specular),
                    ("target_depth", target_depth), ("target_normal", target_normal),
                    ("target_roughness", target_roughness), ("target_f0", target_f0))
            # Only the backward reads the targets (fused loss, shaders.cu:150-166), so host-resident targets are
            # uploaded on a second stream underneath the refit and the forward launches of this call.
            overlap = any(t is not None and t.device.type == "cpu" for _, t in tgts)
            main = torch.cuda.current_stream()
            if overlap:
                if self._upload_stream is None:
                    self._upload_stream = torch.cuda.Stream()
                self._upload_stream.wait_stream(main)  # the previous call's backward is done with the buffers
            with torch.cuda.stream(self._upload_stream if overlap else main):
                for name, tgt in tgts:
                    buf = getattr(fb, name)
                    if tgt is not None:
                        if tgt.device.type == "cpu":
                            # DMA the contiguous CHW image, permute to HWC on the device (a strided host-side
                            # copy would permute 36 MB per view on one CPU core first)
                            tgt = tgt.to(buf.device, non_blocking=True)
                        buf.copy_(tgt.moveaxis(0, -1), non_blocking=True)
                    else:
                        buf.zero_()
        if torch.is_grad_enabled() or force_update_bvh:
            self.cuda_module.update_bvh()
        if overlap:
            from . import _capi

            # primary step first; the rest of the call (bounce steps + backward, which the library overlaps among
            # themselves) needs the targets
            S = self.cuda_module._S
            rest = _capi.EGR_PHASE_FINISH
            for s in range(1, S):
                rest |= _capi.EGR_PHASE_STEP(s)
            self.cuda_module.raytrace_phases(_capi.EGR_PHASE_BEGIN | _capi.EGR_PHASE_STEP(0))
            main.wait_stream(self._upload_stream)
            self.cuda_module.raytrace_phases(rest)
        else:
            self.cuda_module.raytrace()
        if denoise:
            self.cuda_module.denoise()
        if torch.is_grad_enabled() and self.import_gradients:
            self._import_param_gradients()
        return {"render": fb.output_rgb.clone()}


def render(camera, raytracer: GaussianRaytracer, targets_available=True, force_update_bvh=False, denoise=False,
           znear=0.01, zfar=999.9):
    do_backprop = torch.is_grad_enabled()
    names = ("original_image", "diffuse_image", "specular_image", "normal_image", "f0_image", "roughness_image",
             "depth_image")
    if targets_available:
        tgt, t_diff, t_spec, t_n, t_f0, t_r, t_d = (getattr(camera, n, None) for n in names)
    else:
        tgt = t_diff = t_spec = t_n = t_f0 = t_r = t_d = None
    with torch.set_grad_enabled(do_backprop):
        raytracer(camera, target=tgt, target_diffuse=t_diff, target_specular=t_spec, target_depth=t_d,
                  target_normal=t_n, target_roughness=t_r, target_f0=t_f0, force_update_bvh=force_update_bvh,
                  denoise=denoise, znear=znear, zfar=zfar)
    fb = raytracer.cuda_module.get_framebuffer()

    def chw(t):
        return t.clone().detach().moveaxis(-1, 1)

    rgb = chw(fb.output_rgb)
    return SimpleNamespace(
        rgb=rgb, final=chw(fb.output_denoised) if denoise else chw(fb.output_final), depth=chw(fb.output_depth),
        normal=chw(fb.output_normal), roughness=chw(fb.output_roughness), f0=chw(fb.output_f0), target=tgt,
        target_diffuse=t_diff, target_specular=t_spec, target_depth=t_d, target_normal=t_n, target_roughness=t_r,
        target_f0=t_f0)
