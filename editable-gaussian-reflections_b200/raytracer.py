"""Host-side mirror of the reference's TorchScript class ``torch.classes.raytracer.Raytracer``.

Reference surface being mirrored (names, shapes, dtypes, in-place mutation contract):
``editable_gauss_refl/cuda/csrc/raytracer.cpp:24-218`` and the holders in ``csrc/core/*.h``
(camera.h:43-77, config.h:31-102, framebuffer.h:159-288, gaussians.h:30-135, metadata.h:12-38,
stats.h:11-35, per_pixel_linked_list.h:80-137).  PyTorch is used for device memory and streams only; every
device-side operation goes through the C ABI of ``include/egr_b200.h``.  There is no CPU fallback: without the
CUDA extension or without a GPU, construction raises.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from . import _capi

DEFAULT_MAX_BOUNCES = int(os.environ.get("EGR_MAX_BOUNCES", "2"))  # reference: flags.h:5


def _dev():
    if not torch.cuda.is_available():
        raise RuntimeError("egr_b200.Raytracer needs a CUDA device (there is no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


class _Holder:
    """Attribute bag of CUDA tensors; attributes are read-only by name, tensors mutable in place."""

    _fields: tuple = ()

    def _ptr(self, name):
        t = getattr(self, name)
        return t.data_ptr()


class CameraDataHolder(_Holder):  # camera.h:43-77
    def __init__(self, dev):
        f32 = dict(dtype=torch.float32, device=dev)
        self.origin = torch.zeros(3, **f32)
        self.vertical_fov_radians = torch.zeros(1, **f32)
        self.rotation_c2w = torch.zeros(3, 3, **f32)
        self.rotation_w2c = torch.zeros(3, 3, **f32)
        self.znear = torch.zeros(1, **f32)
        self.zfar = torch.zeros(1, **f32)

    def set_pose(self, c2w_origin: torch.Tensor, c2w_rotation: torch.Tensor) -> None:
        if tuple(c2w_rotation.shape) != (3, 3):
            raise RuntimeError("c2w_rotation must be 3x3")
        if tuple(c2w_origin.shape) != (3,):
            raise RuntimeError("c2w_origin must be 3")
        with torch.no_grad():
            # parts that live in host memory travel as kernel arguments of one tiny launch (egr_camera_set_pose):
            # a `copy_` from a host tensor would block the host until the stream has drained
            host_o = self.origin.is_cuda and c2w_origin.device.type == "cpu"
            host_r = self.origin.is_cuda and c2w_rotation.device.type == "cpu"
            if host_o or host_r:
                cam = _capi.EgrCamera()
                for f in _capi.CAMERA_FIELDS:
                    setattr(cam, f, getattr(self, f).data_ptr())
                o = (C.c_float * 3)(*c2w_origin.detach().to(torch.float32).reshape(-1).tolist()) if host_o else None
                r = (C.c_float * 9)(*c2w_rotation.detach().to(torch.float32).reshape(-1).tolist()) if host_r else None
                _capi.check(_capi.load().egr_camera_set_pose(C.byref(cam), o, r,
                                                             torch.cuda.current_stream(self.origin.device).cuda_stream))
            if not host_r:
                self.rotation_c2w.copy_(c2w_rotation)
                self.rotation_w2c.copy_(c2w_rotation.transpose(0, 1))
            if not host_o:
                self.origin.copy_(c2w_origin)


class ConfigDataHolder(_Holder):  # config.h:31-51 (same defaults)
    def __init__(self, dev, max_bounces):
        def f(v):
            return torch.tensor([v], dtype=torch.float32, device=dev)

        self.exp_power = f(3.0)
        self.alpha_threshold = f(0.005)
        self.transmittance_threshold = f(0.01)
        self.accumulate_samples = torch.tensor([False], dtype=torch.bool, device=dev)
        self.jitter_primary_rays = torch.tensor([True], dtype=torch.bool, device=dev)
        self.num_bounces = torch.tensor([2], dtype=torch.int32, device=dev)
        self.global_scale_factor = f(1.0)
        self.loss_weight_diffuse = f(1.0)
        self.loss_weight_specular = f(1.0)
        self.loss_weight_depth = f(1.0)
        self.loss_weight_normal = f(1.0)
        self.loss_weight_f0 = f(1.0)
        self.loss_weight_roughness = f(1.0)
        self.eps_forward_normalization = f(1e-12)
        self.eps_scale_grad = f(1e-12)
        self.eps_ray_surface_offset = f(0.01)
        self.eps_min_roughness = f(0.01)
        self.reflection_invalid_normal_threshold = f(0.7)
        self.backfacing_invalid_normal_threshold = f(0.9)
        self.backfacing_max_dist = f(0.1)


class FramebufferDataHolder(_Holder):  # framebuffer.h:159-216
    def __init__(self, dev, W, H, S, debug_rays):
        f32 = dict(dtype=torch.float32, device=dev)
        for name, ch in (("rgb", 3), ("depth", 1), ("normal", 3), ("f0", 3), ("roughness", 1), ("transmittance", 1),
                         ("total_transmittance", 1), ("ray_origin", 3), ("ray_direction", 3)):
            setattr(self, "output_" + name, torch.zeros(S, H, W, ch, **f32))
        self.output_final = torch.zeros(1, H, W, 3, **f32)
        self.output_denoised = torch.zeros(1, H, W, 3, **f32)
        for name, ch in (("rgb", 3), ("transmittance", 1), ("total_transmittance", 1), ("depth", 1), ("normal", 3),
                         ("f0", 3), ("roughness", 1)):
            setattr(self, "accumulated_" + name, torch.zeros(S, H, W, ch, **f32))
        self.accumulated_sample_count = torch.zeros(1, dtype=torch.int32, device=dev)
        for name, ch in (("diffuse", 3), ("specular", 3), ("depth", 1), ("normal", 3), ("f0", 3), ("roughness", 1)):
            setattr(self, "target_" + name, torch.zeros(H, W, ch, **f32))
        # extras (not in the reference; extra attributes break no caller): the ray traced at each step
        if debug_rays:
            self.debug_ray_origin = torch.zeros(S, H, W, 3, **f32)
            self.debug_ray_direction = torch.zeros(S, H, W, 3, **f32)
            self.debug_step_executed = torch.zeros(S, H, W, dtype=torch.int32, device=dev)
        else:
            self.debug_ray_origin = self.debug_ray_direction = self.debug_step_executed = None


# column of each gradient inside the flat [N, 24] record (include/egr_b200.h, EgrGaussians)
GRAD_COLUMNS = dict(mean=(0, 3), opacity=(3, 1), scale=(4, 3), roughness=(7, 1), rotation=(8, 4), rgb=(12, 3),
                    total_weight=(15, 1), normal=(16, 3), f0=(20, 3))
PARAM_CHANNELS = dict(rgb=3, normal=3, f0=3, roughness=1, opacity=1, scale=3, mean=3, rotation=4)


# Allocator of the flat gradient record.  A multi-GPU job may replace it (distributed.use_registered_gradient_records)
# so that the one buffer the ranks all-reduce comes from NCCL's own allocator and is registered with the communicator.
def _default_record_allocator(n, dev):
    return torch.zeros(n, _capi.EGR_GRAD_STRIDE, dtype=torch.float32, device=dev)


GRAD_RECORD_ALLOCATOR = _default_record_allocator


class GaussianDataHolder(_Holder):  # gaussians.h:30-135
    def __init__(self, dev):
        self._dev = dev
        self.count = 1
        for name, ch in PARAM_CHANNELS.items():
            setattr(self, name, torch.zeros(1, ch, dtype=torch.float32, device=dev))
        self._bind_grads(1)

    def _bind_grads(self, n):
        # ONE flat buffer; the reference's dL_d* tensors and total_weight are strided views of it
        self.grad_records = GRAD_RECORD_ALLOCATOR(n, self._dev)
        for name, (c0, ch) in GRAD_COLUMNS.items():
            view = self.grad_records[:, c0:c0 + ch]
            if name == "total_weight":
                self.total_weight = view
            else:
                setattr(self, "dL_d" + name, view)
                getattr(self, name).grad = view  # gaussians.h:53-62

    def resize(self, n):  # gaussians.h:64-86
        with torch.no_grad():
            self.count = int(n)
            for name, ch in PARAM_CHANNELS.items():
                t = getattr(self, name)
                t.grad = None
                t.resize_(n, ch)
            self._bind_grads(int(n))


class MetaDataHolder(_Holder):  # metadata.h:12-38
    def __init__(self, dev, W, H):
        self.grads_enabled = torch.ones(1, dtype=torch.bool, device=dev)
        self.total_num_calls = torch.zeros(1, dtype=torch.int32, device=dev)
        self.random_seeds = torch.randint(0, 1000000000, (H, W, 1), dtype=torch.int32, device=dev)


class StatsDataHolder(_Holder):  # stats.h:11-35 (+3 planes feeding the roofline formula)
    def __init__(self, dev, W, H):
        i32 = dict(dtype=torch.int32, device=dev)
        self.num_accumulated_per_pixel = torch.zeros(H, W, **i32)
        self.num_traversed_per_pixel = torch.zeros(H, W, **i32)
        self.num_hits_per_pixel = torch.zeros(H, W, **i32)
        self.num_composited_per_pixel = torch.zeros(H, W, **i32)
        self.num_steps_per_pixel = torch.zeros(H, W, **i32)
        self.hitlog_status = torch.zeros(4, **i32)  # chunks used, overflow (this call), overflow (sticky), -


class PPLLDataHolder(_Holder):
    """Placeholder with the reference's attribute names (per_pixel_linked_list.h:80-137).

    The reference logs every hit into a global per-pixel linked list (36 B/entry, 180 M + 120 M entries); this
    implementation keeps hits in a register k-buffer and a compact hit log instead, and nothing outside the
    reference's own device code ever read these tensors, so only ``total_hits`` / ``head_per_pixel`` keep their
    full shape.
    """

    def __init__(self, dev, W, H, requested_capacity):
        i32 = dict(dtype=torch.int32, device=dev)
        f32 = dict(dtype=torch.float32, device=dev)
        self.requested_capacity = int(requested_capacity)
        self.head_per_pixel = torch.full((H, W), self.NULL_PTR(), **i32)
        self.total_hits = torch.zeros(1, **i32)
        self.gaussian_ids = torch.zeros(1, **i32)
        self.previous_entries = torch.zeros(1, **i32)
        self.distances = torch.zeros(1, **f32)
        self.gaussvals = torch.zeros(1, **f32)
        self.alphas = torch.zeros(1, **f32)
        self.transmittances = torch.zeros(1, **f32)
        self.local_hits = torch.zeros(1, 3, **f32)

    @staticmethod
    def NULL_PTR():
        return 2 << 29  # per_pixel_linked_list.h:4


class Raytracer:
    """Drop-in for ``torch.classes.raytracer.Raytracer(W, H, N, ppll_fwd, ppll_bwd)``."""

    def __init__(self, image_width, image_height, num_gaussians, forward_ppll_size=180_000_000,
                 backward_ppll_size=120_000_000, max_bounces=None, debug_rays=False):
        self._lib = _capi.load()
        dev = _dev()
        self._make_holders(dev, image_width, image_height, num_gaussians, forward_ppll_size, backward_ppll_size,
                           max_bounces, debug_rays)
        self._pixel_state = torch.empty(self._lib.egr_pixel_state_bytes(self.width, self.height, self._S),
                                        dtype=torch.uint8, device=dev)
        self.rebuild_bvh()  # raytracer.cpp:70-71 (BVHWrapper ctor builds the TLAS)

    def _make_holders(self, dev, image_width, image_height, num_gaussians, forward_ppll_size, backward_ppll_size,
                      max_bounces, debug_rays):
        """The tensor holders of raytracer.cpp:45-68 (everything the callers read or mutate in place)."""
        self._dev = dev
        self.width, self.height = int(image_width), int(image_height)
        self._max_bounces = DEFAULT_MAX_BOUNCES if max_bounces is None else int(max_bounces)
        S = self._max_bounces + 1
        if not (1 <= S <= _capi.EGR_MAX_STEPS):
            raise RuntimeError(f"max_bounces must be in [0, {_capi.EGR_MAX_STEPS - 1}]")
        self._S = S
        self.camera_data = CameraDataHolder(dev)
        self.config_data = ConfigDataHolder(dev, self._max_bounces)
        self.framebuffer_data = FramebufferDataHolder(dev, self.width, self.height, S, debug_rays)
        self.gaussian_data = GaussianDataHolder(dev)
        self.meta_data = MetaDataHolder(dev, self.width, self.height)
        self.stats_data = StatsDataHolder(dev, self.width, self.height)
        self.ppll_forward_data = PPLLDataHolder(dev, self.width, self.height, forward_ppll_size)
        self.ppll_backward_data = PPLLDataHolder(dev, self.width, self.height, backward_ppll_size)
        if num_gaussians > 0:
            self.gaussian_data.resize(int(num_gaussians))
        self._hitlog_capacity_hits = int(backward_ppll_size)
        self._hitlog = None
        self._bvh_storage = None
        self._bvh_count = -1
        self._params = None
        self._ray_override = None

    # ---- constants (raytracer.cpp:145-148) -------------------------------------------------------------
    def MAX_BOUNCES(self=None):
        return DEFAULT_MAX_BOUNCES if self is None or not isinstance(self, Raytracer) else self._max_bounces

    @staticmethod
    def MAX_ALPHA():
        return float(_capi.load().egr_max_alpha())

    @staticmethod
    def ROUGHNESS_DOWNWEIGHT_GRAD():
        return bool(_capi.load().egr_roughness_downweight_grad())

    @staticmethod
    def ROUGHNESS_DOWNWEIGHT_GRAD_POWER():
        return float(_capi.load().egr_roughness_downweight_grad_power())

    # ---- getters (raytracer.cpp:131-142) -----------------------------------------------------------------
    def get_camera(self):
        return self.camera_data

    def get_config(self):
        return self.config_data

    def get_framebuffer(self):
        return self.framebuffer_data

    def get_gaussians(self):
        return self.gaussian_data

    def get_metadata(self):
        return self.meta_data

    def get_stats(self):
        return self.stats_data

    def get_ppll_forward_data(self):
        return self.ppll_forward_data

    def get_ppll_backward_data(self):
        return self.ppll_backward_data

    # ---- introspection (raytracer.cpp:151-204) -----------------------------------------------------------
    def describe_output_buffers(self):
        return [("output_rgb", 3), ("output_depth", 1), ("output_normal", 3), ("output_f0", 3),
                ("output_roughness", 1), ("output_transmittance", 1), ("output_total_transmittance", 1),
                ("output_brdf", 3), ("output_ray_origin", 3), ("output_ray_direction", 3), ("output_final", 3)]

    def describe_accumulation_buffers(self):
        return [("accumulated_rgb", 3), ("accumulated_transmittance", 1), ("accumulated_total_transmittance", 1),
                ("accumulated_depth", 1), ("accumulated_normal", 3), ("accumulated_f0", 3),
                ("accumulated_roughness", 1)]

    def describe_target_buffers(self):
        return [("target_depth", 1), ("target_normal", 3), ("target_f0", 3), ("target_roughness", 1)]

    def describe_gaussian_attributes(self):
        return [("gaussian_rgb", 3), ("gaussian_normal", 3), ("gaussian_f0", 3), ("gaussian_roughness", 1),
                ("gaussian_opacity", 1), ("gaussian_scale", 3), ("gaussian_mean", 3), ("gaussian_rotation", 4)]

    # ---- plumbing -------------------------------------------------------------------------------------------
    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self._dev).cuda_stream)

    def _gaussians_struct(self):
        g = _capi.EgrGaussians()
        gd = self.gaussian_data
        g.count = gd.count
        for f in _capi.GAUSSIAN_FIELDS:
            setattr(g, f, getattr(gd, f).data_ptr())
        g.grad_records = gd.grad_records.data_ptr()
        return g

    def _config_struct(self):
        c = _capi.EgrConfig()
        for f in _capi.CONFIG_FIELDS:
            setattr(c, f, getattr(self.config_data, f).data_ptr())
        return c

    def _build_params(self):
        """Capture the raw pointers once (raytracer.cpp:59-68); refreshed by resize()/rebuild_bvh()."""
        p = _capi.EgrParams()
        p.image_width, p.image_height, p.num_steps = self.width, self.height, self._S
        p.tile_rows = getattr(self, "_tile_rows", 0)
        p.early_stop_transmittance = float(getattr(self, "_early_stop", 0.0))
        for f in _capi.CAMERA_FIELDS:
            setattr(p.camera, f, getattr(self.camera_data, f).data_ptr())
        p.config = self._config_struct()
        for f in _capi.FRAMEBUFFER_FIELDS:
            setattr(p.framebuffer, f, getattr(self.framebuffer_data, f).data_ptr())
        p.gaussians = self._gaussians_struct()
        for f in _capi.METADATA_FIELDS:
            setattr(p.metadata, f, getattr(self.meta_data, f).data_ptr())
        for f in _capi.STATS_FIELDS:
            setattr(p.stats, f, getattr(self.stats_data, f).data_ptr())
        p.bvh_storage = self._bvh_storage.data_ptr() if self._bvh_storage is not None else None
        p.bvh_storage_bytes = self._bvh_storage.numel() if self._bvh_storage is not None else 0
        p.pixel_state, p.pixel_state_bytes = self._pixel_state.data_ptr(), self._pixel_state.numel()
        if self._hitlog is not None:
            p.hitlog, p.hitlog_bytes = self._hitlog.data_ptr(), self._hitlog.numel()
        fb = self.framebuffer_data
        if fb.debug_ray_origin is not None:
            p.used_ray_origin = fb.debug_ray_origin.data_ptr()
            p.used_ray_direction = fb.debug_ray_direction.data_ptr()
            p.step_executed = fb.debug_step_executed.data_ptr()
        ov = getattr(self, "_ray_override", None)
        if ov is not None:
            p.override_ray_origin, p.override_ray_direction = ov[0].data_ptr(), ov[1].data_ptr()
            p.override_step_executed = ov[2].data_ptr()
        self._params = p

    # ---- methods (raytracer.cpp:81-120) ------------------------------------------------------------------
    def _ensure_hitlog(self, grads):
        if grads and self._hitlog is None:
            nbytes = self._lib.egr_hitlog_bytes_for_hits(self._hitlog_capacity_hits)
            self._hitlog = torch.empty(max(int(nbytes), 1), dtype=torch.uint8, device=self._dev)
            self._build_params()

    def _check_bvh_current(self):
        """resize() changes the Gaussian count; the tree (topology, array offsets) still describes the old count
        until rebuild_bvh() runs, so refitting or tracing against it would read a foreign layout."""
        if self._bvh_count != self.gaussian_data.count:
            raise RuntimeError(f"resize() must be followed by rebuild_bvh(): the BVH was built for {self._bvh_count} "
                               f"Gaussians, the holders now have {self.gaussian_data.count}")

    def hitlog_overflowed(self, clear: bool = False) -> bool:
        """True when a grad-mode raytrace() since construction (or the last ``clear``) ran out of hit-log space: the
        affected 32-ray groups got NO backward (their gradients are missing).  Reads ``stats.hitlog_status[2]`` --
        a host synchronisation, so call it where one happens anyway (pruning interval, logging, end of a bench)."""
        st = self.stats_data.hitlog_status
        over = bool(int(st[2].item()))
        if over and clear:
            st[2] = 0
        return over

    def raytrace(self):
        grads = torch.is_grad_enabled()  # metadata.h:29
        self._check_bvh_current()
        self._ensure_hitlog(grads)
        _capi.check(self._lib.egr_raytrace(C.byref(self._params), int(grads), self._stream()))

    def raytrace_phases(self, phase_mask):
        """raytrace() split into its launches (not in the reference; used by bench.py to time single kernels)."""
        grads = torch.is_grad_enabled()
        self._check_bvh_current()
        self._ensure_hitlog(grads)
        _capi.check(self._lib.egr_raytrace_phases(C.byref(self._params), int(grads), int(phase_mask), self._stream()))

    def denoise(self, passthrough: bool = False):
        """``output_final -> output_denoised``.  The reference runs the OptiX AI denoiser here, a proprietary network
        inside the driver that is unavailable on this box; this is a headless stand-in, NOT that network: an
        edge-avoiding a-trous wavelet filter guided by the primary-step normal and depth (``egr_denoise``,
        include/egr_b200.h).  ``passthrough=True`` copies ``output_final`` unchanged instead."""
        fn = self._lib.egr_denoise_passthrough if passthrough else self._lib.egr_denoise
        _capi.check(fn(C.byref(self._params), self._stream()))

    def reset_accumulators(self):
        _capi.check(self._lib.egr_reset_accumulators(C.byref(self._params), self._stream()))

    def update_bvh(self):
        """Refit (bvh_wrapper.h:32-59): re-pack geometry records and refresh every box; no host sync."""
        self._check_bvh_current()
        g = self._gaussians_struct()
        cfg = self._config_struct()
        _capi.check(self._lib.egr_bvh_refit(C.byref(g), C.byref(cfg), self._bvh_storage.data_ptr(),
                                            self._bvh_storage.numel(), self._stream()))

    def rebuild_bvh(self):
        """Full LBVH rebuild (bvh_wrapper.h:24-30): Morton sort + topology + refit."""
        n = self.gaussian_data.count
        need = int(self._lib.egr_bvh_storage_bytes(n))
        if self._bvh_storage is None or self._bvh_storage.numel() < need:
            self._bvh_storage = torch.empty(max(need, 1), dtype=torch.uint8, device=self._dev)
        scratch = torch.empty(max(int(self._lib.egr_bvh_scratch_bytes(n)), 1), dtype=torch.uint8, device=self._dev)
        g = self._gaussians_struct()
        cfg = self._config_struct()
        _capi.check(self._lib.egr_bvh_build(C.byref(g), C.byref(cfg), self._bvh_storage.data_ptr(),
                                            self._bvh_storage.numel(), scratch.data_ptr(), scratch.numel(),
                                            self._stream()))
        scratch.record_stream(torch.cuda.current_stream(self._dev))
        self._bvh_count = n
        self._build_params()

    def resize(self, new_num_gaussians):
        self.gaussian_data.resize(int(new_num_gaussians))
        self._build_params()

    # ---- test / debug helpers (not in the reference) ---------------------------------------------------------
    def export_parameters_from(self, sources: dict) -> bool:
        """holder.copy_(source) for the eight parameters in one launch (egr_params_export).  Returns False without
        doing anything when a source is not a contiguous float32 CUDA tensor of the holder's shape (the caller then
        falls back to the reference's eight copy_ calls, gaussian_raytracer.py:41-51)."""
        g = self.gaussian_data
        for f in _capi.GRAD_TARGET_FIELDS:
            s_, d_ = sources.get(f), getattr(g, f)
            if (s_ is None or not s_.is_cuda or s_.dtype != torch.float32 or not s_.is_contiguous()
                    or s_.shape != d_.shape or s_.device != d_.device):
                return False
        src, dst = _capi.EgrGradTargets(), _capi.EgrGradTargets()
        for f in _capi.GRAD_TARGET_FIELDS:
            setattr(src, f, sources[f].data_ptr())
            setattr(dst, f, getattr(g, f).data_ptr())
        _capi.check(self._lib.egr_params_export(C.byref(src), C.byref(dst), int(g.count), self._stream()))
        return True

    def accumulate_gradients_into(self, targets: dict):
        """targets[name] += gradient of `name` for the eight parameters, one launch (egr_gradients_accumulate).
        Not in the reference class: its caller does this with eight `add_` (gaussian_raytracer.py:53-62)."""
        t = _capi.EgrGradTargets()
        for f in _capi.GRAD_TARGET_FIELDS:
            setattr(t, f, targets[f].data_ptr() if targets.get(f) is not None else None)
        g = self.gaussian_data
        _capi.check(self._lib.egr_gradients_accumulate(g.grad_records.data_ptr(), int(g.grad_records.shape[0]),
                                                       C.byref(t), self._stream()))

    def set_early_stop(self, transmittance_threshold: float = 0.0):
        """Opt-in early-stop mode (not in the reference; 0 = exact mode, the default).  With a threshold > 0 a ray stops
        collecting hits once its composite has terminated and the running total transmittance is below the threshold:
        ``output_total_transmittance`` is then within the threshold of the exact value, the composited radiance and
        G-buffers are unchanged.  Meant for opaque scenes (trained models, novel-view rendering), where most of a ray's
        Gaussians lie behind the surface it ends on."""
        if not (0.0 <= float(transmittance_threshold) < 1.0):
            raise RuntimeError("early-stop threshold must be in [0, 1)")
        self._early_stop = float(transmittance_threshold)
        self._build_params()

    def set_pixel_band(self, row_begin: int | None = None, row_end: int | None = None):
        """Restrict the following raytrace() calls to the pixel rows [row_begin, row_end) (``row_begin`` a multiple of 4,
        ``row_end`` a multiple of 4 or the image height); ``None`` = the whole image again.  Not in the reference: it
        is how ONE view is split over several GPUs (``distributed.band_for_rank``): the bands of all ranks tile the
        single-GPU image bit for bit, and the all-reduced gradient record equals the single-GPU gradients."""
        if row_begin is None:
            self._tile_rows = 0
        else:
            if row_begin % 4 or (row_end % 4 and row_end != self.height) or not (0 <= row_begin < row_end <= self.height):
                raise RuntimeError("pixel band: rows must be multiples of 4 inside the image")
            self._tile_rows = (row_begin // 4) | (((row_end + 3) // 4) << 16)
        self._build_params()

    def set_ray_override(self, ray_origin=None, ray_direction=None, step_executed=None):
        """Test input: rays of the steps >= 1 ([S,H,W,3] each) and which steps run ([S,H,W] int32), e.g. the bounce
        rays another implementation logged.  ``None`` removes the override."""
        if ray_origin is None:
            self._ray_override = None
        else:
            S, Hh, W = self._S, self.height, self.width
            f = dict(dtype=torch.float32, device=self._dev)
            self._ray_override = (torch.as_tensor(ray_origin, **f).reshape(S, Hh, W, 3).contiguous(),
                                  torch.as_tensor(ray_direction, **f).reshape(S, Hh, W, 3).contiguous(),
                                  torch.as_tensor(step_executed, dtype=torch.int32, device=self._dev).reshape(S, Hh, W).contiguous())
        self._build_params()

    def export_bvh(self):
        """Copies of (sorted_ids [n], records [n,16], nodes [2n-1,8], parents [2n-1])."""
        self._check_bvh_current()
        n = self.gaussian_data.count
        dev = self._dev
        ids = torch.empty(n, dtype=torch.int32, device=dev)
        rec = torch.empty(n, 16, dtype=torch.float32, device=dev)
        nodes = torch.empty(max(2 * n - 1, 0), 8, dtype=torch.float32, device=dev)
        par = torch.empty(max(2 * n - 1, 0), dtype=torch.int32, device=dev)
        _capi.check(self._lib.egr_bvh_export(self._bvh_storage.data_ptr(), self._bvh_storage.numel(), n,
                                             ids.data_ptr(), rec.data_ptr(), nodes.data_ptr(), par.data_ptr(),
                                             self._stream()))
        return ids, rec, nodes, par

    def export_wide_bvh(self):
        """Copies of the 4-wide tree of the bounce walk: (wide [num_wide,32] float32, wide_slot [2n-1] int32)."""
        self._check_bvh_current()
        n = self.gaussian_data.count
        cap = int(self._lib.egr_bvh_wide_capacity(n))
        wide = torch.empty(cap, 32, dtype=torch.float32, device=self._dev)
        slot = torch.empty(max(2 * n - 1, 0), dtype=torch.int32, device=self._dev)
        num = torch.zeros(1, dtype=torch.int32, device=self._dev)
        _capi.check(self._lib.egr_bvh_export_wide(self._bvh_storage.data_ptr(), self._bvh_storage.numel(), n,
                                                  wide.data_ptr(), slot.data_ptr(), num.data_ptr(), self._stream()))
        k = int(num.item())
        return wide[:k], slot

    def export_hit_lists(self, step, cap):
        """Per-pixel composited Gaussian ids / distances of the last grad-mode raytrace() at ``step``."""
        P = self.width * self.height
        ids = torch.full((P, cap), -1, dtype=torch.int32, device=self._dev)
        ts = torch.zeros(P, cap, dtype=torch.float32, device=self._dev)
        counts = torch.zeros(P, dtype=torch.int32, device=self._dev)
        _capi.check(self._lib.egr_hitlog_export(C.byref(self._params), int(step), int(cap), ids.data_ptr(),
                                                ts.data_ptr(), counts.data_ptr(), self._stream()))
        return ids, ts, counts
