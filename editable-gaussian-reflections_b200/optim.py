"""The host step around the tracer as one launch.

Per training iteration the reference runs ``_import_param_gradients`` (8 ``add_``), the scale decay
(``train.py:224-226``), ``torch.optim.Adam.step`` over 8 parameter groups (``scene/gaussian_model.py:333-338``,
eps 1e-15), two ``zero_grad`` passes and three clamps (``train.py:247-254``).  ``FusedAdam.step`` does all of it with
``egr_adam_step``: the gradient is read straight from the tracer's flat ``[N,24]`` record, parameters and moments are
updated in place, the record is zeroed.  Same formulas as ``torch.optim.Adam`` (no weight decay, no amsgrad).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _capi

NAMES = tuple(_capi.GRAD_TARGET_FIELDS)  # mean, opacity, scale, roughness, rotation, rgb, normal, f0
CHANNELS = dict(mean=3, opacity=1, scale=3, roughness=1, rotation=4, rgb=3, normal=3, f0=3)


def _targets(tensors: dict) -> "_capi.EgrGradTargets":
    t = _capi.EgrGradTargets()
    for f in NAMES:
        setattr(t, f, tensors[f].data_ptr() if tensors.get(f) is not None else None)
    return t


class FusedAdam:
    """``params``: name -> contiguous float32 CUDA tensor ``[N, ch]`` (the model's raw parameters; any subset of the
    eight names).  Moments are created like ``torch.optim.Adam`` creates them (zeros) and live in ``exp_avg`` /
    ``exp_avg_sq`` (replace them, e.g. after pruning, with :meth:`rebind`)."""

    def __init__(self, params: dict, betas=(0.9, 0.999), eps: float = 1e-15):
        self.betas, self.eps, self.step_count = (float(betas[0]), float(betas[1])), float(eps), 0
        self.rebind(params)

    def rebind(self, params: dict, exp_avg: dict | None = None, exp_avg_sq: dict | None = None) -> None:
        n = None
        for k, p in params.items():
            if k not in CHANNELS:
                raise RuntimeError(f"unknown parameter {k!r}")
            if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous() or p.dim() != 2 or p.shape[1] != CHANNELS[k]:
                raise RuntimeError(f"{k}: expected a contiguous float32 CUDA tensor [N,{CHANNELS[k]}] (no CPU fallback)")
            n = p.shape[0] if n is None else n
            if p.shape[0] != n:
                raise RuntimeError("all parameters must have the same number of rows")
        self.params = dict(params)
        self.exp_avg = {k: (exp_avg[k] if exp_avg else torch.zeros_like(p)) for k, p in params.items()}
        self.exp_avg_sq = {k: (exp_avg_sq[k] if exp_avg_sq else torch.zeros_like(p)) for k, p in params.items()}
        self.count = int(n or 0)

    @torch.no_grad()
    def step(self, grad_records: torch.Tensor, lr: dict, scale_decay: float = 1.0, clamp: bool = True,
             zero_grads: bool = True, zero_total_weight: bool = False) -> None:
        """One optimiser step from the flat gradient record ``[N,24]`` (``get_gaussians().grad_records``)."""
        if (not grad_records.is_cuda or grad_records.dtype != torch.float32 or not grad_records.is_contiguous()
                or tuple(grad_records.shape) != (self.count, _capi.EGR_GRAD_STRIDE)):
            raise RuntimeError("grad_records must be the contiguous float32 CUDA [N,24] record of the same N")
        self.step_count += 1
        o = _capi.EgrAdam()
        o.lr = (C.c_double * 8)(*[float(lr.get(k, 0.0)) for k in NAMES])
        o.beta1, o.beta2, o.eps, o.step = self.betas[0], self.betas[1], self.eps, self.step_count
        o.scale_decay, o.clamp, o.zero_grads, o.zero_total_weight = float(scale_decay), int(clamp), int(zero_grads), int(zero_total_weight)
        P, M, V = _targets(self.params), _targets(self.exp_avg), _targets(self.exp_avg_sq)
        with torch.cuda.device(grad_records.device):
            _capi.check(_capi.load().egr_adam_step(grad_records.data_ptr(), self.count, C.byref(P), C.byref(M), C.byref(V),
                                                   C.byref(o), torch.cuda.current_stream().cuda_stream))
