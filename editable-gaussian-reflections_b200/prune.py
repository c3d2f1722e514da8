"""Pruning as one stream-ordered pipeline.

The reference builds the mask in ``train.py:238-240`` (``total_weight / pruning_interval < min_weight``) and then
``GaussianModel.prune_points`` / ``_prune_optimizer`` (``scene/gaussian_model.py:479-532``) runs ``tensor[mask]``
about thirty times -- the 8 parameters, their Adam ``exp_avg`` / ``exp_avg_sq`` and six side arrays -- each call
compacting one tensor and synchronising the host to learn the new size.  Here: one mask launch, one scan and ONE
compaction launch for all arrays (``egr_compact_rows``), then a single 4-byte read of the new count.
Results are identical to ``tensor[keep]`` (stable order, bit-exact copies).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _capi


def keep_mask_from_weight(total_weight: torch.Tensor, pruning_interval: float, min_weight: float) -> torch.Tensor:
    """``~(total_weight / pruning_interval < min_weight).squeeze(1)`` of train.py:239 as a uint8 mask (1 = keep)."""
    if not total_weight.is_cuda:
        raise RuntimeError("keep_mask_from_weight expects a CUDA tensor (there is no CPU fallback)")
    tw = total_weight.detach().reshape(-1)
    if tw.dtype != torch.float32 or not tw.is_contiguous():
        tw = tw.to(torch.float32).contiguous()  # (the strided view of the flat gradient record lands here)
    keep = torch.empty(tw.numel(), dtype=torch.uint8, device=tw.device)
    with torch.cuda.device(tw.device):
        _capi.check(_capi.load().egr_prune_keep_mask(tw.data_ptr(), tw.numel(), float(pruning_interval), float(min_weight),
                                                     keep.data_ptr(), torch.cuda.current_stream().cuda_stream))
    return keep


def drop_near_cameras(keep: torch.Tensor, xyz: torch.Tensor, camera_centers: torch.Tensor, znear: torch.Tensor) -> torch.Tensor:
    """Clears ``keep[i]`` for every point closer to any camera centre than that camera's ``znear``
    (``Scene.select_points_to_prune_near_cameras``, scene/scene.py:88-105, one pass instead of one per camera).
    ``keep``: uint8 [N] (in place), ``xyz`` [N,3], ``camera_centers`` [C,3], ``znear`` [C]: float32 CUDA tensors."""
    if not (keep.is_cuda and xyz.is_cuda and camera_centers.is_cuda and znear.is_cuda):
        raise RuntimeError("drop_near_cameras expects CUDA tensors (there is no CPU fallback)")
    if keep.dtype != torch.uint8 or not keep.is_contiguous() or keep.numel() != xyz.shape[0]:
        raise RuntimeError("keep must be a contiguous uint8 [N] tensor")
    x = xyz.detach().to(torch.float32).contiguous()
    c = camera_centers.detach().to(torch.float32).contiguous().reshape(-1, 3)
    z = znear.detach().to(torch.float32).contiguous().reshape(-1)
    if z.numel() != c.shape[0]:
        raise RuntimeError("one znear per camera centre")
    with torch.cuda.device(keep.device):
        _capi.check(_capi.load().egr_prune_near_cameras(x.data_ptr(), x.shape[0], c.data_ptr(), z.data_ptr(), c.shape[0],
                                                        keep.data_ptr(), torch.cuda.current_stream().cuda_stream))
    return keep


def compact(tensors, keep: torch.Tensor):
    """``[t[keep] for t in tensors]`` for CUDA tensors that share their first dimension, with one launch and one host
    read.  ``keep``: bool or uint8 ``[N]``.  Element sizes must make every row a multiple of 4 bytes (<= 64)."""
    tensors = list(tensors)
    if not keep.is_cuda or any(not t.is_cuda for t in tensors):
        raise RuntimeError("compact expects CUDA tensors (there is no CPU fallback)")
    n = int(keep.numel())
    if len(tensors) > _capi.EGR_COMPACT_MAX_ARRAYS:
        raise RuntimeError(f"at most {_capi.EGR_COMPACT_MAX_ARRAYS} arrays per call")
    k8 = keep.reshape(-1)
    k8 = k8.view(torch.uint8) if k8.dtype == torch.bool else k8.to(torch.uint8)
    k8 = k8.contiguous()
    srcs, dsts, rows = [], [], []
    for t in tensors:
        if t.shape[0] != n:
            raise RuntimeError("every tensor must have keep.numel() rows")
        s = t.detach().contiguous()
        srcs.append(s), dsts.append(torch.empty_like(s)), rows.append(s[0].numel() * s.element_size() if n else 4)
    lib = _capi.load()
    count = torch.zeros(1, dtype=torch.int32, device=keep.device)
    m = len(tensors)
    src_p = (C.c_void_p * max(m, 1))(*[s.data_ptr() for s in srcs])
    dst_p = (C.c_void_p * max(m, 1))(*[d.data_ptr() for d in dsts])
    row_b = (C.c_uint32 * max(m, 1))(*rows)
    with torch.cuda.device(keep.device):
        scratch = torch.empty(int(lib.egr_compact_scratch_bytes(n)) + 256, dtype=torch.uint8, device=keep.device)
        _capi.check(lib.egr_compact_rows(k8.data_ptr(), n, m, src_p, dst_p, row_b, scratch.data_ptr(),
                                         scratch.numel(), count.data_ptr(), torch.cuda.current_stream().cuda_stream))
    c = int(count.item())  # the one host read of the pipeline
    return [d[:c] for d in dsts], c
