"""``distCUDA2``: mean squared distance of every point to its three nearest neighbours.

Mirror of ``simple_knn._C.distCUDA2`` (reference: ``submodules/simple-knn/ext.cpp``, ``spatial.cu``,
``simple_knn.cu:46-222``), the one native dependency besides the tracer that ``train.py`` needs to start from a
point cloud (``scene/gaussian_model.py:197-200, 246-249`` take ``sqrt`` of it as the initial scale).  Same call
shape: a CUDA float tensor ``[N,3]`` in, a CUDA float tensor ``[N]`` out.  The search runs through the C ABI
(``egr_knn3_mean_dist2``) over the tracer's Morton order + LBVH; there is no CPU fallback.
"""
from __future__ import annotations

import torch

from . import _capi


def distCUDA2(points: torch.Tensor) -> torch.Tensor:
    if not isinstance(points, torch.Tensor) or points.dim() != 2 or points.shape[1] != 3:
        raise RuntimeError("distCUDA2 expects a [N,3] tensor")
    if not points.is_cuda:
        raise RuntimeError("distCUDA2 expects a CUDA tensor (there is no CPU fallback)")
    lib = _capi.load()
    pts = points.detach().to(torch.float32).contiguous()
    n = int(pts.shape[0])
    out = torch.zeros(n, dtype=torch.float32, device=pts.device)  # the reference returns torch.full((P,), 0.0)
    if n == 0:
        return out
    with torch.cuda.device(pts.device):
        scratch = torch.empty(int(lib.egr_knn3_scratch_bytes(n)), dtype=torch.uint8, device=pts.device)
        _capi.check(lib.egr_knn3_mean_dist2(pts.data_ptr(), n, out.data_ptr(), scratch.data_ptr(), scratch.numel(),
                                            torch.cuda.current_stream().cuda_stream))
    return out
