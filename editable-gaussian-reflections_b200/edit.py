"""The edit path on the device.

Mirrors what ``EditableGaussianModel`` (``scene/editable_gaussian_model.py``) does with N-sized torch / kornia op
chains and boolean-mask indexing: the selection masks of ``make_editable`` (``:30-50``), the translate / scale /
rotate of the selected Gaussians in ``get_xyz`` / ``_get_scaling`` / ``_get_rotation`` (``:201-279``) and
``remove_object`` (``:324-328``).  Here each is one launch through the C ABI, in place on CUDA tensors (typically the
tracer's own ``get_gaussians()`` tensors right before ``update_bvh()``), with no host synchronisation.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np
import torch

from . import _capi


def axis_angle_to_matrix_and_quaternion(rx_deg: float, ry_deg: float, rz_deg: float):
    """The edit rotation of the reference: ``deg2rad([rotate_x, rotate_y, rotate_z])`` taken as an axis-angle vector
    (``kornia.geometry.axis_angle_to_rotation_matrix``, editable_gaussian_model.py:223-225): Rodrigues' formula.
    Returns (3x3 float32 row-major matrix, unit quaternion (r,x,y,z) float32) -- host scalars, computed once per edit."""
    v = np.deg2rad(np.array([rx_deg, ry_deg, rz_deg], np.float64))
    th = float(np.linalg.norm(v))
    if th < 1e-12:
        return np.eye(3, dtype=np.float32), np.array([1, 0, 0, 0], np.float32)
    k = v / th
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    R = np.eye(3) + math.sin(th) * K + (1 - math.cos(th)) * (K @ K)
    q = np.concatenate([[math.cos(th / 2)], math.sin(th / 2) * k])
    return R.astype(np.float32), q.astype(np.float32)


def _f3(v):
    return (C.c_float * 3)(*[float(x) for x in v])


def _select(fn_name: str, xyz: torch.Tensor, bmin, bmax) -> torch.Tensor:
    if not xyz.is_cuda:
        raise RuntimeError("selection expects a CUDA tensor (there is no CPU fallback)")
    if xyz.dim() != 2 or xyz.shape[1] != 3 or xyz.dtype != torch.float32 or not xyz.is_contiguous():
        raise RuntimeError("selection expects a contiguous float32 [N,3] tensor")
    n = int(xyz.shape[0])
    mask = torch.empty(n, dtype=torch.uint8, device=xyz.device)
    with torch.cuda.device(xyz.device):
        _capi.check(getattr(_capi.load(), fn_name)(xyz.data_ptr(), n, _f3(bmin), _f3(bmax), mask.data_ptr(),
                                                   torch.cuda.current_stream().cuda_stream))
    return mask


def select_box(xyz: torch.Tensor, bmin, bmax) -> torch.Tensor:
    """``(xyz - min >= 0).all(-1) & (xyz - max <= 0).all(-1)`` (editable_gaussian_model.py:45-48) as uint8 [N]."""
    return _select("egr_select_box", xyz, bmin, bmax)


def select_cylinder(xyz: torch.Tensor, bmin, bmax) -> torch.Tensor:
    """The ``"cyl"`` selection: elliptic cylinder inscribed in the box (editable_gaussian_model.py:30-44)."""
    return _select("egr_select_cylinder", xyz, bmin, bmax)


def apply_edit(mask: torch.Tensor, *, mean=None, scale=None, rotation=None, opacity=None, translate=(0.0, 0.0, 0.0),
               center=(0.0, 0.0, 0.0), scale_factor: float = 1.0, rotate_deg=(0.0, 0.0, 0.0), kill: bool = False) -> None:
    """In place on the selected rows: the transform of ``get_xyz`` / ``_get_scaling`` / ``_get_rotation`` for one edit
    key, and ``remove_object`` when ``kill``.  ``center`` is the centre of the object's bounding box; ``scale`` holds
    raw log-scales, ``rotation`` raw quaternions (r,x,y,z).  Any of the four tensors may be omitted."""
    if not mask.is_cuda:
        raise RuntimeError("apply_edit expects CUDA tensors (there is no CPU fallback)")
    n = int(mask.numel())
    m8 = mask.reshape(-1)
    m8 = (m8.view(torch.uint8) if m8.dtype == torch.bool else m8.to(torch.uint8)).contiguous()
    ptrs = []
    for t, ch in ((mean, 3), (scale, 3), (rotation, 4), (opacity, 1)):
        if t is None:
            ptrs.append(None)
            continue
        if not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous() or t.numel() != n * ch:
            raise RuntimeError("apply_edit expects contiguous float32 CUDA tensors with one row per mask entry")
        ptrs.append(t.data_ptr())
    R, q = axis_angle_to_matrix_and_quaternion(*rotate_deg)
    e = _capi.EgrEdit()
    e.translate, e.center, e.scale = _f3(translate), _f3(center), float(scale_factor)
    e.rotation = (C.c_float * 9)(*R.reshape(-1).tolist())
    e.quaternion = (C.c_float * 4)(*q.tolist())
    e.kill = int(bool(kill))
    with torch.cuda.device(mask.device), torch.no_grad():
        _capi.check(_capi.load().egr_edit_apply(m8.data_ptr(), n, *ptrs, C.byref(e),
                                                torch.cuda.current_stream().cuda_stream))


def duplicate(tensors, mask: torch.Tensor, add_index: int = -1, add_row=None):
    """``[torch.cat((t, t[mask])) for t in tensors]`` with ``add_row`` added to the appended copy of
    ``tensors[add_index]`` -- ``EditableGaussianModel.duplicate_object`` (editable_gaussian_model.py:284-322: the
    means get ``offset + delta_xyz``) -- as one scan + one launch for all arrays and one host read (the number of
    selected rows, needed to size the results).  Returns (list of new tensors, number of appended rows)."""
    tensors = list(tensors)
    if not mask.is_cuda or any(not t.is_cuda for t in tensors):
        raise RuntimeError("duplicate expects CUDA tensors (there is no CPU fallback)")
    n = int(mask.numel())
    m8 = mask.reshape(-1)
    m8 = (m8.view(torch.uint8) if m8.dtype == torch.bool else m8.to(torch.uint8)).contiguous()
    extra = int(m8.count_nonzero().item())  # the one host read
    srcs, dsts, rows = [], [], []
    for t in tensors:
        if t.shape[0] != n:
            raise RuntimeError("every tensor must have mask.numel() rows")
        s_ = t.detach().contiguous()
        srcs.append(s_)
        dsts.append(torch.empty((n + extra,) + tuple(s_.shape[1:]), dtype=s_.dtype, device=s_.device))
        rows.append(s_[0].numel() * s_.element_size() if n else 4)
    lib = _capi.load()
    k = len(tensors)
    src_p = (C.c_void_p * max(k, 1))(*[s_.data_ptr() for s_ in srcs])
    dst_p = (C.c_void_p * max(k, 1))(*[d.data_ptr() for d in dsts])
    row_b = (C.c_uint32 * max(k, 1))(*rows)
    add = None
    if add_index >= 0:
        if tensors[add_index].dtype != torch.float32:
            raise RuntimeError("the offset row applies to a float32 array")
        vals = [float(v) for v in add_row]
        if len(vals) * 4 != rows[add_index]:
            raise RuntimeError("add_row must have one value per column of tensors[add_index]")
        add = (C.c_float * len(vals))(*vals)
    count = torch.zeros(1, dtype=torch.int32, device=mask.device)
    with torch.cuda.device(mask.device), torch.no_grad():
        scratch = torch.empty(int(lib.egr_compact_scratch_bytes(n)) + 256, dtype=torch.uint8, device=mask.device)
        _capi.check(lib.egr_duplicate_rows(m8.data_ptr(), n, k, src_p, dst_p, row_b, int(add_index), add,
                                           scratch.data_ptr(), scratch.numel(), count.data_ptr(),
                                           torch.cuda.current_stream().cuda_stream))
    return dsts, extra
