"""TEST INFRASTRUCTURE ONLY (never imported by the product): CPU restatement of the reference's edit path, row f2.

Follows ``editable_gauss_refl/scene/editable_gaussian_model.py`` operation by operation in float32 numpy (torch
runs each op as its own kernel, so every intermediate is rounded to float32):
  select_box       :45-48      select_cylinder  :30-44
  edit_xyz         :201-238    edit_scaling     :240-254    edit_rotation :256-279    remove :324-328
Pinning: the reference module cannot be imported here (it needs ``kornia``, absent from this image, and a CUDA
device).  Everything in it that is plain torch -- the masks, the translate / scale part of get_xyz, the scaling edit,
remove_object and duplicate_object -- is pinned by running the reference's own torch expressions on the GPU next to
the kernels (tests/test_edit.py::test_kernels_match_the_reference_torch_expressions_run_on_the_gpu, bit for bit).
The three kornia conversions it calls are restated from their published definitions (Rodrigues' formula,
unit-quaternion <-> matrix) and pinned to hand-computed known answers (quarter turns about each axis, a 120-degree
turn about (1,1,1), the composition order R_edit @ R(q): test_rotation_conversions_known_answers); tests compare
rotations as matrices (a unit quaternion and its negative are the same rotation) within a stated tolerance.  The 3x3 product of edit_xyz is a GEMM
in the reference (summation order unspecified): compared within tolerance as well.  Everything else is bit-exact.
"""
import numpy as np

f32 = np.float32


def select_box(xyz, bmin, bmax):
    xyz = np.asarray(xyz, f32)
    d1 = xyz - np.asarray(bmin, f32)
    d2 = xyz - np.asarray(bmax, f32)
    return (d1 >= 0).all(-1) & (d2 <= 0).all(-1)


def select_cylinder(xyz, bmin, bmax):
    xyz = np.asarray(xyz, f32)
    lo, hi = np.asarray(bmin, f32), np.asarray(bmax, f32)
    c = f32(0.5) * (lo[:2] + hi[:2])
    h = f32(0.5) * (hi[:2] - lo[:2])
    with np.errstate(divide="ignore", invalid="ignore"):
        nrm = (xyz[:, :2] - c) / h
    sq = nrm * nrm  # normalized**2
    inside = (sq[:, 0] + sq[:, 1]) <= f32(1.0)
    return inside & (xyz[:, 2] >= lo[2]) & (xyz[:, 2] <= hi[2])


def axis_angle_to_matrix(deg3):
    """kornia.geometry.axis_angle_to_rotation_matrix of deg2rad(deg3) (published definition: Rodrigues)."""
    v = np.deg2rad(np.asarray(deg3, np.float64))
    th = np.linalg.norm(v)
    if th < 1e-12:
        return np.eye(3)
    k = v / th
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)


def quaternion_to_matrix(q):
    """Rotation matrix of the NORMALISED quaternion (r,x,y,z) (kornia normalises first), float64."""
    q = np.asarray(q, np.float64)
    q = q / np.linalg.norm(q, axis=-1, keepdims=True)
    r, x, y, z = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
    return np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                     2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                     2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], -1).reshape(q.shape[:-1] + (3, 3))


def edit_xyz(xyz, sel, translate, bbox_center, scale, rot_deg):
    """get_xyz for one edit key (:208-236)."""
    xyz = np.array(xyz, f32)
    t = np.asarray(translate, f32)
    xyz[sel] = xyz[sel] + t
    oc = np.asarray(bbox_center, f32) + t
    xyz[sel] = (xyz[sel] - oc) * f32(scale) + oc
    R = axis_angle_to_matrix(rot_deg).astype(f32)
    xyz[sel] = ((xyz[sel] - oc).astype(np.float64) @ R.T.astype(np.float64)).astype(f32) + oc
    return xyz


def edit_scaling(scaling_raw, sel, scale):
    """_get_scaling (:242-253): log(exp(raw) * scale) on the selection (exp/log of every row are round trips there)."""
    s = np.array(scaling_raw, f32)
    s[sel] = np.log(np.exp(s[sel]) * f32(scale))
    return s


def edit_rotation_matrices(rotation_raw, sel, rot_deg):
    """_get_rotation (:256-279) as rotation matrices of the result: R_edit @ R(q) on the selection, R(q) elsewhere."""
    M = quaternion_to_matrix(rotation_raw)
    M[sel] = axis_angle_to_matrix(rot_deg) @ M[sel]
    return M


def remove(opacity, sel):
    o = np.array(opacity, f32)
    o[sel] = o[sel] * f32(0.0) - f32(100000000.0)
    return o
