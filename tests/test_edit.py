"""Row f2: selection masks and edit transforms, CUDA (through the C ABI) against oracle/edit_oracle.py.
Masks, translate+scale of the means, the scaling edit and the delete are compared bit for bit (same float32
operations in the same order); the rotated means within 2e-6 of the object's size (the reference's GEMM has no
specified summation order) and the composed rotations as matrices within 2e-6 (unit quaternions may differ in sign)."""
import numpy as np
import pytest

from oracle import edit_oracle as eo


def _cloud(n=200_003, seed=0):
    g = np.random.default_rng(seed)
    xyz = g.uniform(-2, 2, (n, 3)).astype(np.float32)
    xyz[:64] = np.array([0.5, -0.25, 1.0], np.float32)  # exactly on the faces / rim of the boxes below
    scale = g.normal(-4.0, 0.7, (n, 3)).astype(np.float32)
    rot = g.normal(size=(n, 4)).astype(np.float32) * g.uniform(0.2, 3.0, (n, 1)).astype(np.float32)  # un-normalised
    opac = g.normal(size=(n, 1)).astype(np.float32)
    return xyz, scale, rot, opac


BOXES = [((-0.5, -1.25, -1.0), (0.5, -0.25, 1.0)), ((0.1, 0.2, 0.3), (1.7, 0.9, 1.1)), ((-3, -3, -3), (3, 3, 3)),
         ((0.5, 0.5, 0.5), (0.5, 0.5, 0.5))]


def test_oracle_selection_sanity():
    xyz = np.array([[0, 0, 0], [1, 0, 0], [0.5, 0.5, 2.0], [0.99, 0.0, 0.5]], np.float32)
    assert eo.select_box(xyz, (0, 0, 0), (1, 1, 1)).tolist() == [True, True, False, True]
    assert eo.select_cylinder(np.array([[0.5, 0.5, 0.5], [0.5, 1.0, 0.5], [0.9, 0.9, 0.5]], np.float32),
                              (0, 0, 0), (1, 1, 1)).tolist() == [True, True, False]
    R = eo.axis_angle_to_matrix((0, 0, 90))
    assert np.allclose(R @ np.array([1.0, 0, 0]), [0, 1, 0], atol=1e-12)
    q = np.array([[2.0, 0, 0, 0], [0.7071068, 0, 0, 0.7071068]])
    assert np.allclose(eo.quaternion_to_matrix(q)[0], np.eye(3)) and np.allclose(eo.quaternion_to_matrix(q)[1], R, atol=1e-6)


@pytest.mark.gpu
def test_selection_masks_bit_exact():
    import torch

    from egr_b200 import edit

    xyz, *_ = _cloud()
    t = torch.from_numpy(xyz).cuda()
    for lo, hi in BOXES:
        assert np.array_equal(edit.select_box(t, lo, hi).cpu().numpy().astype(bool), eo.select_box(xyz, lo, hi)), (lo, hi)
        assert np.array_equal(edit.select_cylinder(t, lo, hi).cpu().numpy().astype(bool),
                              eo.select_cylinder(xyz, lo, hi)), (lo, hi)
    assert int(edit.select_box(t, *BOXES[0]).sum()) > 1000
    assert edit.select_box(torch.zeros(0, 3, device="cuda"), (0, 0, 0), (1, 1, 1)).shape == (0,)
    with pytest.raises(RuntimeError):
        edit.select_box(torch.zeros(5, 3), (0, 0, 0), (1, 1, 1))


@pytest.mark.gpu
@pytest.mark.parametrize("rot", [(0.0, 0.0, 0.0), (10.0, -35.0, 80.0)])
def test_edit_transform_matches_the_reference_formulas(rot):
    import torch

    from egr_b200 import edit

    xyz, scale, rotq, opac = _cloud(seed=1)
    lo, hi = BOXES[1]
    sel = eo.select_box(xyz, lo, hi)
    center = [(lo[i] + hi[i]) / 2 for i in range(3)]
    tr, sc = (0.25, -0.5, 0.125), 1.3
    d = {k: torch.from_numpy(v.copy()).cuda() for k, v in dict(mean=xyz, scale=scale, rotation=rotq, opacity=opac).items()}
    mask = edit.select_box(d["mean"], lo, hi)
    edit.apply_edit(mask, mean=d["mean"], scale=d["scale"], rotation=d["rotation"], translate=tr, center=center,
                    scale_factor=sc, rotate_deg=rot)
    want_xyz = eo.edit_xyz(xyz, sel, tr, center, sc, rot)
    got_xyz = d["mean"].cpu().numpy()
    assert np.array_equal(got_xyz[~sel], xyz[~sel])
    if rot == (0.0, 0.0, 0.0):
        assert np.array_equal(got_xyz, want_xyz)  # identity rotation: every operation is exact or identically rounded
    else:
        assert np.abs(got_xyz - want_xyz).max() <= 2e-6 * 4.0
    # the scaling edit on the selection; untouched rows keep their raw values (the reference's exp/log round trip of
    # unselected rows is a property-level artefact: the tracer takes exp() of either)
    want_s = eo.edit_scaling(scale, sel, sc)
    got_s = d["scale"].cpu().numpy()
    assert np.array_equal(got_s[~sel], scale[~sel])
    assert np.abs(got_s[sel] - want_s[sel]).max() <= 2e-6 * np.abs(want_s[sel]).max()  # CUDA vs numpy exp/log ulps
    # rotations, as matrices
    got_M = eo.quaternion_to_matrix(d["rotation"].cpu().numpy())
    want_M = eo.edit_rotation_matrices(rotq, sel, rot)
    assert np.abs(got_M - want_M).max() <= 2e-6
    qn = np.linalg.norm(d["rotation"].cpu().numpy()[sel].astype(np.float64), axis=1)
    assert np.abs(qn - 1).max() <= 1e-6
    assert np.array_equal(d["rotation"].cpu().numpy()[~sel], rotq[~sel])
    # delete
    edit.apply_edit(mask, opacity=d["opacity"], kill=True)
    assert np.array_equal(d["opacity"].cpu().numpy(), eo.remove(opac, sel))


@pytest.mark.gpu
def test_edit_then_refit_renders_like_an_edited_upload():
    """The device edit feeding the refit directly == exporting host-edited parameters (the reference's route)."""
    import torch

    import helpers as H
    from egr_b200 import edit, scenes

    W, Hh = 48, 32
    sc = scenes.random_scene(3000, seed=2)
    cam = scenes.random_camera(W, Hh)
    lo, hi = (-0.6, -0.6, -0.6), (0.4, 0.5, 0.6)
    center = [(lo[i] + hi[i]) / 2 for i in range(3)]
    sel = eo.select_box(sc["mean"], lo, hi)
    assert 50 < sel.sum() < 2900
    # primary step only: a bounce amplifies last-bit differences of the records (different tree -> different product order)
    rt = H.make_gpu_raytracer(sc, cam, dict(num_bounces=0, jitter_primary_rays=False))
    g = rt.get_gaussians()
    mask = edit.select_box(g.mean, lo, hi)
    edit.apply_edit(mask, mean=g.mean, scale=g.scale, rotation=g.rotation, translate=(0.1, 0.0, 0.0), center=center,
                    scale_factor=1.2, rotate_deg=(0.0, 0.0, 0.0))
    rt.update_bvh()
    with torch.no_grad():
        rt.raytrace()
    a = H.gpu_outputs(rt)["rgb"]
    sc2 = dict(sc)
    sc2["mean"] = eo.edit_xyz(sc["mean"], sel, (0.1, 0.0, 0.0), center, 1.2, (0, 0, 0))
    sc2["scale"] = g.scale.cpu().numpy()  # (exp/log ulps of the scaling edit: take the device values)
    sc2["rotation"] = g.rotation.cpu().numpy()  # (selected rows come back normalised)
    rt2 = H.make_gpu_raytracer(sc2, cam, dict(num_bounces=0, jitter_primary_rays=False))
    with torch.no_grad():
        rt2.raytrace()
    b = H.gpu_outputs(rt2)["rgb"]
    assert float(np.abs(a - b).max()) <= 1e-4 and float(np.abs(a).max()) > 0
