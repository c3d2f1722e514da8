"""Row f3 (`distCUDA2`: mean squared distance to the three nearest neighbours).

CPU tier: the oracle restatement (oracle/egr_oracle.cpp::orc_knn3_mean_dist2, brute force) against scipy's exact
k-d tree and the degenerate cases of the reference's definition.  GPU tier: the CUDA path through the C ABI against
the oracle bit for bit, against the REFERENCE's own simple_knn.cu compiled for sm_100 (oracle/_ref/libref_knn.so,
when shipped), and at one million points against scipy."""
import numpy as np
import pytest

from oracle import oracle as orc


def _clouds():
    g = np.random.default_rng(5)
    yield "gauss", g.normal(size=(6000, 3)).astype(np.float32)
    yield "lattice", np.stack(np.meshgrid(*[np.arange(14, dtype=np.float32)] * 3), -1).reshape(-1, 3) * 0.25  # many ties
    p = g.uniform(-1, 1, (3000, 3)).astype(np.float32)
    yield "duplicates", np.concatenate([p, p[:700], p[:50]])  # distance-0 neighbours count
    yield "plane", np.concatenate([g.uniform(0, 1, (2500, 2)), np.zeros((2500, 1))], 1).astype(np.float32)
    yield "clusters", (g.normal(size=(4000, 3)) * 1e-3 + g.integers(0, 5, (4000, 1)) * 100.0).astype(np.float32)


def test_oracle_matches_exact_kdtree():
    from scipy.spatial import cKDTree

    for name, p in _clouds():
        o = orc.knn3_mean_dist2(p)
        d, _ = cKDTree(p.astype(np.float64)).query(p.astype(np.float64), k=4)
        ref = (d[:, 1:] ** 2).mean(axis=1)
        # float32 coordinates differences lose bits for far-apart clusters: compare against the coordinate scale
        tol = 4e-6 * max(1.0, float(np.abs(p).max()) ** 2) * 1e-0
        assert np.abs(o - ref).max() <= tol + 2e-6 * ref.max(), name


def test_oracle_small_counts_follow_the_reference_definition():
    big = np.float32(np.finfo(np.float32).max)
    assert np.isinf(orc.knn3_mean_dist2(np.zeros((1, 3), np.float32))[0])  # FLT_MAX + FLT_MAX overflows
    two = orc.knn3_mean_dist2(np.array([[0, 0, 0], [3, 0, 0]], np.float32))
    assert np.isinf(two).all()
    three = orc.knn3_mean_dist2(np.array([[0, 0, 0], [1, 0, 0], [0, 2, 0]], np.float32))
    assert np.array_equal(three, np.array([(np.float32(1) + np.float32(4) + big) / np.float32(3),
                                           (np.float32(1) + np.float32(5) + big) / np.float32(3),
                                           (np.float32(4) + np.float32(5) + big) / np.float32(3)], np.float32))
    four = orc.knn3_mean_dist2(np.array([[0, 0, 0], [1, 0, 0], [0, 2, 0], [0, 0, 3]], np.float32))
    assert np.array_equal(four[0], np.float32((1 + 4 + 9) / 3))


def _golden():
    import os

    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "knn_ref.npz"))


def test_oracle_reproduces_the_committed_reference_vectors():
    """tests/golden/knn_ref.npz = outputs of the reference's own simple_knn.cu run on a B200
    (tests/golden/make_golden_knn.py): the oracle matches them bit for bit."""
    g = _golden()
    for k in ("gauss", "duplicates", "plane", "tiny3"):
        assert np.array_equal(orc.knn3_mean_dist2(g["points_" + k]), g["ref_" + k]), k


@pytest.mark.gpu
def test_cuda_knn_reproduces_the_committed_reference_vectors():
    import torch

    import egr_b200

    g = _golden()
    for k in ("gauss", "duplicates", "plane", "tiny3"):
        got = egr_b200.distCUDA2(torch.from_numpy(g["points_" + k]).cuda()).cpu().numpy()
        assert np.array_equal(got, g["ref_" + k]), k


@pytest.mark.gpu
def test_cuda_knn_is_bit_identical_to_the_oracle():
    import torch

    import egr_b200

    for name, p in _clouds():
        got = egr_b200.distCUDA2(torch.from_numpy(p).cuda()).cpu().numpy()
        want = orc.knn3_mean_dist2(p)
        assert np.array_equal(got, want), (name, np.abs(got - want).max())
    for n in (1, 2, 3, 4, 5, 9):
        p = np.random.default_rng(n).normal(size=(n, 3)).astype(np.float32)
        got = egr_b200.distCUDA2(torch.from_numpy(p).cuda()).cpu().numpy()
        assert np.array_equal(got, orc.knn3_mean_dist2(p)), n
    assert egr_b200.distCUDA2(torch.zeros(0, 3, device="cuda")).shape == (0,)
    with pytest.raises(RuntimeError):
        egr_b200.distCUDA2(torch.zeros(4, 3))  # host tensor: no CPU fallback
    with pytest.raises(RuntimeError):
        egr_b200.distCUDA2(torch.zeros(4, 2, device="cuda"))


@pytest.mark.gpu
def test_cuda_knn_against_the_reference_simple_knn_on_the_gpu():
    """The reference's own simple_knn.cu (compiled in place for sm_100 by `make -C oracle refknn`) run on this GPU."""
    import torch

    import egr_b200

    if not orc.ref_knn_available(try_build=False):
        pytest.skip("oracle/_ref/libref_knn.so was not shipped")
    for name, p in _clouds():
        pts = torch.from_numpy(p).cuda()
        ref = torch.zeros(p.shape[0], device="cuda")
        orc.ref_knn3_mean_dist2(pts.data_ptr(), p.shape[0], ref.data_ptr())
        ref = ref.cpu().numpy()
        ours = egr_b200.distCUDA2(pts).cpu().numpy()
        o = orc.knn3_mean_dist2(p)
        assert np.array_equal(ours, ref), (name, np.abs(ours - ref).max())  # bit-identical
        assert np.array_equal(o, ref), name  # pins the oracle itself


@pytest.mark.gpu
def test_cuda_knn_one_million_points():
    import torch
    from scipy.spatial import cKDTree

    import egr_b200
    from egr_b200 import scenes

    p = scenes.kitchen_scene(1_000_000, seed=0)["mean"]
    got = egr_b200.distCUDA2(torch.from_numpy(p).cuda()).cpu().numpy()
    idx = np.random.default_rng(0).choice(p.shape[0], 20000, replace=False)
    d, _ = cKDTree(p.astype(np.float64)).query(p[idx].astype(np.float64), k=4)
    ref = (d[:, 1:] ** 2).mean(axis=1)
    assert np.abs(got[idx] - ref).max() <= 1e-5 * ref.max() + 1e-9
    assert np.isfinite(got).all() and (got >= 0).all()
