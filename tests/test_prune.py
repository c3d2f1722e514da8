"""Row f4: pruning mask and stable multi-array compaction against their torch definitions
(train.py:239, scene/gaussian_model.py:479-532: `tensor[mask]`).  Bit-exact."""
import numpy as np
import pytest


def test_prune_symbols_and_argument_checks_without_gpu():
    import ctypes as C

    from egr_b200 import _capi

    lib = _capi.load()
    assert lib.egr_compact_scratch_bytes(1000) >= 4000
    assert lib.egr_compact_scratch_bytes(2_000_000) < 2_000_000 * 4 + (8 << 20)
    cnt = C.c_uint32(0)
    # count_out NULL / too many arrays are rejected before any CUDA call
    assert lib.egr_compact_rows(None, 10, 0, None, None, None, None, 0, None, None) == -1  # EGR_EINVAL
    assert lib.egr_compact_rows(C.addressof(cnt), 10, 41, None, None, None, None, 0, C.addressof(cnt), None) == -1  # EGR_EINVAL
    assert b"at most" in lib.egr_last_error()


@pytest.mark.gpu
def test_keep_mask_matches_the_reference_expression():
    import torch

    from egr_b200 import prune

    g = torch.Generator(device="cuda").manual_seed(0)
    tw = torch.rand(100_003, 1, device="cuda", generator=g) * 3.0
    tw[5] = float("nan")
    tw[7] = 0.0
    for interval, mw in ((125.0, 0.01), (1.0, 1.5), (3.0, 0.0)):
        want = ~((tw / interval < mw).squeeze(1))
        got = prune.keep_mask_from_weight(tw, interval, mw)
        assert torch.equal(got.bool(), want)
    # the strided total_weight view of the flat gradient record
    rec = torch.rand(5000, 24, device="cuda", generator=g)
    assert torch.equal(prune.keep_mask_from_weight(rec[:, 15:16], 125.0, 0.004).bool(), ~(rec[:, 15] / 125.0 < 0.004))


@pytest.mark.gpu
@pytest.mark.parametrize("n,p_keep", [(1, 1.0), (1, 0.0), (257, 0.5), (100_003, 0.9), (100_003, 0.0), (1_000_000, 0.37)])
def test_compaction_equals_boolean_indexing(n, p_keep):
    import torch

    from egr_b200 import prune

    g = torch.Generator(device="cuda").manual_seed(n)
    keep = torch.rand(n, device="cuda", generator=g) < p_keep
    shapes = [(n, 3), (n, 1), (n, 4), (n, 3), (n, 3), (n, 1), (n, 3), (n, 3)]
    params = [torch.randn(s, device="cuda", generator=g) for s in shapes]
    tensors = params + [p * 2 for p in params] + [p.abs() for p in params]  # parameter, exp_avg, exp_avg_sq
    tensors += [torch.randint(0, 1 << 30, (n,), device="cuda", generator=g, dtype=torch.int32),  # _round_counter
                torch.randn(n, 16, device="cuda", generator=g), torch.randn(n, 2, device="cuda", generator=g).double()]
    out, count = prune.compact(tensors, keep)
    assert count == int(keep.sum())
    for t, o in zip(tensors, out):
        assert o.dtype == t.dtype and torch.equal(o, t[keep])
    out8, c8 = prune.compact(tensors[:2], keep.to(torch.uint8))
    assert c8 == count and torch.equal(out8[1], tensors[1][keep])


@pytest.mark.gpu
def test_compaction_rejects_what_it_cannot_do():
    import torch

    from egr_b200 import prune

    keep = torch.ones(8, dtype=torch.bool, device="cuda")
    with pytest.raises(RuntimeError):
        prune.compact([torch.zeros(8, 3, dtype=torch.uint8, device="cuda")], keep)  # 3-byte rows
    with pytest.raises(RuntimeError):
        prune.compact([torch.zeros(9, 3, device="cuda")], keep)
    with pytest.raises(RuntimeError):
        prune.compact([torch.zeros(8, 3)], keep)  # host tensor


@pytest.mark.gpu
def test_near_camera_cull_matches_the_per_camera_loop():
    """scene.py:88-105 run as written (one norm pass per camera, OR-ed) on the same GPU."""
    import torch

    from egr_b200 import prune

    g = torch.Generator(device="cuda").manual_seed(9)
    n, nc = 300_007, 301
    pts = torch.rand(n, 3, device="cuda", generator=g) * 4 - 2
    centers = torch.rand(nc, 3, device="cuda", generator=g) * 4 - 2
    znear = torch.rand(nc, device="cuda", generator=g) * 0.12
    want = torch.zeros(n, dtype=torch.bool, device="cuda")
    for k in range(nc):
        want |= (pts - centers[k]).norm(dim=1) < znear[k]
    assert 100 < int(want.sum()) < n // 2
    keep = torch.ones(n, dtype=torch.uint8, device="cuda")
    keep[::5] = 0  # already dropped entries stay dropped
    prune.drop_near_cameras(keep, pts, centers, znear)
    expect = torch.ones(n, dtype=torch.bool, device="cuda")
    expect[::5] = False
    expect &= ~want
    assert torch.equal(keep.bool(), expect)
