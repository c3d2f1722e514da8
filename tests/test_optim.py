"""Row f1: the fused step (gradient import + scale decay + Adam + clamps + zero_grad) against the reference's own
sequence of torch calls (train.py:224-226, 247-254; torch.optim.Adam with eps 1e-15 as scene/gaussian_model.py:333-338)
run on the same GPU.  Tolerance 2e-6 of each tensor's scale (fma contraction and division order differ by ulps)."""
import numpy as np
import pytest


def test_adam_argument_checks_without_gpu():
    import ctypes as C

    from egr_b200 import _capi

    lib = _capi.load()
    t = _capi.EgrGradTargets()
    o = _capi.EgrAdam()
    o.beta1, o.beta2, o.step = 0.9, 0.999, 0
    buf = (C.c_float * 24)()
    assert lib.egr_adam_step(C.addressof(buf), 1, C.byref(t), C.byref(t), C.byref(t), C.byref(o), None) == -1
    assert b"step must be" in lib.egr_last_error()
    assert lib.egr_adam_step(None, 1, C.byref(t), C.byref(t), C.byref(t), C.byref(o), None) == -1


@pytest.mark.gpu
@pytest.mark.parametrize("scale_decay", [1.0, 0.999])
def test_fused_step_matches_the_reference_sequence(scale_decay):
    import torch

    from egr_b200 import optim

    n = 50_003
    g = torch.Generator(device="cuda").manual_seed(3)
    names = optim.NAMES
    col = dict(mean=0, opacity=3, scale=4, roughness=7, rotation=8, rgb=12, normal=16, f0=20)
    init = {k: torch.randn(n, optim.CHANNELS[k], device="cuda", generator=g) * (0.3 if k != "scale" else 1.0) for k in names}
    init["roughness"] = init["roughness"].abs().clamp(max=1)
    lr = dict(mean=1.6e-4, opacity=0.05, scale=5e-3, roughness=2.5e-3, rotation=1e-3, rgb=2.5e-3, normal=2.5e-3, f0=2.5e-3)
    # reference sequence
    ref = {k: torch.nn.Parameter(v.clone()) for k, v in init.items()}
    opt = torch.optim.Adam([{"params": [ref[k]], "lr": lr[k], "name": k} for k in names], lr=0.0, eps=1e-15, betas=(0.9, 0.999))
    ours = {k: v.clone() for k, v in init.items()}
    fused = optim.FusedAdam(ours, betas=(0.9, 0.999), eps=1e-15)
    rec = torch.zeros(n, 24, device="cuda")
    for it in range(4):
        grads = {k: torch.randn(n, optim.CHANNELS[k], device="cuda", generator=g) * (10.0 ** (it - 2)) for k in names}
        grads["rgb"][::7] = 0.0  # untouched Gaussians: zero gradient
        rec.zero_()
        for k in names:
            rec[:, col[k]:col[k] + optim.CHANNELS[k]] = grads[k]
        rec[:, 15] = 1.5  # total_weight survives the zeroing
        with torch.no_grad():
            for k in names:
                ref[k].grad = grads[k].clone()                     # _import_param_gradients into zeroed .grad
            if scale_decay < 1.0:
                ref["scale"].copy_(torch.log(torch.exp(ref["scale"]) * scale_decay))   # train.py:224-226
            opt.step()                                              # train.py:247
            ref["rgb"].clamp_(min=0.0), ref["roughness"].clamp_(min=0.0, max=1.0), ref["f0"].clamp_(min=0.0, max=1.0)
        fused.step(rec, lr, scale_decay=scale_decay, clamp=True, zero_grads=True)
        assert float(rec[:, :15].abs().max()) == 0.0 and float(rec[:, 16:].abs().max()) == 0.0
        assert torch.equal(rec[:, 15], torch.full((n,), 1.5, device="cuda"))
        for k in names:
            st = opt.state[ref[k]]
            for a, b, what in ((ours[k], ref[k].data, "param"), (fused.exp_avg[k], st["exp_avg"], "exp_avg"),
                               (fused.exp_avg_sq[k], st["exp_avg_sq"], "exp_avg_sq")):
                scale = float(b.abs().max()) + 1e-30
                assert float((a - b).abs().max()) <= 2e-6 * scale, (it, k, what, float((a - b).abs().max()), scale)
    fused.step(rec, lr, zero_total_weight=True)
    assert float(rec.abs().max()) == 0.0


@pytest.mark.gpu
def test_fused_step_subset_and_errors():
    import torch

    from egr_b200 import optim

    n = 1000
    p = {"rgb": torch.rand(n, 3, device="cuda"), "mean": torch.randn(n, 3, device="cuda")}
    before = {k: v.clone() for k, v in p.items()}
    f = optim.FusedAdam(p)
    rec = torch.ones(n, 24, device="cuda")
    f.step(rec, dict(rgb=0.01, mean=0.0), zero_grads=False)
    assert float((p["rgb"] - before["rgb"]).abs().max()) > 0 and torch.equal(p["mean"], before["mean"])  # lr 0: unchanged
    assert float(rec.min()) == 1.0
    with pytest.raises(RuntimeError):
        f.step(torch.ones(n, 23, device="cuda"), dict(rgb=0.01))
    with pytest.raises(RuntimeError):
        optim.FusedAdam({"rgb": torch.rand(n, 3)})
