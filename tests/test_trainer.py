"""The training iteration around the tracer (egr_b200.trainer.Trainer) against the reference's sequence of calls
(train.py:211-260) run with the package's reference-shaped facade + torch.optim.Adam on the same GPU, and the
prune -> rebind -> resize -> export -> rebuild pipeline against the reference's `prune_points` semantics."""
import numpy as np
import pytest

from egr_b200 import scenes

pytestmark = pytest.mark.gpu

ATTRS = ("_xyz", "_opacity", "_scaling", "_roughness", "_rotation", "_diffuse", "_normal", "_f0")


def _setup(n=3000, W=64, Hh=48, seed=0):
    import egr_b200

    sc = scenes.random_scene(n, seed=seed)
    cams = [scenes.random_camera(W, Hh, distance=d) for d in (4.0, 3.6, 4.4)]
    tgs = [scenes.random_targets(W, Hh, seed=10 + i) for i in range(3)]
    pc = scenes.SyntheticModel(sc)
    views = [scenes.SyntheticCamera(c, t) for c, t in zip(cams, tgs)]
    rtr = egr_b200.GaussianRaytracer(pc, W, Hh, ppll_forward_size=1000, ppll_backward_size=4_000_000)
    rtr.cuda_module.get_config().num_bounces.fill_(1)
    return sc, pc, views, rtr


def test_iteration_matches_the_reference_call_sequence():
    import torch

    import egr_b200
    from egr_b200 import trainer

    _, pc_a, views, rtr_a = _setup()
    _, pc_b, _, rtr_b = _setup()
    lr, decay = dict(trainer.DEFAULT_LR), 0.999
    tr = trainer.Trainer(pc_a, rtr_a, lr=lr, scale_decay=decay, pruning_interval=10 ** 9)
    # reference: render (with gradient import) -> scale decay -> Adam over 8 groups -> zero_grad x2 -> clamps
    params = {k: torch.nn.Parameter(getattr(pc_b, a)) for k, a in trainer.MODEL_ATTR.items()}
    for k, a in trainer.MODEL_ATTR.items():
        params[k].grad = torch.zeros_like(params[k])
        setattr(pc_b, a, params[k])
    opt = torch.optim.Adam([{"params": [params[k]], "lr": lr[k], "name": k} for k in params], lr=0.0, eps=1e-15,
                           betas=(0.9, 0.999))
    for it in range(4):
        v = views[it % len(views)]
        tr.step(v)
        with torch.enable_grad():
            egr_b200.render(v, rtr_b, denoise=False)                                   # train.py:222
        with torch.no_grad():
            pc_b._scaling.copy_(torch.log(torch.exp(pc_b._scaling) * decay))          # :224-226
            opt.step()                                                                 # :247
            opt.zero_grad(set_to_none=False)                                           # :248
            rtr_b.zero_grad()                                                          # :249
            pc_b._diffuse.data.clamp_(min=0.0)                                         # :251-254
            pc_b._roughness.data.clamp_(min=0.0, max=1.0)
            pc_b._f0.data.clamp_(min=0.0, max=1.0)
        for a in ATTRS:
            x, y = getattr(pc_a, a), getattr(pc_b, a).data
            scale = float(y.abs().max()) + 1e-30
            # float atomics reorder the gradient sums (~1e-6 relative); Adam's first steps normalise the gradient, so
            # that noise enters the update at the size of lr times the relative gradient error
            assert float((x - y).abs().max()) <= 1e-5 * scale, (it, a, float((x - y).abs().max()), scale)
    assert int(rtr_a.cuda_module.get_metadata().total_num_calls) == 4
    # total_weight is carried outside the record and equals the reference's accumulating holder
    tw_b = rtr_b.cuda_module.get_gaussians().total_weight
    assert float((tr.total_weight - tw_b).abs().max()) <= 1e-4 * float(tw_b.abs().max())
    assert float(rtr_a.cuda_module.get_gaussians().grad_records.abs().max()) == 0.0


def test_prune_pipeline_matches_prune_points_semantics():
    import torch

    import egr_b200
    import helpers as H
    from egr_b200 import trainer

    sc, pc, views, rtr = _setup(seed=1)
    tr = trainer.Trainer(pc, rtr, pruning_interval=3, pruning_start_iter=0, min_weight=0.02)
    before = None
    for it in range(3):
        if it == 2:  # state right before the pruning iteration's decision
            before = {a: getattr(pc, a).clone() for a in ATTRS}
            m_before = {k: v.clone() for k, v in tr.optimizer.exp_avg.items()}
        tr.step(views[it])
    n0, n1 = sc["mean"].shape[0], pc._xyz.shape[0]
    assert 0 < n1 < n0, (n0, n1)  # something was pruned, something survived
    rt = rtr.cuda_module
    assert rt.get_gaussians().count == n1 and rt.get_gaussians().mean.shape == (n1, 3)
    assert tr.optimizer.count == n1 and all(v.shape[0] == n1 for v in tr.optimizer.exp_avg.values())
    assert tr.total_weight.shape == (n1, 1) and float(tr.total_weight.abs().max()) == 0.0
    # the survivors are exactly the rows `tensor[keep]` keeps, in order: re-derive keep from what the step saw.
    # (the step after the prune ran Adam with zero gradients on the survivors: compare the rotation, whose lr-scaled
    # update is tiny, row-wise against the rows before)
    d = (before["_rotation"][:, None, :] - pc._rotation[None, :8, :]).abs().sum(-1)  # first 8 survivors
    idx = d.argmin(0)
    assert bool((idx[1:] > idx[:-1]).all())  # stable order
    # a tracer built from scratch on the pruned model renders the same image bit for bit
    with torch.no_grad():
        # (the pruning iteration's optimiser step moved the survivors after its rebuild: rebuild once more so that both
        # tracers sort the same means -- the transmittance product is taken in Morton order)
        rtr.rebuild_bvh()
        egr_b200.render(views[0], rtr, targets_available=False)
        a = rt.get_framebuffer().output_rgb.clone()
    sc2 = {k: getattr(pc, attr).detach().cpu().numpy() for k, attr in trainer.MODEL_ATTR.items()}
    cam = scenes.random_camera(64, 48, distance=4.0)
    rt2 = H.make_gpu_raytracer(sc2, cam, dict(num_bounces=1, alpha_threshold=0.005, transmittance_threshold=0.01,
                                              loss_weight_diffuse=5.0))
    rt2.get_metadata().total_num_calls.copy_(rt.get_metadata().total_num_calls - 1)
    with torch.no_grad():
        rt2.raytrace()
    assert torch.equal(a, rt2.get_framebuffer().output_rgb)
    # training goes on after the prune
    for it in range(2):
        tr.step(views[it])
    assert bool(torch.isfinite(pc._xyz).all())


def test_point_insertion_rebinds_and_rebuilds():
    import torch

    from egr_b200 import trainer

    sc, pc, views, rtr = _setup(seed=2)
    tr = trainer.Trainer(pc, rtr, pruning_interval=10 ** 9)
    tr.step(views[0])
    n0 = pc._xyz.shape[0]
    gen = torch.Generator(device="cuda").manual_seed(5)
    new = tr.farfield_points(500, extent=1.0, generator=gen)  # scene/gaussian_model.py:233-265
    assert new["scale"].shape == (500, 3) and bool(torch.isfinite(new["scale"]).all())
    m_old = tr.optimizer.exp_avg["rgb"].clone()
    n1 = tr.add_points(new)
    assert n1 == n0 + 500 and rtr.cuda_module.get_gaussians().count == n1
    assert torch.equal(pc._xyz[n0:], new["mean"]) and torch.equal(tr.optimizer.exp_avg["rgb"][:n0], m_old)
    assert float(tr.optimizer.exp_avg["rgb"][n0:].abs().max()) == 0.0  # cat_tensors_to_optimizer: zero moments
    tr.step(views[1])
    assert bool(torch.isfinite(pc._xyz).all()) and int(rtr.cuda_module.hitlog_overflowed()) == 0
