"""One training iteration of the reference's loop (``train.py:211-260``) around the tracer, single- or multi-GPU.

Reference sequence per iteration: ``render(view)`` (parameter export, refit, trace with the fused loss, gradient
import) -> scale decay (``train.py:224-226``) -> every ``pruning_interval`` iterations prune by accumulated weight and
``rebuild_bvh`` (``:238-245``) -> ``optimizer.step()`` + two ``zero_grad`` + three clamps (``:247-254``); once, when
bounces are switched on, far-field points are appended and the BVH rebuilt (``:256-260``,
``scene/gaussian_model.py:233-265``).

Here the same sequence, with the host-side op chains fused (rows f1 / f4 of SURVEY 8): the optimiser reads the
tracer's flat gradient record directly (``optim.FusedAdam``: import + decay + Adam + clamps + zeroing, one launch),
pruning is one mask launch + one compaction launch for all parameter / moment arrays (``prune``), followed by
rebind -> resize -> export -> ``rebuild_bvh``.

Multi-GPU (one process per GPU, ``torch.distributed`` initialised by the caller): the scene and its BVH are
replicated, every rank renders a different view of the same optimiser step, and the flat ``[N,24]`` gradient record
is summed over ranks with ONE all-reduce (the only exchange step of the path).  Every rank then applies the identical
fused step, so replicas stay bit-identical without parameter broadcasts; ``total_num_calls`` advances once per step
on every rank; pruning decisions derive from the all-reduced weights and are therefore identical; points inserted
during training are generated on rank 0 and broadcast.
"""
from __future__ import annotations

import math

import torch
import torch.distributed as dist

from . import distributed, optim, prune
from .knn import distCUDA2

# model attribute of each parameter (scene/gaussian_model.py:53-75) <-> tracer / optimiser name
MODEL_ATTR = dict(mean="_xyz", opacity="_opacity", scale="_scaling", roughness="_roughness", rotation="_rotation",
                  rgb="_diffuse", normal="_normal", f0="_f0")
# config.py:62-74 (the learning rates of training_setup, scene/gaussian_model.py:300-338)
DEFAULT_LR = dict(mean=0.00016, normal=0.0025, roughness=0.0025, f0=0.0025, rgb=0.005, opacity=0.025, scale=0.005,
                  rotation=0.001)


def _world(group=None):
    return dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1


def _rank(group=None):
    return dist.get_rank(group) if dist.is_available() and dist.is_initialized() else 0


class Trainer:
    """``pc``: the model (duck-typed ``GaussianModel``: raw parameters ``_xyz ... _f0`` as contiguous float32 CUDA
    tensors).  ``raytracer``: the ``GaussianRaytracer`` built on it."""

    def __init__(self, pc, raytracer, lr: dict | None = None, scale_decay: float = 0.9999, pruning_interval: int = 125,
                 pruning_start_iter: int = 1250, min_weight: float = 0.1, betas=(0.9, 0.999), eps: float = 1e-15,
                 group=None):
        self.pc, self.raytracer, self.group = pc, raytracer, group
        self.lr = dict(DEFAULT_LR if lr is None else lr)
        self.scale_decay, self.pruning_interval = float(scale_decay), int(pruning_interval)
        self.pruning_start_iter, self.min_weight = int(pruning_start_iter), float(min_weight)
        self.iteration = 0
        self.optimizer = optim.FusedAdam(self._params(), betas=betas, eps=eps)
        n = pc._xyz.shape[0]
        # weight accumulated since the last pruning decision (the reference keeps it in the tracer's total_weight
        # holder; here that column of the record is all-reduced with the gradients every step, so the running sum
        # lives outside the record)
        self.total_weight = torch.zeros(n, 1, dtype=torch.float32, device=pc._xyz.device)
        # the fused optimiser reads the record itself: the facade's eight `.grad.add_` imports are not needed
        raytracer.import_gradients = False

    def _params(self):
        return {k: getattr(self.pc, a) for k, a in MODEL_ATTR.items()}

    # ---- one iteration (train.py:211-254) ----------------------------------------------------------------------
    def step(self, view, events: dict | None = None) -> None:
        """``events``: optional {"render" | "allreduce" | "optimizer": (start, end)} CUDA event pairs recorded around
        the three phases (bench.py's per-phase split)."""
        from .renderer import render

        def mark(name, i):
            if events is not None and name in events:
                events[name][i].record()

        self.iteration += 1
        rt = self.raytracer.cuda_module
        mark("render", 0)
        with torch.enable_grad():
            render(view, self.raytracer, denoise=False)
        mark("render", 1)
        rec = rt.get_gaussians().grad_records
        mark("allreduce", 0)
        distributed.allreduce_gradients(rec, group=self.group)
        mark("allreduce", 1)
        mark("optimizer", 0)
        with torch.no_grad():
            self.total_weight.add_(rec[:, 15:16])  # (the fused step below zeroes the whole record, this column too)
            if self.iteration % self.pruning_interval == 0:
                if self.iteration > self.pruning_start_iter and self.min_weight > 0:
                    keep = prune.keep_mask_from_weight(self.total_weight, self.pruning_interval, self.min_weight)
                    self.prune(keep)
                    rec = rt.get_gaussians().grad_records  # resize() made a new (zeroed) record, as prune_points does
                else:
                    self.raytracer.rebuild_bvh()
                self.total_weight.zero_()
        self.optimizer.step(rec, self.lr, scale_decay=self.scale_decay, clamp=True, zero_grads=True,
                            zero_total_weight=True)
        mark("optimizer", 1)

    # ---- prune -> rebind -> resize -> export -> rebuild (scene/gaussian_model.py:479-532, train.py:238-245) -------
    @torch.no_grad()
    def prune(self, keep: torch.Tensor) -> int:
        """Keeps the rows with ``keep != 0`` in every parameter, both Adam moments and the weight accumulator (one
        compaction launch, one 4-byte host read), rebinds the optimiser, resizes the tracer's holders and rebuilds the
        BVH.  Returns the new count.  ``keep`` must be identical on all ranks (it is when derived from all-reduced
        data)."""
        names = list(MODEL_ATTR)
        opt = self.optimizer
        arrays = ([getattr(self.pc, MODEL_ATTR[k]) for k in names] + [opt.exp_avg[k] for k in names] +
                  [opt.exp_avg_sq[k] for k in names] + [self.total_weight])
        out, count = prune.compact(arrays, keep)
        n = len(names)
        for i, k in enumerate(names):
            t = out[i].contiguous()
            t.grad = torch.zeros_like(t)  # prune_points re-creates zeroed .grad tensors
            setattr(self.pc, MODEL_ATTR[k], t)
        opt.rebind(self._params(), {k: out[n + i].contiguous() for i, k in enumerate(names)},
                   {k: out[2 * n + i].contiguous() for i, k in enumerate(names)})
        self.total_weight = out[3 * n].contiguous()
        self.raytracer.rebuild_bvh()  # resize + export + rebuild
        return count

    # ---- point insertion (train.py:256-260, scene/gaussian_model.py:233-265) ------------------------------------
    @torch.no_grad()
    def add_points(self, new: dict) -> int:
        """Appends Gaussians (``new``: name -> [M, ch] raw parameter rows, names of ``MODEL_ATTR``) with zeroed Adam
        moments (``cat_tensors_to_optimizer``), then rebuilds.  In a multi-GPU job the rows of rank 0 are used on
        every rank (broadcast), so that replicas stay identical even if the rows were generated randomly."""
        dev = self.pc._xyz.device
        names = list(MODEL_ATTR)
        rows = {k: new[k].to(dev, torch.float32).contiguous() for k in names}
        if _world(self.group) > 1:
            m = torch.tensor([rows["mean"].shape[0]], dtype=torch.int64, device=dev)
            dist.broadcast(m, src=0, group=self.group)
            for k in names:
                if _rank(self.group) != 0:
                    rows[k] = torch.empty(int(m.item()), getattr(self.pc, MODEL_ATTR[k]).shape[1], dtype=torch.float32,
                                          device=dev)
                dist.broadcast(rows[k], src=0, group=self.group)
        opt = self.optimizer
        m_new, v_new = {}, {}
        for k in names:
            t = torch.cat([getattr(self.pc, MODEL_ATTR[k]), rows[k]]).contiguous()
            t.grad = torch.zeros_like(t)
            setattr(self.pc, MODEL_ATTR[k], t)
            m_new[k] = torch.cat([opt.exp_avg[k], torch.zeros_like(rows[k])]).contiguous()
            v_new[k] = torch.cat([opt.exp_avg_sq[k], torch.zeros_like(rows[k])]).contiguous()
        opt.rebind(self._params(), m_new, v_new)
        self.total_weight = torch.cat([self.total_weight, torch.zeros(rows["mean"].shape[0], 1, device=dev)])
        self.raytracer.rebuild_bvh()
        return int(self.pc._xyz.shape[0])

    @torch.no_grad()
    def farfield_points(self, num: int, extent: float, init_scale: float = 0.1, init_opacity: float = 0.1,
                        init_diffuse: float = 0.2, radius: float = 4.0, generator=None) -> dict:
        """The rows ``GaussianModel.add_farfield_points`` creates (scene/gaussian_model.py:233-265): clamped normal
        positions, scale from the 3-NN distance (``distCUDA2``), identity rotation, constant appearance."""
        dev = self.pc._xyz.device
        xyz = torch.randn(num, 3, device=dev, generator=generator).clamp(-3, 3) * float(extent) * float(radius)
        d2 = torch.clamp_min(distCUDA2(xyz), 1e-7)
        scale = torch.log(torch.sqrt(d2) * init_scale)[..., None].repeat(1, 3)
        rot = torch.zeros(num, 4, device=dev)
        rot[:, 0] = 1
        opa = torch.full((num, 1), math.log(init_opacity / (1 - init_opacity)), device=dev)  # inverse_sigmoid
        return dict(mean=xyz, scale=scale, rotation=rot, opacity=opa, rgb=torch.full((num, 3), init_diffuse, device=dev),
                    normal=torch.zeros(num, 3, device=dev), f0=torch.full((num, 3), 0.04, device=dev),
                    roughness=torch.zeros(num, 1, device=dev))
