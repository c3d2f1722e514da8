#!/usr/bin/env python
"""Runs the BASELINE.json configs that are parity / capability cases rather than the bench line:

  C1  50k random Gaussians, 256x256, 1 bounce, fwd+bwd, checked against the CPU oracle at full size
  C3  kitchen scene, 1920x1080, 3 reflection bounces (max_bounces=3), forward only, 16 accumulated calls
  C4  3M anisotropic Gaussians, 1024x768 views sharded over the ranks of a torchrun job (one view per rank per step),
      train iteration with one gradient all-reduce; run as
      python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tests/run_configs.py c4
  C5  10M Gaussians: bounding-box selection move / delete / duplicate, refit / rebuild + 1080p 3-bounce re-render

Prints one JSON line per config.  Usage: python tests/run_configs.py [c1] [c3] [c4] [c5] [--n5 N]
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))  # (this file lives in tests/: it is test infrastructure and may use the oracle)

import torch  # noqa: E402

import egr_b200  # noqa: E402
from egr_b200 import scenes  # noqa: E402


def ev():
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    return e


def c1():
    import helpers as H
    from oracle import oracle as orc

    W = Hh = 256
    sc = scenes.random_scene(50_000, seed=0)
    cam = scenes.random_camera(W, Hh)
    tg = scenes.random_targets(W, Hh, seed=1)
    cfg = dict(num_bounces=1, jitter_primary_rays=False, loss_weight_diffuse=5.0, loss_weight_specular=3.0,
               loss_weight_depth=2.5, loss_weight_normal=2.5)
    rt = H.make_gpu_raytracer(sc, cam, cfg, hit_capacity=40_000_000)
    H.set_targets(rt, tg)
    with torch.no_grad():
        rt.raytrace()
    out = H.gpu_outputs(rt)
    with torch.enable_grad():
        for _ in range(3):
            rt.raytrace()
        rt.get_gaussians().grad_records.zero_()
        rt.get_metadata().total_num_calls.fill_(1)
        a = ev()
        rt.raytrace()
        b = ev()
    torch.cuda.synchronize()
    out2 = H.gpu_outputs(rt)
    g = H.gpu_grads(rt)
    ocfg = orc.make_config(num_bounces=1, jitter_primary_rays=0, loss_weight_diffuse=5.0, loss_weight_specular=3.0,
                           loss_weight_depth=2.5, loss_weight_normal=2.5)
    ids, rec, _, _ = rt.export_bvh()
    rec_by_id = np.zeros_like(rec.cpu().numpy())
    rec_by_id[ids.cpu().numpy()] = rec.cpu().numpy()
    t0 = time.perf_counter()
    ref = orc.render(sc, ocfg, H.oracle_camera(cam), W, Hh, total_num_calls=1)  # fully independent oracle
    t_cpu = time.perf_counter() - t0
    ov = dict(ray_origin=out2["used_ray_origin"], ray_direction=out2["used_ray_direction"],
              step_executed=out2["step_executed"], records=rec_by_id)
    refg = orc.render(sc, ocfg, H.oracle_camera(cam), W, Hh, total_num_calls=2, grads=True, targets=tg, override=ov,
                      hit_cap=96)
    keys = ("rgb", "depth", "normal", "f0", "roughness", "transmittance", "total_transmittance")
    # (a) oracle pinned to the GPU's geometry records (libm ulps in exp/log/pow cannot reorder nearly coincident
    # hits): the max-abs criterion; (b) fully independent oracle: PSNR and the number of pixels beyond 1e-4
    ov0 = dict(ray_origin=out["used_ray_origin"], ray_direction=out["used_ray_direction"],
               step_executed=out["step_executed"], records=rec_by_id)
    refp = orc.render(sc, ocfg, H.oracle_camera(cam), W, Hh, total_num_calls=1, override=ov0)
    planes = {k: float(np.abs(out[k] - refp[k]).max() / max(1.0, np.abs(refp[k]).max())) for k in keys}
    mse = float(np.mean((out["rgb"][0] - ref["rgb"][0]) ** 2))
    psnr = 10 * np.log10(1.0 / max(mse, 1e-30))
    outliers = int((np.abs(out["rgb"][0] - ref["rgb"][0]).max(-1) > 1e-4).sum())
    grads = {k: H.rel_err(g[k], refg[k]) for k in list(H.GRAD_MAP) + ["total_weight"]}
    id_mismatch = 0
    for step in range(2):
        hid, _, cnt = rt.export_hit_lists(step, 96)
        id_mismatch += int((hid.cpu().numpy().reshape(Hh, W, 96) != refg["hit_ids"][step]).any(-1).sum())
    rays = int(out2["num_steps_per_pixel"].sum())
    print(json.dumps({"config": "C1 random-50k 256x256 1 bounce fwd+bwd", "plane_max_err_all_steps_pinned": planes,
                      "independent_oracle_step0_rgb_psnr_db": psnr, "independent_oracle_pixels_over_1e-4": outliers,
                      "grad_rel_err": grads, "rays_with_hit_id_mismatch": id_mismatch,
                      "gpu_ms_fwd_bwd": a.elapsed_time(b), "gpu_mrays_s": rays / a.elapsed_time(b) / 1e3,
                      "cpu_oracle_forward_s": t_cpu, "cpu_threads": orc.num_threads(),
                      "pass": max(planes.values()) <= 1e-4 and max(grads.values()) <= 1e-3 and id_mismatch == 0}))


def c3():
    N, W, Hh = 1_000_000, 1920, 1080
    sc = scenes.kitchen_scene(N, seed=0)
    # a scene that has seen some optimisation: non-uniform opacities so that rays terminate
    g = np.random.default_rng(4)
    sc["opacity"] = g.uniform(-1.0, 4.0, (N, 1)).astype(np.float32)
    pc = scenes.SyntheticModel(sc)
    cams = scenes.kitchen_cameras(8, W, Hh)
    rtr = egr_b200.GaussianRaytracer(pc, W, Hh, max_bounces=3)
    rt = rtr.cuda_module
    rt.get_config().num_bounces.fill_(3)
    with torch.no_grad():
        view = scenes.SyntheticCamera(cams[0])
        for _ in range(3):
            egr_b200.render(view, rtr, targets_available=False)
        a = ev()
        n = 8
        for i in range(n):
            egr_b200.render(scenes.SyntheticCamera(cams[i % len(cams)]), rtr, targets_available=False)
        b = ev()
        torch.cuda.synchronize()
        single = a.elapsed_time(b) / n
        rt.get_config().accumulate_samples.copy_(torch.tensor([True]))
        rt.reset_accumulators()
        a = ev()
        for _ in range(16):
            egr_b200.render(view, rtr, targets_available=False)
        b = ev()
        torch.cuda.synchronize()
        fb = rt.get_framebuffer()
        st = rt.get_stats()
        steps = float(st.num_steps_per_pixel.float().mean())
        ok = bool(torch.isfinite(fb.output_final).all()) and int(fb.accumulated_sample_count.item()) == 16
    print(json.dumps({"config": "C3 novel-1080p 1M Gaussians 3 bounces forward", "ms_per_frame_single_call": single,
                      "fps_measure_fps_style": 1e3 / single, "ms_per_accumulated_call": a.elapsed_time(b) / 16,
                      "ray_steps_per_pixel": steps, "mrays_s": steps * W * Hh / (a.elapsed_time(b) / 16) / 1e3,
                      "output_shape": list(fb.output_rgb.shape), "pass": ok}))


def c5(n5):
    N, W, Hh = n5, 1920, 1080
    sc = scenes.kitchen_scene(N, seed=0)
    g = np.random.default_rng(4)
    sc["opacity"] = g.uniform(-1.0, 4.0, (N, 1)).astype(np.float32)
    pc = scenes.SyntheticModel(sc)
    cam = scenes.kitchen_cameras(4, W, Hh)[1]
    view = scenes.SyntheticCamera(cam)
    t0 = time.perf_counter()
    rtr = egr_b200.GaussianRaytracer(pc, W, Hh, max_bounces=3)
    torch.cuda.synchronize()
    t_build = time.perf_counter() - t0
    rt = rtr.cuda_module
    rt.get_config().num_bounces.fill_(3)
    res = {"config": f"C5 edit-{N // 1_000_000}M 1080p 3 bounces", "construct_s": t_build}

    def render_ms(force):
        a = ev()
        egr_b200.render(view, rtr, targets_available=False, force_update_bvh=force)
        b = ev()
        torch.cuda.synchronize()
        return a.elapsed_time(b)

    with torch.no_grad():
        render_ms(False)
        res["render_ms"] = render_ms(False)
        # bounding-box selection (~5 % of the points) and move, as editable_gaussian_model.py:30-50,201-279
        a = ev()
        sel = ((pc._xyz[:, 0] > 0.5) & (pc._xyz[:, 0] < 1.5) & (pc._xyz[:, 1] > 0.2) & (pc._xyz[:, 1] < 1.4) &
               (pc._xyz[:, 2] < 1.0))
        pc._xyz[sel] += torch.tensor([0.1, 0.0, 0.0], device="cuda")
        b = ev()
        torch.cuda.synchronize()
        res["selected_fraction"] = float(sel.float().mean())
        res["select_move_ms_torch_ops"] = a.elapsed_time(b)
        # the same selection + move through the device edit path (row f2): one launch each, no host sync
        from egr_b200 import edit
        pc._xyz[sel] -= torch.tensor([0.1, 0.0, 0.0], device="cuda")  # undo, then redo with the kernels
        torch.cuda.synchronize()
        a = ev()
        m8 = edit.select_box(pc._xyz, (0.5, 0.2, -1e30), (1.5, 1.4, 1.0))
        edit.apply_edit(m8, mean=pc._xyz, translate=(0.1, 0.0, 0.0), center=(1.0, 0.8, 0.5))
        b = ev()
        torch.cuda.synchronize()
        res["select_move_ms"] = a.elapsed_time(b)
        res["selected_by_kernel_vs_torch"] = [int(m8.sum()), int(sel.sum())]  # (>=,<= vs >,<; undo rounding: boundary points)
        res["move_refit_render_ms"] = render_ms(True)
        a = ev()
        rt.update_bvh()
        b = ev()
        torch.cuda.synchronize()
        res["refit_ms"] = a.elapsed_time(b)
        pc._opacity[sel] = -1e8  # delete (editable_gaussian_model.py:324-328)
        res["delete_refit_render_ms"] = render_ms(True)
        # duplicate: append the selection (means offset), then rebuild (gaussian_viewer.py:323-327,
        # editable_gaussian_model.py:284-322) -- one scan + one launch for all eight arrays (edit.duplicate)
        names = ("_xyz", "_scaling", "_rotation", "_opacity", "_diffuse", "_normal", "_roughness", "_f0")
        pc._opacity[sel] = 2.0  # (undo the delete so that the copy is visible)
        torch.cuda.synchronize()
        a = ev()  # the reference's op chain: per array `t[sel].clone()` (+ offset) and `torch.cat`
        ref_new = [torch.cat((getattr(pc, n_), getattr(pc, n_)[sel].clone() + (torch.tensor([0.0, 0.3, 0.0], device="cuda") if n_ == "_xyz" else 0)))
                   for n_ in names]
        b = ev()
        torch.cuda.synchronize()
        res["duplicate_ms_torch_ops"] = a.elapsed_time(b)
        del ref_new
        a = ev()
        new, extra = edit.duplicate([getattr(pc, n_) for n_ in names], sel, add_index=0, add_row=(0.0, 0.3, 0.0))
        b = ev()
        torch.cuda.synchronize()
        res["duplicate_ms"] = a.elapsed_time(b)
        res["duplicated_rows"] = extra
        for n_, t in zip(names, new):
            setattr(pc, n_, t)
        a = ev()
        rtr.rebuild_bvh()
        b = ev()
        torch.cuda.synchronize()
        res["duplicate_rebuild_ms"] = a.elapsed_time(b)
        res["num_gaussians_after"] = int(pc._xyz.shape[0])
        res["render_after_rebuild_ms"] = render_ms(False)
        res["pass"] = bool(torch.isfinite(rt.get_framebuffer().output_final).all())
    print(json.dumps(res))


def c4(n4=3_000_000, steps=10, warmup=3):
    import torch.distributed as dist

    from egr_b200 import distributed

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    W, Hh = 1024, 768
    sc = scenes.catacaustics_scene(n4, seed=0)
    pc = scenes.SyntheticModel(sc, device=dev, cfg=dict(loss_weight_depth=0.0, loss_weight_specular=0.01))
    cams = scenes.kitchen_cameras(64, W, Hh)
    tg = scenes.random_targets(W, Hh, seed=3)
    views = [scenes.SyntheticCamera(c, tg, device=dev) for c in cams[:16]]
    if world > 1:
        distributed.use_registered_gradient_records()
    rtr = egr_b200.GaussianRaytracer(pc, W, Hh, ppll_forward_size=1000, ppll_backward_size=120_000_000)
    rt = rtr.cuda_module
    rt.get_config().num_bounces.fill_(1)
    g = rt.get_gaussians()
    rays = torch.zeros((), dtype=torch.int64, device=dev)
    from egr_b200 import trainer as trainer_mod

    tr = trainer_mod.Trainer(pc, rtr, pruning_interval=10 ** 9)  # render + all-reduce + fused optimiser step

    def step(i):
        tr.step(views[distributed.shard_views(len(views), rank, world, i)])
        rays.add_(rt.get_stats().num_steps_per_pixel.sum())

    for i in range(warmup):
        step(i)
    rays.zero_()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    a = ev()
    for i in range(steps):
        step(warmup + i)
    b = ev()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = distributed.max_over_ranks(a.elapsed_time(b), dev)
    if world > 1:
        dist.all_reduce(rays)
    st = rt.get_stats()
    res = {"config": f"C4 catacaustics-{n4 // 1_000_000}M 1024x768 1 bounce train iteration, {world} rank(s)",
           "n_gpus": world, "ms_per_step": ms / steps, "views_per_s": world * steps / (ms * 1e-3),
           "mrays_s": int(rays.item()) / (ms * 1e-3) / 1e6, "hitlog_overflow": int(st.hitlog_status[2].item()),
           "hits_per_primary_ray": float(st.num_hits_per_pixel.sum().item()) / (W * Hh),
           "iteration": "render(view) + gradient all-reduce + fused Adam step (trainer.Trainer.step)",
           "pass": bool(torch.isfinite(pc._xyz).all()) and int(st.hitlog_status[2].item()) == 0}
    if rank == 0:
        print(json.dumps(res))
    if world > 1:
        del tr, rtr, rt, g, pc
        import gc

        gc.collect()
        torch.cuda.synchronize()
        distributed.release_registered_pools()
        torch.cuda.empty_cache()
        dist.destroy_process_group()


if __name__ == "__main__":
    args = [a.lower() for a in sys.argv[1:]]
    n5 = 10_000_000
    if "--n5" in args:
        n5 = int(args[args.index("--n5") + 1])
    which = [a for a in args if a in ("c1", "c3", "c4", "c5")] or ["c1", "c3", "c5"]
    if "c4" not in which:
        torch.cuda.set_device(0)
    for w in which:
        {"c1": c1, "c3": c3, "c4": c4, "c5": lambda: c5(n5)}[w]()
