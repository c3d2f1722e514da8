#!/usr/bin/env python
"""Benchmark of the hot path: Mrays/s forward+backward (1 bounce, 1 M Gaussians), BASELINE.json config C2.

A "step" is one training-iteration pass of the tracer over one 800x800 view of the synthetic kitchen-1M scene:
BVH refit + forward (primary + 1 bounce) + fused-loss backward.  A "ray" is one executed ray-step of one pixel
(reference: total_effective_steps, shaders.cu:101-105).

  python bench.py [--gpus N] [--steps K] [--warmup W]            our CUDA path (one process per GPU under torchrun)
  python bench.py --impl reference ...                            the reference's shader code on host cores

JSON keys are those of the driver contract; `roofline` is computed for the slowest kernel of the step from CUDA
events recorded around every launch inside the timed region; `cpu_baseline` is a bounded CPU sample.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mrays/s fwd+bwd (1-bounce, 1M Gaussians)"
WORKLOAD = "C2 kitchen-1M: 1,000,000 Gaussians, 800x800 views, train iteration (refit+fwd+bwd), 1 bounce"


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.path = tempfile.mktemp(suffix=".csv")
        self.proc = None
        self.idx = gpu_index

    def start(self):
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            p = [x.strip() for x in line.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1])), mx.append(float(p[2]))
            except ValueError:
                continue
            for n, v in zip(names, p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.path)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        # samples taken under load = upper half of the sorted clocks (the sampler also sees idle edges)
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def cpu_sample(kind, scene, cam, n_pixels, seconds_hint=None):
    """Time the CPU tracer (fwd+bwd, 1 bounce) on a bounded block of pixels of one C2 view, all host threads."""
    from oracle import oracle as orc

    W, Hh = cam["width"], cam["height"]
    cfg = orc.make_config(num_bounces=1, jitter_primary_rays=1, loss_weight_diffuse=5.0, loss_weight_specular=3.0,
                          loss_weight_depth=2.5, loss_weight_normal=2.5)
    ocam = orc.make_camera(cam["origin"], cam["c2w"], cam["fovy"])
    from egr_b200 import scenes

    tg = scenes.random_targets(W, Hh, seed=1)
    pb = (Hh // 2) * W + W // 4
    rng = (pb, pb + n_pixels)
    use_ref = kind == "reference" and orc.ref_available(try_build=False)
    fn = orc.ref_render if use_ref else orc.render
    t0 = time.perf_counter()
    fn(scene, cfg, ocam, W, Hh, total_num_calls=1, grads=True, targets=tg, pixel_range=rng)
    dt = time.perf_counter() - t0
    # executed ray-steps of the sample (the reference build does not export them; both agree on step counts)
    steps = orc.render(scene, cfg, ocam, W, Hh, total_num_calls=1, pixel_range=rng)["num_steps_per_pixel"]
    rays = int(steps.sum())
    cores = orc.ref_lib().ref_num_threads() if use_ref else orc.num_threads()
    return dict(value=rays / dt / 1e6, unit="Mrays/s", cores=int(cores), kind="reference" if use_ref else "port",
                sample=f"{n_pixels} pixels of one 800x800 C2 view, fwd+bwd, 1 bounce, brute-force candidates over "
                       f"1M Gaussians ({rays} ray-steps in {dt:.1f} s)",
                seconds=dt, rays=rays)


def run_reference(args):
    """The reference's own shader code (oracle/_ref, compiled from /root/reference for the host; else the CPU port)
    on the box's host cores.  The reference has no CPU tracer of its own: OptiX traversal is replaced by a
    brute-force candidate loop."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # torchrun exports OMP_NUM_THREADS=1 to its workers; this arm is the one CPU job of the launch and is meant to
    # use every host core (the OpenMP runtime reads the variable when the oracle library is loaded, below)
    if os.environ.get("OMP_NUM_THREADS") == "1" and int(os.environ.get("WORLD_SIZE", "1")) > 1:
        os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    from egr_b200 import scenes

    scene = scenes.kitchen_scene(args.num_gaussians, seed=0)
    cams = scenes.kitchen_cameras(100, args.res, args.res)
    n_pix = args.cpu_pixels
    vals = []
    for i in range(args.warmup + args.steps):
        r = cpu_sample("reference", scene, cams[i % len(cams)], n_pix)
        if i >= args.warmup:
            vals.append(r)
    v = sum(x["rays"] for x in vals) / sum(x["seconds"] for x in vals) / 1e6
    ms = 1e3 * sum(x["seconds"] for x in vals) / len(vals)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample_per_step": f"{n_pix} pixels of one view (CPU-bounded sample)"},
        "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": vals[0]["cores"], "kind": vals[0]["kind"],
                         "sample": vals[0]["sample"]},
        "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def run_ours(args):
    import torch
    import torch.distributed as dist

    import egr_b200
    from egr_b200 import _capi, distributed, scenes

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    peak_gbs, peak_kind = _peaks()

    N, R = args.num_gaussians, args.res
    scene = scenes.kitchen_scene(N, seed=0)
    cams = scenes.kitchen_cameras(100, R, R)
    n_tg = 4
    targets = [scenes.random_targets(R, R, seed=1 + i) for i in range(n_tg)]
    pc = scenes.SyntheticModel(scene, device=dev)
    views_dev = [scenes.SyntheticCamera(cams[i], targets[i % n_tg], device=dev) for i in range(len(cams))]
    rtr = egr_b200.GaussianRaytracer(pc, R, R, ppll_forward_size=1000, ppll_backward_size=args.hitlog_hits)
    rt = rtr.cuda_module
    rt.get_config().num_bounces.fill_(1)
    S = rt._S
    fb, cam_h = rt.get_framebuffer(), rt.get_camera()
    g = rt.get_gaussians()

    tkeys = (("diffuse", "diffuse_image"), ("specular", "specular_image"), ("depth", "depth_image"),
             ("normal", "normal_image"), ("f0", "f0_image"), ("roughness", "roughness_image"))
    for v in views_dev:  # poses resident on the device too; targets resident in the tracer's HWC layout
        Rm = torch.from_numpy(v.R).to(dev)
        v.c2w_dev = -Rm
        v.c2w_dev[:, 0] = -v.c2w_dev[:, 0]
        v.hwc = {k: getattr(v, attr).moveaxis(0, -1).contiguous() for k, attr in tkeys}
    torch.cuda.synchronize()

    def set_view_device(v):
        """Inputs already resident in HBM: pose + targets set from device tensors, no host traffic."""
        with torch.no_grad():
            cam_h.znear.fill_(0.01), cam_h.zfar.fill_(999.9), cam_h.vertical_fov_radians.fill_(v.FoVy)
            cam_h.set_pose(v.camera_center, v.c2w_dev)
            for k, _ in tkeys:
                getattr(fb, "target_" + k).copy_(v.hwc[k])

    # one event pair per launch where launches are serial; the bounce steps and the backward go in ONE call because
    # the library runs the backward of step s-1 underneath the forward of step s (side stream) within a call
    rest = _capi.EGR_PHASE_FINISH
    for s in range(1, S):
        rest |= _capi.EGR_PHASE_STEP(s)
    phases = [("begin", _capi.EGR_PHASE_BEGIN), ("trace_step0", _capi.EGR_PHASE_STEP(0)), ("bounce+backward", rest)]
    rays_dev = torch.zeros((), dtype=torch.int64, device=dev)

    def new_pair():
        return torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def hot_step(view_idx, events=None):
        """refit + forward + backward of one resident view (+ the gradient all-reduce when world > 1).
        `events`: pre-created {phase: (start, end)} pairs recorded around every launch of this step."""
        with torch.enable_grad():
            if events is not None:
                events["refit"][0].record()
            rt.update_bvh()
            if events is not None:
                events["refit"][1].record()
            for name, mask in phases:
                if events is not None:
                    events[name][0].record()
                rt.raytrace_phases(mask)
                if events is not None:
                    events[name][1].record()
        if world > 1:
            if events is not None:
                events["allreduce"][0].record()
            distributed.allreduce_gradients(g.grad_records)
            if events is not None:
                events["allreduce"][1].record()
        rays_dev.add_(rt.get_stats().num_steps_per_pixel.sum())

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-resident throughput (`value`) ----------------
    K, Wm = args.steps, args.warmup
    my_views = [distributed.shard_views(len(cams), rank, world, i) for i in range(Wm + K)]
    for i in range(Wm):
        set_view_device(views_dev[my_views[i]])
        hot_step(my_views[i])
    rtr.zero_grad()
    rays_dev.zero_()
    # timed region: exactly K steps between one pair of events (barrier + synchronize on both sides); selecting the
    # step's view = device-to-device copies of pose and targets that are already resident in HBM
    clock = ClockSampler(local) if rank == 0 and not os.environ.get("EGR_BENCH_NOCLOCK") else None  # one sampler per job
    if clock:
        clock.start()
    sync()
    names = ["refit"] + [n for n, _ in phases] + (["allreduce"] if world > 1 else [])
    ev_steps = [{n: new_pair() for n in names} for _ in range(K)]  # created before the timed region
    t_a, t_b = new_pair()
    t_a.record()
    for i in range(K):
        set_view_device(views_dev[my_views[Wm + i]])
        hot_step(my_views[Wm + i], ev_steps[i])
    t_b.record()
    sync()
    clocks = clock.stop() if clock else None
    total_ms = distributed.max_over_ranks(t_a.elapsed_time(t_b), dev)
    rays = rays_dev.clone()
    if world > 1:
        dist.all_reduce(rays)
    rays = int(rays.item())
    value = rays / (total_ms * 1e-3) / 1e6
    per_phase = {}
    for evs in ev_steps:
        for name, (a, b) in evs.items():
            per_phase.setdefault(name, []).append(a.elapsed_time(b))
    phase_ms = {k: sum(v) / len(v) for k, v in per_phase.items()}

    # ---------------- roofline of the slowest kernel ----------------
    roofline = None
    if rank == 0:
        with torch.no_grad():
            v = views_dev[my_views[Wm + K - 1]]
            set_view_device(v)
            st = rt.get_stats()
            rt.get_config().num_bounces.fill_(0)
            rt.raytrace()
            h0, c0 = int(st.num_hits_per_pixel.sum().item()), int(st.num_composited_per_pixel.sum().item())
            rt.get_config().num_bounces.fill_(1)
            rt.get_metadata().total_num_calls.sub_(1)
            rt.raytrace()
            h1 = int(st.num_hits_per_pixel.sum().item()) - h0
            c1 = int(st.num_composited_per_pixel.sum().item()) - c0
        P = R * R
        # the launches that overlap inside the timed region (bounce forward, backward) are also timed one by one in a
        # short serial pass (one call per launch: no overlap), so that every kernel has a stand-alone duration
        serial = {}
        ser_phases = [("trace_step1", _capi.EGR_PHASE_STEP(1))] if S > 1 else []
        ser_phases += [(f"trace_step{s}", _capi.EGR_PHASE_STEP(s)) for s in range(2, S)] + [("backward", _capi.EGR_PHASE_FINISH)]
        for i in range(3):
            set_view_device(views_dev[my_views[Wm + i % K]])
            with torch.enable_grad():
                rt.update_bvh()
                rt.raytrace_phases(_capi.EGR_PHASE_BEGIN | _capi.EGR_PHASE_STEP(0))
                for name, mask in ser_phases:
                    a, b = new_pair()
                    a.record()
                    rt.raytrace_phases(mask)
                    b.record()
                    serial.setdefault(name, []).append((a, b))
        torch.cuda.synchronize()
        rtr.zero_grad()
        serial_ms = {k: sum(a.elapsed_time(b) for a, b in v) / len(v) for k, v in serial.items()}
        alg = {"trace_step0": 44 * h0 + 40 * c0, "trace_step1": 44 * h1 + 40 * c1,
               "backward": 260 * c0 + 176 * c1 + 56 * P}
        kernel_ms = dict(serial_ms)
        kernel_ms["trace_step0"] = phase_ms["trace_step0"]  # serial inside the timed region: measured there
        dom = max(alg, key=lambda k: kernel_ms.get(k, 0.0))
        achieved = alg[dom] / (kernel_ms[dom] * 1e-3) / 1e9
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            try:
                traffic = json.load(open(tpath)).get(dom)
            except Exception:
                traffic = None
        roofline = {"bound": "hbm", "kernel": {"trace_step0": "k_trace_step(step=0)", "trace_step1": "k_trace_step(step=1)",
                                               "backward": "k_backward"}[dom],
                    "achieved": achieved, "peak": peak_gbs, "unit": "GB/s", "frac": achieved / peak_gbs,
                    "traffic": traffic, "peak_source": peak_kind + " (MEASURED_PEAKS.json hbm_gbs)",
                    "algorithmic_bytes_per_launch": alg[dom], "avg_launch_ms": kernel_ms[dom],
                    "timed_where": "timed region" if dom == "trace_step0" else "serial pass after the timed region",
                    "phase_ms": {k: round(v, 4) for k, v in phase_ms.items()},
                    "serial_kernel_ms": {k: round(v, 4) for k, v in serial_ms.items()},
                    "hits_per_ray": {"accepted_step0": h0 / P, "composited_step0": c0 / P,
                                     "accepted_step1": h1 / P, "composited_step1": c1 / P}}

    # ---------------- end to end through the renderer facade with HOST buffers (`e2e`) ----------------
    views_host = [scenes.SyntheticCamera(cams[i], targets[i % n_tg], device="cpu", pin=True) for i in range(n_tg)]
    for v in views_host:
        v.camera_center = v.camera_center.pin_memory()
    h2d = sum(getattr(views_host[0], a).numel() * 4 for a in ("diffuse_image", "specular_image", "depth_image",
                                                              "normal_image", "f0_image", "roughness_image")) + 12 + 36
    out_host = torch.zeros(2, dtype=torch.float64).pin_memory()
    d2h = out_host.numel() * 8

    def e2e_step(i):
        v = views_host[i % n_tg]
        with torch.enable_grad():
            egr_b200.render(v, rtr, denoise=False)  # host->device copies of pose + 6 target images happen inside
        if world > 1:
            distributed.allreduce_gradients(g.grad_records)
        res = torch.stack([rt.get_stats().num_steps_per_pixel.sum().double(), pc._xyz.grad.abs().sum().double()])
        out_host.copy_(res, non_blocking=False)  # device->host read of the step's result
        return float(out_host[0])

    for i in range(max(Wm, 3)):
        e2e_step(i)
    sync()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    e_rays = 0.0
    for i in range(K):
        e_rays += e2e_step(i)
    b.record()
    sync()
    e_ms = distributed.max_over_ranks(a.elapsed_time(b), dev)
    e_rays_t = torch.tensor([e_rays], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e_rays_t)
    e2e = {"value": float(e_rays_t.item()) / (e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": int(h2d),
           "d2h_bytes_per_step": int(d2h), "ms_per_step": e_ms / K,
           "api": "egr_b200.render(view, GaussianRaytracer) == reference renderer.render (export, refit, trace, import)"}

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu:
            cpu = cpu_sample("reference", scene, cams[0], args.cpu_pixels)
            cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}
        overflow = int(rt.get_stats().hitlog_status[2].item())
        line = {
            "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "num_gaussians": N, "resolution": [R, R], "views": len(cams),
                       "views_per_step": world, "sharding": "views sharded over ranks, scene+BVH replicated, one "
                       "all-reduce of the flat [N,24] gradient record" if world > 1 else "single GPU",
                       "l2": "inputs larger than L2: >400 MB of records/nodes/params/state/hit log touched per step, "
                             "a different view every step", "hitlog_overflow": overflow},
            "clocks": clocks, "e2e": e2e, "gpu_launches": (1 + 2 + S + (S - 1) + S + 1) * K,  # refit, pack+begin, S steps, S-1 scans, S backward, end
            "roofline": roofline, "cpu_baseline": cpu, "rays_per_step": rays / K / world,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--num-gaussians", type=int, default=1_000_000)
    ap.add_argument("--res", type=int, default=800)
    ap.add_argument("--hitlog-hits", type=int, default=120_000_000)
    ap.add_argument("--cpu-pixels", type=int, default=2048)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
