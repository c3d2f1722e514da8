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
# --config: the driver's line is C2 (BASELINE.json configs[1], the one `metric` is quoted on); C3 / C4 run the same
# machinery on the other named shapes (BASELINE.json configs[2] / configs[3]) and label their lines accordingly.
CONFIGS = {
    "c2": dict(metric=METRIC, workload=WORKLOAD, scene="kitchen", n=1_000_000, w=800, h=800, bounces=1, grads=True,
               max_bounces=2),
    "c3": dict(metric="Mrays/s forward (3-bounce, 1M Gaussians, 1920x1080)",
               workload="C3 novel-view 1080p: 1,000,000 Gaussians (spread opacities), 1920x1080, 3 reflection bounces, "
                        "forward only (render.py / measure_fps.py)",
               scene="kitchen_varied", n=1_000_000, w=1920, h=1080, bounces=3, grads=False, max_bounces=3),
    "c4": dict(metric="Mrays/s fwd+bwd (1-bounce, 3M anisotropic Gaussians)",
               workload="C4 catacaustics-3M: 3,000,000 anisotropic Gaussians, 1024x768 views, train iteration "
                        "(refit+fwd+bwd), 1 bounce, views sharded over ranks",
               scene="catacaustics", n=3_000_000, w=1024, h=768, bounces=1, grads=True, max_bounces=2),
}


def kernel_sources_sha256():
    """Hash of the CUDA sources the library is built from (ties profiles/traffic.json to the kernels it was captured from)."""
    import hashlib

    from egr_b200 import build as _b

    h = hashlib.sha256()
    for f in sorted(_b.SOURCES + _b.HEADERS):
        with open(os.path.join(_b.CSRC, f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()


def make_scene(kind, n):
    from egr_b200 import scenes

    if kind == "catacaustics":
        return scenes.catacaustics_scene(n, seed=0)
    sc = scenes.kitchen_scene(n, seed=0)
    if kind == "kitchen_varied":  # a scene that has seen some optimisation: non-uniform opacities so that rays terminate
        sc["opacity"] = np.random.default_rng(4).uniform(-1.0, 4.0, (n, 1)).astype(np.float32)
    return sc


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

    scene = scenes.kitchen_scene(args.num_gaussians or 1_000_000, seed=0)
    cams = scenes.kitchen_cameras(100, args.res or 800, args.res or 800)
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
    registered = False
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        if args.nccl_register:
            registered = distributed.use_registered_gradient_records()
    peak_gbs, peak_kind = _peaks()

    cf = CONFIGS[args.config]
    N = args.num_gaussians or cf["n"]
    Wd, Hd = (args.res, args.res) if args.res else (cf["w"], cf["h"])
    P = Wd * Hd
    NB, GRADS = cf["bounces"], cf["grads"]
    scene = make_scene(cf["scene"], N)
    cams = scenes.kitchen_cameras(100, Wd, Hd)
    n_tg = 4
    targets = [scenes.random_targets(Wd, Hd, seed=1 + i) for i in range(n_tg)] if GRADS else [None] * n_tg
    pc = scenes.SyntheticModel(scene, device=dev)
    views_dev = [scenes.SyntheticCamera(cams[i], targets[i % n_tg], device=dev) for i in range(len(cams))]
    rtr = egr_b200.GaussianRaytracer(pc, Wd, Hd, ppll_forward_size=1000, ppll_backward_size=args.hitlog_hits,
                                     max_bounces=cf["max_bounces"])
    rt = rtr.cuda_module
    rt.get_config().num_bounces.fill_(NB)
    if args.early_stop > 0:  # opt-in mode, never the driver's line: labelled in `metric` and `config`
        rt.set_early_stop(args.early_stop)
    S = rt._S
    fb, cam_h = rt.get_framebuffer(), rt.get_camera()
    g = rt.get_gaussians()
    grad_ctx = torch.enable_grad if GRADS else torch.no_grad

    tkeys = (("diffuse", "diffuse_image"), ("specular", "specular_image"), ("depth", "depth_image"),
             ("normal", "normal_image"), ("f0", "f0_image"), ("roughness", "roughness_image"))
    for v in views_dev:  # poses resident on the device too; targets resident in the tracer's HWC layout
        Rm = torch.from_numpy(v.R).to(dev)
        v.c2w_dev = -Rm
        v.c2w_dev[:, 0] = -v.c2w_dev[:, 0]
        v.hwc = {k: getattr(v, attr).moveaxis(0, -1).contiguous() for k, attr in tkeys} if GRADS else {}
    torch.cuda.synchronize()

    def set_view_device(v):
        """Inputs already resident in HBM: pose + targets set from device tensors, no host traffic."""
        with torch.no_grad():
            cam_h.znear.fill_(0.01), cam_h.zfar.fill_(999.9), cam_h.vertical_fov_radians.fill_(v.FoVy)
            cam_h.set_pose(v.camera_center, v.c2w_dev)
            for k in v.hwc:
                getattr(fb, "target_" + k).copy_(v.hwc[k])

    # one event pair per launch where launches are serial; with gradients the bounce steps and the backward go in ONE
    # call because the library runs the backward of step s-1 underneath the forward of step s (side stream) within a call
    rest = _capi.EGR_PHASE_FINISH
    for s in range(1, S):
        rest |= _capi.EGR_PHASE_STEP(s)
    if GRADS:
        phases = [("begin", _capi.EGR_PHASE_BEGIN), ("trace_step0", _capi.EGR_PHASE_STEP(0)), ("bounce+backward", rest)]
    else:
        phases = ([("begin", _capi.EGR_PHASE_BEGIN)] + [(f"trace_step{s}", _capi.EGR_PHASE_STEP(s)) for s in range(NB + 1)] +
                  [("resolve", _capi.EGR_PHASE_FINISH)])
    rays_dev = torch.zeros((), dtype=torch.int64, device=dev)

    def new_pair():
        return torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def hot_step(view_idx, events=None):
        """refit + forward (+ backward) of one resident view (+ the gradient all-reduce when world > 1).
        `events`: pre-created {phase: (start, end)} pairs recorded around every launch of this step."""
        with grad_ctx():
            if GRADS:  # the reference refits in grad mode only (gaussian_raytracer.py:139-140)
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
        if world > 1 and GRADS:
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
    names = (["refit"] if GRADS else []) + [n for n, _ in phases] + (["allreduce"] if world > 1 and GRADS else [])
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
        hh, cc = [], []  # accepted / composited hits of each step (cumulative stats planes, read step by step)
        with torch.no_grad():
            v = views_dev[my_views[Wm + K - 1]]
            set_view_device(v)
            st = rt.get_stats()
            calls0 = rt.get_metadata().total_num_calls.clone()
            prev_h = prev_c = 0
            for nb in range(NB + 1):
                rt.get_config().num_bounces.fill_(nb)
                rt.get_metadata().total_num_calls.copy_(calls0)
                rt.raytrace()
                h_, c_ = int(st.num_hits_per_pixel.sum().item()), int(st.num_composited_per_pixel.sum().item())
                hh.append(h_ - prev_h), cc.append(c_ - prev_c)
                prev_h, prev_c = h_, c_
            rt.get_config().num_bounces.fill_(NB)
        # the launches that overlap inside the timed region (bounce forward, backward) are also timed one by one in a
        # short serial pass (one call per launch: no overlap), so that every kernel has a stand-alone duration
        serial = {}
        kernel_ms = {}
        if GRADS:
            ser_phases = [(f"trace_step{s}", _capi.EGR_PHASE_STEP(s)) for s in range(1, S)] + [("backward", _capi.EGR_PHASE_FINISH)]
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
            kernel_ms = {k: sum(a.elapsed_time(b) for a, b in v) / len(v) for k, v in serial.items()}
            kernel_ms["trace_step0"] = phase_ms["trace_step0"]  # serial inside the timed region: measured there
        else:
            kernel_ms = {f"trace_step{s}": phase_ms[f"trace_step{s}"] for s in range(NB + 1)}
        serial_ms = {k: v for k, v in kernel_ms.items() if k != "trace_step0" or not GRADS}
        alg = {f"trace_step{s}": 44 * hh[s] + 40 * cc[s] for s in range(NB + 1)}
        if GRADS:
            alg["backward"] = 260 * cc[0] + 176 * sum(cc[1:]) + 56 * P
        dom = max(alg, key=lambda k: kernel_ms.get(k, 0.0))
        achieved = alg[dom] / (kernel_ms[dom] * 1e-3) / 1e9
        # ncu dram__bytes_read+write of the same kernels (profiles/traffic.json, written from the round's ncu capture);
        # reported only while the library is not newer than that capture
        traffic, traffic_note = None, None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath) and args.config == "c2":
            try:
                tj = json.load(open(tpath))
                traffic = tj.get(dom)
                if tj.get("kernel_sources_sha256") != kernel_sources_sha256():
                    traffic_note = "profiles/traffic.json was captured from other kernel sources than the ones built now"
            except Exception:
                traffic = None
        kname = {f"trace_step{s}": f"k_trace_step(step={s})" for s in range(S)}
        kname["backward"] = "k_backward"
        roofline = {"bound": "hbm", "kernel": kname[dom],
                    "achieved": achieved, "peak": peak_gbs, "unit": "GB/s", "frac": achieved / peak_gbs,
                    "traffic": traffic, "traffic_note": traffic_note,
                    "peak_source": peak_kind + " (MEASURED_PEAKS.json hbm_gbs)",
                    "algorithmic_bytes_per_launch": alg[dom], "avg_launch_ms": kernel_ms[dom],
                    "timed_where": "timed region" if (dom == "trace_step0" or not GRADS) else "serial pass after the timed region",
                    "phase_ms": {k: round(v, 4) for k, v in phase_ms.items()},
                    "serial_kernel_ms": {k: round(v, 4) for k, v in serial_ms.items()},
                    "frac_per_kernel": {k: round(alg[k] / (kernel_ms[k] * 1e-3) / 1e9 / peak_gbs, 4) for k in alg if kernel_ms.get(k)},
                    "hits_per_ray": dict([(f"accepted_step{s}", hh[s] / P) for s in range(NB + 1)] +
                                         [(f"composited_step{s}", cc[s] / P) for s in range(NB + 1)])}

    # ---------------- end to end through the renderer facade with HOST buffers (`e2e`) ----------------
    views_host = [scenes.SyntheticCamera(cams[i], targets[i % n_tg], device="cpu", pin=True) for i in range(n_tg)]
    for v in views_host:
        v.camera_center = v.camera_center.pin_memory()
    h2d = 12 + 36
    if GRADS:
        h2d += sum(getattr(views_host[0], a).numel() * 4 for _, a in tkeys)
    out_host = torch.zeros(2, dtype=torch.float64).pin_memory()
    d2h = out_host.numel() * 8

    def e2e_step(i):
        v = views_host[i % n_tg]
        with grad_ctx():
            egr_b200.render(v, rtr, denoise=False, targets_available=GRADS)  # host->device copies of pose (+ 6 target images) happen inside
        if world > 1 and GRADS:
            distributed.allreduce_gradients(g.grad_records)
        probe = pc._xyz.grad.abs().sum().double() if GRADS else fb.output_final.sum().double()
        res = torch.stack([rt.get_stats().num_steps_per_pixel.sum().double(), probe])
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

    # ---------------- train iterations per second (the second half of BASELINE.json's `metric`) ----------------
    # One iteration = train.py:211-254: render() with the pose and the six target images in pinned HOST memory ->
    # all-reduce of the gradient record (world > 1) -> the fused optimiser step (gradient import + scale decay + Adam
    # + clamps + zeroing).  Each step's parameters feed the next export.  The prune -> compact -> rebind -> resize ->
    # export -> rebuild pipeline of every `pruning_interval`-th iteration is timed once and amortised.
    train = tr = None
    if GRADS and not args.no_train:
        from egr_b200 import trainer as trainer_mod

        rtr.zero_grad()
        tr = trainer_mod.Trainer(pc, rtr, pruning_interval=10 ** 9)  # the prune is timed separately below
        tviews = [views_host[(i * world + rank) % n_tg] for i in range(Wm + K)]  # a different view on every rank
        for i in range(max(Wm, 3)):
            tr.step(tviews[i % len(tviews)])
        sync()
        tnames = ["render", "allreduce", "optimizer"]
        tev = [{n: new_pair() for n in tnames} for _ in range(K)]
        a, b = new_pair()
        a.record()
        for i in range(K):
            tr.step(tviews[i], tev[i])
        b.record()
        sync()
        it_ms = distributed.max_over_ranks(a.elapsed_time(b), dev) / K
        tphase = {n: sum(e[n][0].elapsed_time(e[n][1]) for e in tev) / K for n in tnames}
        # pruning iteration: the 10 % of the Gaussians with the lowest accumulated weight go (same on all ranks: the
        # weights are all-reduced); the mask is made outside the timed region, train.py makes it with one compare
        tw = tr.total_weight.reshape(-1)
        n_before = int(tw.numel())
        keep = torch.ones(n_before, dtype=torch.uint8, device=dev)
        keep[torch.argsort(tw, stable=True)[: n_before // 10]] = 0  # the 10 % with the lowest accumulated weight
        tr.prune(torch.ones(n_before, dtype=torch.uint8, device=dev))  # untimed: the allocator grows its pools once
        sync()
        a, b = new_pair()
        a.record()
        n_after = tr.prune(keep)
        b.record()
        sync()
        prune_ms = distributed.max_over_ranks(a.elapsed_time(b), dev)
        for i in range(2):  # the model still trains after the prune (rebound optimiser state, rebuilt BVH)
            tr.step(tviews[i])
        sync()
        interval = 125  # config.py:75 pruning_interval
        amort = it_ms + prune_ms / interval
        train = {"iter_per_s": 1e3 / amort, "views_per_s": world * 1e3 / amort, "ms_per_iter": it_ms,
                 "phase_ms": {k: round(v, 4) for k, v in tphase.items()},
                 "prune_rebuild_ms": prune_ms, "pruned": [n_before, n_after], "pruning_interval": interval,
                 "iter_per_s_without_prune": 1e3 / it_ms,
                 "what": "render(host targets) + all-reduce + fused Adam step per iteration; prune+compact+rebind+"
                         "resize+export+rebuild timed once and amortised over the pruning interval",
                 "finite": bool(torch.isfinite(pc._xyz).all().item())}

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu and args.config == "c2":
            cpu = cpu_sample("reference", scene, cams[0], args.cpu_pixels)
            cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}
        overflow = int(rt.hitlog_overflowed())
        # grads: refit, pack+begin, S forward launches, S-1 scans, S backward launches, end; forward only: pack+begin,
        # NB+1 forward launches, NB scans, resolve, end
        n_launch = (1 + 2 + S + (S - 1) + S + 1) if GRADS else (2 + (NB + 1) + NB + 1 + 1)
        line = {
            "metric": cf["metric"] + (" [early-stop mode]" if args.early_stop > 0 else ""), "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": cf["workload"], "num_gaussians": N, "resolution": [Wd, Hd], "views": len(cams),
                       "views_per_step": world, "sharding": "views sharded over ranks, scene+BVH replicated, one "
                       "all-reduce of the flat [N,24] gradient record" + (" (NCCL-registered buffer)" if registered else "")
                       if world > 1 else "single GPU",
                       "l2": "inputs larger than L2: >400 MB of records/nodes/params/state/hit log touched per step, "
                             "a different view every step", "hitlog_overflow": overflow,
                       "early_stop_transmittance": args.early_stop},
            "clocks": clocks, "e2e": e2e, "gpu_launches": n_launch * K,
            "roofline": roofline, "cpu_baseline": cpu, "rays_per_step": rays / K / world, "train": train,
        }
        print(json.dumps(line))
    if world > 1:
        if registered:
            # the registered pool frees its memory through NCCL: drop every tensor that lives in it first
            sys.stdout.flush()
            del g, tr, rtr, rt, pc
            import gc

            gc.collect()
            torch.cuda.synchronize()
            distributed.release_registered_pools()
            torch.cuda.empty_cache()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c2", choices=sorted(CONFIGS), help="named shape of BASELINE.json (default: the metric's C2)")
    ap.add_argument("--num-gaussians", type=int, default=0, help="override the config's Gaussian count")
    ap.add_argument("--res", type=int, default=0, help="override the config's resolution (square)")
    ap.add_argument("--no-train", action="store_true", help="skip the train-iteration leg")
    ap.add_argument("--early-stop", type=float, default=0.0,
                    help="opt-in early-stop mode: total-transmittance threshold (0 = exact mode, the default and the metric's)")
    ap.add_argument("--nccl-register", type=int, default=1, help="1: gradient record from NCCL's allocator, registered (multi-GPU)")
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
