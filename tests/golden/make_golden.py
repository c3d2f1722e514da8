#!/usr/bin/env python
"""Generates tests/golden/ref_small.npz: known-answer vectors produced by the REFERENCE'S OWN shader sources
(editable_gauss_refl/cuda/csrc/{shaders,forward_pass,backward_pass}.cu + utils + core) compiled for the host
through oracle/ref_shim (needs the read-only checkout at /root/reference; run once, commit the .npz).

The reference's own tests hold no golden vector for the tracer (tests/test_gaussian_tracing.py only constructs
the object), so these fixtures are what pins oracle/egr_oracle.cpp.  Also stores the oracle's ordered hit-ID
lists on the same inputs as a regression pin (the reference build does not export them)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as orc  # noqa: E402
import egr_b200  # noqa: E402,F401
from egr_b200 import scenes  # noqa: E402

W, H, N = 24, 16, 2500
CFG = dict(num_bounces=2, jitter_primary_rays=1, loss_weight_diffuse=5.0, loss_weight_specular=3.0,
           loss_weight_depth=2.5, loss_weight_normal=2.5)
CALLS = 7


def main():
    assert orc.ref_available(), "needs /root/reference to build oracle/_ref"
    # larger, mostly translucent Gaussians with coherent normals: a 24x16 image then exercises several 16-wide
    # batches per ray (up to 42 composited hits), early termination (T < 0.01 on ~17 % of the rays), first and
    # second bounces
    sc = scenes.random_scene(N, seed=11)
    sc["scale"] += 1.0
    g5 = np.random.default_rng(5)
    sc["opacity"] = g5.uniform(-3.0, 0.0, (N, 1)).astype(np.float32)
    nz = sc["mean"] * 0.3 + np.array([0, 0, 1.0], np.float32)
    nz /= np.linalg.norm(nz, axis=1, keepdims=True)
    sc["normal"] = (nz * g5.uniform(0.85, 1, (N, 1))).astype(np.float32)
    cam = scenes.random_camera(W, H)
    tg = scenes.random_targets(W, H, seed=12)
    cfg = orc.make_config(**CFG)
    ocam = orc.make_camera(cam["origin"], cam["c2w"], cam["fovy"])
    out = {}
    for k in scenes.PARAM_KEYS:
        out["scene_" + k] = sc[k]
    for k, v in tg.items():
        out["target_" + k] = v
    out["cam_origin"], out["cam_c2w"], out["cam_fovy"] = cam["origin"], cam["c2w"], np.float32(cam["fovy"])
    fwd = orc.ref_render(sc, cfg, ocam, W, H, total_num_calls=CALLS)
    for k in list(orc.OUT_PLANES) + ["final", "random_seeds", "num_accumulated"]:
        out["ref_" + k] = fwd[k]
    bwd = orc.ref_render(sc, cfg, ocam, W, H, total_num_calls=CALLS, grads=True, targets=tg)
    for k in orc.GRAD_KEYS:
        out["ref_" + k] = bwd[k]
    # accumulation: two calls with accumulate_samples
    cfg_a = orc.make_config(accumulate_samples=1, **CFG)
    S = cfg.max_bounces + 1
    acc = {k: np.zeros((S, H, W, ch), np.float32) for k, ch in orc.ACC_PLANES.items()}
    acc["acc_count"] = 0
    orc.ref_render(sc, cfg_a, ocam, W, H, total_num_calls=1, accum=acc)
    acc["acc_count"] = 1
    a2 = orc.ref_render(sc, cfg_a, ocam, W, H, total_num_calls=2, accum=acc)
    out["ref_acc2_rgb"], out["ref_acc2_final"], out["ref_acc2_accumulated_rgb"] = a2["rgb"], a2["final"], acc["acc_rgb"]
    # oracle's own ordered hit-ID lists (regression pin)
    o = orc.render(sc, cfg, ocam, W, H, total_num_calls=CALLS, hit_cap=64)
    out["orc_hit_counts"], out["orc_hit_ids"] = o["hit_counts"], o["hit_ids"].astype(np.int16)
    out["orc_num_hits_per_pixel"] = o["num_hits_per_pixel"]
    # RNG known answers
    out["tea4"] = np.array([[a, b, orc.lib().orc_tea4(a, b)] for a, b in [(0, 0), (1, 1), (383, 7), (65535, 123456),
                                                                         (4294967295, 2)]], np.uint64)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_small.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
