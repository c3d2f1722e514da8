"""Two-rank NCCL job on one box (skipped below two GPUs; the round's own run is kept in profiles/): views sharded over
the ranks, scene + BVH replicated, ONE all-reduce of the flat gradient record per iteration.

Checks (a) the all-reduced record equals the single-GPU sum over the same two views (float atomics reorder sums),
(b) after five iterations -- including a pruning iteration and an insertion of random points generated differently on
each rank and broadcast from rank 0 -- parameters, Adam moments and the BVH's Morton order are bit-identical on both
ranks, and (c) total_num_calls advanced in lock-step."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist

    import egr_b200
    from egr_b200 import distributed, scenes, trainer

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    W, Hh, n = 64, 48, 3000
    sc = scenes.random_scene(n, seed=0)
    cams = [scenes.random_camera(W, Hh, distance=d) for d in (4.0, 3.6, 4.4, 3.8)]
    tgs = [scenes.random_targets(W, Hh, seed=10 + i) for i in range(4)]
    views = [scenes.SyntheticCamera(c, t, device=dev) for c, t in zip(cams, tgs)]
    pc = scenes.SyntheticModel(sc, device=dev)
    rtr = egr_b200.GaussianRaytracer(pc, W, Hh, ppll_forward_size=1000, ppll_backward_size=4_000_000)
    rt = rtr.cuda_module
    rt.get_config().num_bounces.fill_(1)
    res = {}
    # (a) one iteration's all-reduced record vs the single-GPU sum over the same two views
    rtr.import_gradients = False
    with torch.enable_grad():
        egr_b200.render(views[distributed.shard_views(len(views), rank, world, 0)], rtr)
    rec = rt.get_gaussians().grad_records
    distributed.allreduce_gradients(rec)
    summed = rec.clone()
    rec.zero_()
    if rank == 0:
        single = torch.zeros_like(summed)
        for r in range(world):
            rt.get_metadata().total_num_calls.zero_()
            with torch.enable_grad():
                egr_b200.render(views[distributed.shard_views(len(views), r, world, 0)], rtr)
            single += rec
            rec.zero_()
        scale = single.abs().amax(dim=0).clamp_min(1e-20)
        res["allreduce_vs_single_gpu_sum_rel"] = float(((summed - single).abs() / scale).max())
        rt.get_metadata().total_num_calls.fill_(1)
    # (b) five iterations with a pruning iteration (3rd) and an insertion (after the 4th)
    tr = trainer.Trainer(pc, rtr, pruning_interval=3, pruning_start_iter=0, min_weight=0.02)
    for it in range(5):
        tr.step(views[distributed.shard_views(len(views), rank, world, it + 1)])
        if it == 3:
            gen = torch.Generator(device=dev).manual_seed(100 + rank)  # DIFFERENT rows on every rank ...
            tr.add_points(tr.farfield_points(200 + 50 * rank, extent=1.0, generator=gen))  # ... rank 0's win
    ids, _, _, _ = rt.export_bvh()
    res["calls_after_training"] = int(rt.get_metadata().total_num_calls)
    mine = [getattr(pc, a) for a in trainer.MODEL_ATTR.values()] + list(tr.optimizer.exp_avg.values()) + \
           list(tr.optimizer.exp_avg_sq.values()) + [ids.float(), rt.get_metadata().total_num_calls.float()]
    flat = torch.cat([t.reshape(-1).float() for t in mine])
    sizes = torch.tensor([flat.numel()], device=dev)
    all_sizes = [torch.zeros_like(sizes) for _ in range(world)]
    dist.all_gather(all_sizes, sizes)
    res["sizes"] = [int(s) for s in all_sizes]
    if len(set(res["sizes"])) == 1:
        other = [torch.zeros_like(flat) for _ in range(world)]
        dist.all_gather(other, flat)
        res["replicas_bit_identical"] = all(bool(torch.equal(other[0], o)) for o in other[1:])
    else:
        res["replicas_bit_identical"] = False
    # (d) ONE view split into horizontal bands over the ranks: gathered planes == the single-GPU image, bit for bit
    rtr.import_gradients = False
    v0 = views[0]
    with torch.no_grad():
        rt.get_metadata().total_num_calls.zero_()
        rt.set_pixel_band(*distributed.band_for_rank(Hh, rank, world))
        egr_b200.render(v0, rtr, targets_available=False, force_update_bvh=True)
        fb = rt.get_framebuffer()
        banded = {k: distributed.gather_bands(getattr(fb, k), Hh) for k in ("output_rgb", "output_depth", "output_final")}
        rt.set_pixel_band(None)
        rt.get_metadata().total_num_calls.zero_()
        egr_b200.render(v0, rtr, targets_available=False)
        res["bands_equal_full_image"] = all(bool(torch.equal(banded[k], getattr(fb, k))) for k in banded)
    res["count"] = int(pc._xyz.shape[0])
    res["calls"] = int(rt.get_metadata().total_num_calls)
    res["finite"] = bool(torch.isfinite(pc._xyz).all())
    res["overflow"] = int(rt.hitlog_overflowed())
    if rank == 0:
        import json

        with open(os.path.join(out_dir, "result.json"), "w") as f:
            json.dump(res, f)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_nccl_training_keeps_replicas_identical(tmp_path):
    import json

    import torch
    import torch.multiprocessing as mp

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run under `gpurun --gpus 2`)")
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    res = json.load(open(tmp_path / "result.json"))
    print("two-rank result:", res)
    assert res["allreduce_vs_single_gpu_sum_rel"] <= 1e-4, res
    assert res["replicas_bit_identical"] and len(set(res["sizes"])) == 1, res
    assert res["bands_equal_full_image"], res
    assert res["finite"] and res["overflow"] == 0 and res["calls_after_training"] == 6, res
    assert 3000 - 2500 < res["count"] < 3000 + 200, res  # pruned some, then rank 0's 200 rows were inserted
