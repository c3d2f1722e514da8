"""CPU tier: the C-ABI library loads and exports every symbol include/egr_b200.h declares, argument checking
works without a GPU, host-side layouts / scenes / LBVH oracle / multi-rank logic behave."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import egr_b200
from egr_b200 import _capi, scenes
from oracle import lbvh_oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "egr_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(egr_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = _capi.load()
    names = _declared_symbols()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), n
        assert n in _capi.SYMBOLS, f"{n} declared in the header but not bound in _capi.SYMBOLS"
    assert set(_capi.SYMBOLS) == set(names)
    assert lib.egr_abi_version() == _capi.EGR_ABI_VERSION


def test_constants_match_reference_flags():
    lib = _capi.load()
    assert abs(lib.egr_max_alpha() - 0.9999) < 1e-7          # flags.h:8
    assert lib.egr_roughness_downweight_grad() == 1            # flags.h:12
    assert lib.egr_roughness_downweight_grad_power() == 3.0    # flags.h:13
    assert lib.egr_buffer_size() == 16 and lib.egr_max_iterations() == 99  # flags.h:16-17
    assert egr_b200.Raytracer.MAX_BOUNCES() == 2               # flags.h:5
    assert egr_b200.Raytracer.MAX_ALPHA() == pytest.approx(0.9999)


def test_struct_layout_matches_header():
    # every field of the C structs is one pointer (or documented scalar): sizes must agree with the header's layout
    assert C.sizeof(_capi.EgrCamera) == 6 * 8
    assert C.sizeof(_capi.EgrConfig) == 20 * 8
    assert C.sizeof(_capi.EgrFramebuffer) == 25 * 8
    assert C.sizeof(_capi.EgrGaussians) == 8 + 9 * 8
    assert C.sizeof(_capi.EgrMetadata) == 3 * 8 and C.sizeof(_capi.EgrStats) == 6 * 8
    assert C.sizeof(_capi.EgrParams) == 16 + (6 + 20 + 25 + 10 + 3 + 6) * 8 + 6 * 8 + 3 * 8 + 3 * 8 + 8  # ... + 3 debug outputs + 3 test inputs + early-stop threshold


def test_size_queries_and_argument_errors_without_gpu():
    lib = _capi.load()
    assert lib.egr_bvh_storage_bytes(0) < lib.egr_bvh_storage_bytes(1) < lib.egr_bvh_storage_bytes(1000)
    per = (lib.egr_bvh_storage_bytes(2_000_000) - lib.egr_bvh_storage_bytes(1_000_000)) / 1e6
    assert 64 + 32 + 48 + 64 + 4 * 4 <= per <= 310  # records + raw geometry + appearance + 2 nodes + index arrays + wide nodes (worst case n/2 x 144 B) per Gaussian
    assert lib.egr_pixel_state_bytes(800, 800, 3) >= 40 * 3 * 640000 * 4
    assert lib.egr_hitlog_bytes_for_hits(512) == 128 + 16 * 32 * 8  # {leaf, T_after} per logged hit
    assert lib.egr_raytrace(None, 0, None) == -1 and b"null" in lib.egr_last_error()
    g = _capi.EgrGaussians()
    g.count = 5
    cfg = _capi.EgrConfig()
    assert lib.egr_bvh_refit(C.byref(g), C.byref(cfg), None, 0, None) == -1
    assert b"null parameter pointer" in lib.egr_last_error()
    with pytest.raises(RuntimeError):
        _capi.check(-1)
    p = _capi.EgrParams()
    p.image_width, p.image_height, p.num_steps = 8, 8, 99
    assert lib.egr_raytrace(C.byref(p), 0, None) == -1 and b"num_steps" in lib.egr_last_error()
    p.num_steps = 3
    assert lib.egr_raytrace(C.byref(p), 0, None) == -2  # pixel_state missing -> ENOMEM-class error, no launch


def test_raytracer_fails_loudly_without_cuda():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        egr_b200.make_raytracer(8, 8, 4)


def test_set_pose_shape_errors_mirror_reference():
    import torch

    from egr_b200.raytracer import CameraDataHolder

    cam = CameraDataHolder(torch.device("cpu"))
    cam.set_pose(torch.ones(3), torch.eye(3))  # tests/test_gaussian_tracing.py:15-20 of the reference
    assert torch.equal(cam.rotation_w2c, torch.eye(3))
    with pytest.raises(RuntimeError, match="3x3"):
        cam.set_pose(torch.ones(3), torch.eye(4))
    with pytest.raises(RuntimeError, match="must be 3"):
        cam.set_pose(torch.ones(4), torch.eye(3))


def test_gaussian_holder_gradient_views_cpu():
    import torch

    from egr_b200.raytracer import GaussianDataHolder

    g = GaussianDataHolder(torch.device("cpu"))
    g.resize(7)
    assert g.grad_records.shape == (7, 24) and g.grad_records.is_contiguous()
    for name, ch in (("rgb", 3), ("normal", 3), ("f0", 3), ("roughness", 1), ("opacity", 1), ("scale", 3), ("mean", 3),
                     ("rotation", 4)):
        t = getattr(g, name)
        assert t.shape == (7, ch) and t.grad.shape == (7, ch)
        assert t.grad.data_ptr() == getattr(g, "dL_d" + name).data_ptr()
    g.grad_records.fill_(1.0)
    g.rgb.grad.zero_()
    assert float(g.grad_records.sum()) == 7 * 24 - 21 and float(g.total_weight.sum()) == 7
    cols = sorted(c for v in egr_b200.raytracer.GRAD_COLUMNS.values() for c in range(v[0], v[0] + v[1]))
    assert cols == [c for c in range(24) if c not in (19, 23)]  # 22 floats + 2 pads, no overlap


def test_lbvh_oracle_invariants():
    sc = scenes.random_scene(500, seed=2)
    sc["mean"][50:60] = sc["mean"][50]  # duplicate keys
    t = lbvh_oracle.build(sc["mean"])
    n = 500
    assert sorted(t["sorted_ids"].tolist()) == list(range(n))
    keys = lbvh_oracle.morton_keys(sc["mean"])[t["sorted_ids"]]
    assert np.all(np.diff(keys.astype(np.uint64)) >= 0)
    dup = t["sorted_ids"][np.isin(t["sorted_ids"], np.arange(50, 60))]
    assert dup.tolist() == list(range(50, 60))  # stable: equal keys keep id order
    skip, leaf, par = t["skip"], t["leaf"], t["parents"]
    assert skip[0] == 2 * n - 1 and par[0] == lbvh_oracle.NO_PARENT
    assert sorted(leaf[leaf != lbvh_oracle.INTERNAL].tolist()) == list(range(n))
    for i in range(2 * n - 1):
        if leaf[i] == lbvh_oracle.INTERNAL:
            l, r = i + 1, skip[i + 1]
            assert par[l] == i and par[r] == i and skip[r] == skip[i]
        else:
            assert skip[i] == i + 1
    one = lbvh_oracle.build(sc["mean"][:1])
    assert one["skip"].tolist() == [1] and one["leaf"].tolist() == [0]


def test_scenes_have_the_named_shapes():
    k = scenes.kitchen_scene(20000, seed=0)
    assert all(k[p].shape[0] == 20000 for p in scenes.PARAM_KEYS)
    assert np.allclose(np.linalg.norm(k["normal"], axis=1), 1.0, atol=1e-5)
    assert np.abs(k["mean"][:, :2]).max() <= 2.0 + 1e-5 and k["mean"][:, 2].min() >= -1e-6
    assert np.allclose(1 / (1 + np.exp(-k["opacity"])), 0.1, atol=1e-6)
    cams = scenes.kitchen_cameras(100, 800, 800)
    assert len(cams) == 100
    for c in cams[:5]:
        R = c["c2w"]
        assert np.allclose(R.T @ R, np.eye(3), atol=1e-5) and np.linalg.det(R) > 0
        back = scenes.c2w_to_reference_R(R)
        c2 = -back
        c2[:, 0] = -c2[:, 0]
        assert np.allclose(c2, R)
    r = scenes.random_scene(50_000, seed=0)
    assert r["mean"].shape == (50_000, 3) and r["rotation"].shape == (50_000, 4)


def _gloo_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist

    from egr_b200 import distributed

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        grads = torch.full((10, 24), float(rank + 1))
        distributed.allreduce_gradients(grads)
        views = [distributed.shard_views(7, rank, world, s) for s in range(7)]
        t = distributed.max_over_ranks(float(rank + 1), torch.device("cpu"))
        # one view split into horizontal bands: every rank holds only the rows of its own band
        Hh, W = 22, 5  # 6 tile rows (the last one ragged) over 2 ranks
        full = torch.arange(2 * Hh * W * 3, dtype=torch.float32).reshape(2, Hh, W, 3)
        a, b = distributed.band_for_rank(Hh, rank, world)
        mine = torch.full_like(full, -1.0)
        mine[:, a:b] = full[:, a:b]
        band_ok = bool(torch.equal(distributed.gather_bands(mine, Hh), full))
        q.put((rank, float(grads[0, 0]), float(grads.sum()), views, t, (a, b), band_ok))
    finally:
        dist.destroy_process_group()


def test_view_sharding_and_gradient_allreduce_gloo_world2():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, v00, total, views, t, band, band_ok in res:
        assert v00 == 3.0 and total == 3.0 * 240 and t == 2.0  # sum over ranks, max over ranks
        assert band_ok and band == ((0, 12), (12, 22))[rank]  # whole 4-pixel tile rows, together the whole image
    seen = sorted(res[0][3] + res[1][3])
    assert seen == sorted(list(range(7)) * 2)  # every view visited equally often, ranks never collide in a step
    assert all(a != b for a, b in zip(res[0][3], res[1][3]))
