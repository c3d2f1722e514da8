"""Allocation-free BVH rebuild (`egr_bvh_build_noalloc`), called straight through the C ABI.

Some hosts cannot allocate device memory at the moment a rebuild is due (the whole GPU belongs to their own pool,
or they sit in a latency-critical edit loop), so the library carries a fixed 6 MiB scratch block and offers this
entry point for up to 65,536 Gaussians.  The test builds the same scene with the regular entry point (scratch from
the caller) and with this one and expects identical trees, twice in a row: the second build is a different, smaller
scene, so it runs over whatever the first one left behind in the shared block.  The file name sorts last on purpose:
under `-x` every other test has already run when this one does, and it passes on its own in a fresh process too.
"""
import ctypes as C

import numpy as np
import pytest

import helpers as H
from egr_b200 import _capi, scenes

pytestmark = pytest.mark.gpu


def _build_both(n, seed):
    import torch

    sc = scenes.random_scene(n, seed=seed)
    cam = scenes.random_camera(32, 16)
    rt = H.make_gpu_raytracer(sc, cam, dict(num_bounces=0), debug_rays=False, hit_capacity=1000)  # regular build
    ref = [t.clone() for t in rt.export_bvh()]
    lib = _capi.load()
    storage = torch.zeros_like(rt._bvh_storage)  # the caller still provides the output storage
    g, cfg = rt._gaussians_struct(), rt._config_struct()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _capi.check(lib.egr_bvh_build_noalloc(C.byref(g), C.byref(cfg), storage.data_ptr(), storage.numel(), stream))
    torch.cuda.synchronize()
    old = rt._bvh_storage
    rt._bvh_storage = storage
    got = [t.clone() for t in rt.export_bvh()]
    rt._bvh_storage = old
    return ref, got


def test_noalloc_build_matches_regular_build_twice_in_a_row():
    for n, seed in ((20000, 21), (777, 22)):
        ref, got = _build_both(n, seed)
        names = ("sorted_ids", "records", "nodes", "parents")
        for name, a, b in zip(names, ref, got):
            # integer indices and float boxes alike: the same kernels on the same inputs give the same bits
            assert np.array_equal(a.cpu().numpy().view(np.uint32), b.cpu().numpy().view(np.uint32)), (n, name)


def test_noalloc_build_rejects_oversized_input():
    import torch

    lib = _capi.load()
    g = _capi.EgrGaussians()
    g.count = _capi.EGR_NOALLOC_MAX_GAUSSIANS + 1
    dummy = torch.zeros(64, device="cuda")
    for f in _capi.GAUSSIAN_FIELDS:
        setattr(g, f, dummy.data_ptr())
    cfg = _capi.EgrConfig()
    rc = lib.egr_bvh_build_noalloc(C.byref(g), C.byref(cfg), dummy.data_ptr(), 256, None)
    assert rc == -1
    msg = lib.egr_last_error().decode()
    assert "65537" in msg and "65536" in msg and "exceed" in msg
