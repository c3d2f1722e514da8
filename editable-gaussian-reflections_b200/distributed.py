"""Multi-GPU: rays shard, the scene and its BVH are replicated, ONE collective sums the gradients.

The reference is single-GPU (no torch.distributed anywhere).  Here each rank (one process per GPU) renders a
different training view per optimizer step; afterwards the flat ``[N, 24]`` gradient record (the eight ``dL_d*``
tensors and ``total_weight`` are views of it) is summed over ranks with a single all-reduce over NCCL/NVLink (or
gloo on CPU for tests).  Every rank then applies the identical optimizer step, so replicas stay bit-identical
without broadcasting parameters.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_views(num_views: int, rank: int, world_size: int, step: int):
    """Index of the view rank ``rank`` renders at optimizer step ``step`` (rank-strided walk, replaces the
    reference's single-process random pop, train.py:218-220)."""
    return (step * world_size + rank) % num_views


def band_for_rank(image_height: int, rank: int, world_size: int):
    """Rows [begin, end) of the horizontal band rank ``rank`` traces when ONE view is split over ``world_size`` GPUs
    (the secondary split of SURVEY 8e: re-render latency of an edit, exact single-view semantics).  Bands are whole
    4-pixel tile rows, as even as possible; feed them to ``Raytracer.set_pixel_band``."""
    rows = (image_height + 3) // 4
    lo, hi = rows * rank // world_size, rows * (rank + 1) // world_size
    return 4 * lo, min(4 * hi, image_height)


def gather_bands(plane: torch.Tensor, image_height: int, group=None) -> torch.Tensor:
    """Assemble a framebuffer plane ``[S,H,W,C]`` (or ``[1,H,W,C]``) whose rows were traced band-wise: every rank
    contributes the rows of its own band (one all-gather of equal-sized padded bands)."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return plane
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    bands = [band_for_rank(image_height, r, world) for r in range(world)]
    hmax = max(b - a for a, b in bands)
    a, b = bands[rank]
    mine = torch.zeros(plane.shape[0], hmax, *plane.shape[2:], dtype=plane.dtype, device=plane.device)
    mine[:, : b - a] = plane[:, a:b]
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine, group=group)
    out = torch.empty_like(plane)
    for (a, b), part in zip(bands, parts):
        out[:, a:b] = part[:, : b - a]
    return out


def allreduce_gradients(grad_records: torch.Tensor, group=None, async_op: bool = False):
    """Sum the flat gradient record over all ranks, in place.  One call covers all 22 floats per Gaussian."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return None
    if not grad_records.is_contiguous():
        raise RuntimeError("grad_records must be the contiguous flat [N,24] buffer")
    return dist.all_reduce(grad_records, op=dist.ReduceOp.SUM, group=group, async_op=async_op)


def max_over_ranks(value_ms: float, device) -> float:
    """Max over ranks of a device-timed duration (multi-GPU numbers are never wall-clock)."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size() == 1:
        return value_ms
    t = torch.tensor([value_ms], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


_pools = []  # registered pools stay alive (and registered) for the life of the process group


def use_registered_gradient_records(group=None) -> bool:
    """From now on the tracer's flat gradient record is allocated from NCCL's allocator (``ncclMemAlloc``) in a memory
    pool registered with the communicator, which lets NCCL run the all-reduce on user buffers (no staging copies) and
    through the NVSwitch reduction (NVLS) where the fabric offers it.  Returns False -- and changes nothing -- when the
    backend does not expose the allocator (CPU / gloo, older builds).  Call it after ``init_process_group`` and before
    the ``Raytracer`` is constructed; ``resize()`` allocates the new record from the same pool."""
    from . import _capi, raytracer

    if not dist.is_available() or not dist.is_initialized() or not torch.cuda.is_available():
        return False
    try:
        dev = torch.device("cuda", torch.cuda.current_device())
        backend = (group or dist.group.WORLD)._get_backend(dev)
        pool = torch.cuda.MemPool(backend.mem_allocator)
    except Exception:
        return False

    def alloc(n, device):
        with torch.cuda.use_mem_pool(pool):
            t = torch.zeros(n, _capi.EGR_GRAD_STRIDE, dtype=torch.float32, device=device)
        try:
            backend.register_mem_pool(pool)  # (re-)registers the pool's segments, the new one included
        except Exception:
            pass
        return t

    _pools.append((pool, backend))
    raytracer.GRAD_RECORD_ALLOCATOR = alloc
    return True


def release_registered_pools() -> None:
    """Deregister the pools made by :func:`use_registered_gradient_records` and restore the default allocator.  Call it
    before ``destroy_process_group`` (the pools free their memory through NCCL, which needs a live communicator);
    tensors allocated from the pools must not be used afterwards."""
    from . import raytracer

    raytracer.GRAD_RECORD_ALLOCATOR = raytracer._default_record_allocator
    while _pools:
        pool, backend = _pools.pop()
        try:
            backend.deregister_mem_pool(pool)
        except Exception:
            pass
        del pool
