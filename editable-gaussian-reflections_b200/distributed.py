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
