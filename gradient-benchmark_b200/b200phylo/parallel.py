"""Site-pattern sharding across the GPUs of one box.

Patterns are independent through the whole traversal; the only coupling is the final
sums ``logL = sum_p ...`` and ``g_b = sum_p ...``.  Each rank owns a contiguous slice
of patterns, runs the full traversal on it, and ONE all-reduce of ``2N-1`` doubles
(``[logL, g_0 .. g_{2N-3}]``) over NCCL/NVLink finishes an evaluation.  There is no
other data-path collective.

The reference has no counterpart (every contender is one CPU process,
/root/reference/nextflow.config:9); this is the one data-parallel axis BASELINE.json's
north_star adds.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

__all__ = ["shard_bounds", "rank_world", "allreduce_sum_"]


def shard_bounds(n_patterns: int, world_size: int, rank: int) -> Tuple[int, int]:
    """Contiguous, balanced-by-count pattern slice ``[start, stop)`` of ``rank``.

    The first ``n_patterns % world_size`` ranks get one extra pattern.  A rank may end
    up with an empty slice only when ``n_patterns < world_size`` (rejected by callers).
    """
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError(f"bad rank/world: {rank}/{world_size}")
    base, extra = divmod(int(n_patterns), int(world_size))
    start = rank * base + min(rank, extra)
    stop = start + base + (1 if rank < extra else 0)
    return start, stop


def rank_world(group=None) -> Tuple[int, int]:
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def allreduce_sum_(buf: torch.Tensor, group=None) -> torch.Tensor:
    """In-place sum of ``buf`` ([logL, grad...], float64) over all ranks.

    Runs on the tensor's own device/stream: NCCL for CUDA tensors (the product path),
    gloo for CPU tensors (host-logic tests).  A no-op without an initialised group.
    """
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    return buf
