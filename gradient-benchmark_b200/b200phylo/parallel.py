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

__all__ = ["shard_bounds", "rank_world", "allreduce_sum_", "attach_peers"]


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


def attach_peers(engine, group=None) -> bool:
    """Turn the engine's final reduction into the all-reduce (``b2p_peer_*``): every rank
    exports its receive buffer, the 64-byte CUDA IPC handles are all-gathered through the
    process group, and each rank maps its peers' buffers.  Returns False (and leaves the
    NCCL all-reduce in charge) when there is no group, a single rank, more than 8 ranks,
    ``B2P_NCCL_ALLREDUCE=1``, or when ANY rank fails to export/map — the decision is agreed
    across ranks so that all of them take the same path."""
    import os

    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return False
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    if world < 2 or world > 8 or os.environ.get("B2P_NCCL_ALLREDUCE") == "1":
        return False
    dev = torch.device(f"cuda:{engine.device}")
    try:
        handle = engine.peer_export(world)
    except Exception:
        handle = None
    ok = torch.tensor([1 if handle is not None else 0], dtype=torch.int32, device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
    if int(ok.item()) == 0:
        return False
    mine = torch.tensor(list(handle), dtype=torch.uint8, device=dev)
    every = [torch.empty(64, dtype=torch.uint8, device=dev) for _ in range(world)]
    dist.all_gather(every, mine, group=group)
    try:
        engine.peer_attach(rank, world, [bytes(t.cpu().tolist()) for t in every])
        attached = 1
    except Exception:
        attached = 0
    ok = torch.tensor([attached], dtype=torch.int32, device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
    if int(ok.item()) == 0:
        if attached:
            raise RuntimeError("peer attach succeeded on this rank but failed on another; cannot mix paths")
        return False
    dist.barrier(group=group)
    return True
