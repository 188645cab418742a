"""Multi-GPU plumbing of the request-sharded path: one replica per GPU, no collective on the step path.  torch.distributed is
used only to line the ranks up (barrier) and to take the max over ranks of a device-side time; backend nccl on GPUs, gloo in
the CPU tests."""
from __future__ import annotations

import os
from typing import List, Sequence

import torch
import torch.distributed as dist


def env_rank_world():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def shard_requests(n_requests: int, rank: int, world: int) -> range:
    """Contiguous, balanced shard of request ids for this rank (sizes differ by at most one; every id exactly once)."""
    base, extra = divmod(n_requests, world)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


def max_over_ranks(values: Sequence[float], device="cpu") -> List[float]:
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


def sum_over_ranks(values: Sequence[float], device="cpu") -> List[float]:
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.tolist()


def barrier() -> None:
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()
