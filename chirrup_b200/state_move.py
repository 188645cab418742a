"""State-slot movement: admission / extraction / migration of one request's RWKV state.

Replaces the reference's host round trips (``chirrup/worker.py:427-429``, ``:468-470`` slot -> CPU; ``:586-597`` CPU -> slot) and
its three-copy physical slot swap (``:304-360``).  One slot = ``state[0][:, :, s]`` (L,2,C) + ``state[1][:, s]`` (L,H,64,64) fp16 +
``state[2][s]`` int32 = 16.5 MiB at 7.2B, 31.5 MiB at 13.3B.  Between GPUs the copy goes device to device (NVLink P2P when the
driver enables peer access; torch issues cudaMemcpyPeerAsync) — never through host memory.  No arithmetic happens here.
"""
from __future__ import annotations

from typing import List

import torch


def slot_bytes(state: List[torch.Tensor]) -> int:
    s0, s1, _ = state
    return s0[:, :, 0].numel() * 2 + s1[:, 0].numel() * 2 + 4


def extract_slot(state: List[torch.Tensor], slot: int, device=None, non_blocking: bool = True) -> List[torch.Tensor]:
    """Copy of one slot in the reference's single-request layout ((L,2,1,C), (L,1,H,64,64), (1,)) — what Chirrup caches
    (``chirrup/utils/state_cache.py``) and re-injects through ``Task.state``."""
    dev = device if device is not None else state[0].device
    out = [state[0][:, :, slot:slot + 1].to(dev, non_blocking=non_blocking, copy=True),
           state[1][:, slot:slot + 1].to(dev, non_blocking=non_blocking, copy=True),
           state[2][slot:slot + 1].to(dev, non_blocking=non_blocking, copy=True)]
    return out


def inject_slot(state: List[torch.Tensor], slot: int, packed: List[torch.Tensor], non_blocking: bool = True) -> None:
    """``batch_state[i][..., slot, ...] = task.state[i]`` (``chirrup/worker.py:595-597``); source may live on another GPU or on
    the host."""
    state[0][:, :, slot:slot + 1].copy_(packed[0].view(state[0].shape[0], 2, 1, -1), non_blocking=non_blocking)
    state[1][:, slot:slot + 1].copy_(packed[1].view(state[1].shape[0], 1, *state[1].shape[2:]), non_blocking=non_blocking)
    state[2][slot:slot + 1].copy_(packed[2].view(1), non_blocking=non_blocking)


def clear_slot(state: List[torch.Tensor], slot: int) -> None:
    state[0][:, :, slot].zero_()
    state[1][:, slot].zero_()
    state[2][slot] = 0


def migrate_slot(src: List[torch.Tensor], src_slot: int, dst: List[torch.Tensor], dst_slot: int) -> int:
    """Move one request's state between two workers' slot tables (possibly on different GPUs) with strided device-to-device
    copies; returns the bytes moved.  Ordered on the destination's current stream after the source's current stream."""
    sdev, ddev = src[0].device, dst[0].device
    if sdev != ddev and sdev.type == "cuda" and ddev.type == "cuda":
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(sdev))
        torch.cuda.current_stream(ddev).wait_event(ev)
    with torch.cuda.device(ddev) if ddev.type == "cuda" else _Null():
        dst[0][:, :, dst_slot].copy_(src[0][:, :, src_slot], non_blocking=True)
        dst[1][:, dst_slot].copy_(src[1][:, src_slot], non_blocking=True)
        dst[2][dst_slot:dst_slot + 1].copy_(src[2][src_slot:src_slot + 1], non_blocking=True)
    return slot_bytes(src)


class _Null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
