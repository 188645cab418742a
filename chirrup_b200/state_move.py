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
    dev = torch.device(device) if device is not None else state[0].device
    if dev.type == "cuda" and state[0].is_cuda:                 # device target: strided DMA copies (b200_slot_copy)
        s0, s1 = state[0], state[1]
        out = [torch.empty((s0.shape[0], 2, 1, s0.shape[3]), dtype=s0.dtype, device=dev),
               torch.empty((s1.shape[0], 1) + tuple(s1.shape[2:]), dtype=s1.dtype, device=dev),
               state[2][slot:slot + 1].to(dev, non_blocking=non_blocking, copy=True)]
        if dev != s0.device:
            _enable_peer(s0.device, dev)
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(s0.device))
            torch.cuda.current_stream(dev).wait_event(ev)
        with torch.cuda.device(dev):
            _copy_rows(out[0][:, :, 0], s0[:, :, slot])
            _copy_rows(out[1][:, 0], s1[:, slot])
        return out
    out = [state[0][:, :, slot:slot + 1].to(dev, non_blocking=non_blocking, copy=True),
           state[1][:, slot:slot + 1].to(dev, non_blocking=non_blocking, copy=True),
           state[2][slot:slot + 1].to(dev, non_blocking=non_blocking, copy=True)]
    return out


def inject_slot(state: List[torch.Tensor], slot: int, packed: List[torch.Tensor], non_blocking: bool = True) -> None:
    """``batch_state[i][..., slot, ...] = task.state[i]`` (``chirrup/worker.py:595-597``); source may live on another GPU or on
    the host."""
    p0 = packed[0].view(state[0].shape[0], 2, 1, -1)
    p1 = packed[1].view(state[1].shape[0], 1, *state[1].shape[2:])
    if p0.is_cuda and state[0].is_cuda and p0.device == state[0].device and p0.dtype == state[0].dtype:
        with torch.cuda.device(state[0].device):
            _copy_rows(state[0][:, :, slot], p0[:, :, 0])
            _copy_rows(state[1][:, slot], p1[:, 0])
    else:
        state[0][:, :, slot:slot + 1].copy_(p0, non_blocking=non_blocking)
        state[1][:, slot:slot + 1].copy_(p1, non_blocking=non_blocking)
    state[2][slot:slot + 1].copy_(packed[2].view(1), non_blocking=non_blocking)


def clear_slot(state: List[torch.Tensor], slot: int) -> None:
    state[0][:, :, slot].zero_()
    state[1][:, slot].zero_()
    state[2][slot] = 0


_peers_enabled = set()


def _enable_peer(a: torch.device, b: torch.device) -> None:
    """torch turns peer access on the first time it copies between two GPUs; do that once per pair so that the strided DMA
    copies below go over NVLink instead of being staged through host memory by the driver."""
    key = (a.index, b.index)
    if key in _peers_enabled:
        return
    torch.zeros(1, device=a).to(b)
    torch.zeros(1, device=b).to(a)
    _peers_enabled.add(key)
    _peers_enabled.add((b.index, a.index))


def _rows_view(t: torch.Tensor):
    """(pointer, pitch bytes, width bytes, rows) of a view that is `rows` contiguous pieces at one pitch, else None."""
    if t.dim() < 2 or t.dtype != torch.float16:
        return None
    # state[1][:, s]: (L, H, 64, 64), inner block contiguous; state[0][:, :, s]: (L, 2, C), rows (l, which) at a uniform pitch
    if t.dim() == 4 and t[0].is_contiguous():
        return t.data_ptr(), t.stride(0) * 2, t[0].numel() * 2, t.shape[0]
    if t.dim() == 3 and t.stride(2) == 1 and t.stride(0) == 2 * t.stride(1) and t.shape[1] == 2:
        return t.data_ptr(), t.stride(1) * 2, t.shape[2] * 2, 2 * t.shape[0]
    return None


def _copy_rows(dst: torch.Tensor, src: torch.Tensor) -> None:
    """dst <- src for one slot's rows: one strided DMA copy (``b200_slot_copy``) on the current stream of dst's device when both
    are CUDA views with the slot-table layout, else a torch copy (host tiers)."""
    dv, sv = _rows_view(dst), _rows_view(src)
    if dst.is_cuda and src.is_cuda and dv and sv and dv[2:] == sv[2:]:
        from . import _lib
        _lib.check(_lib.lib().b200_slot_copy(dv[0], dv[1], sv[0], sv[1], dv[2], dv[3],
                                             torch.cuda.current_stream(dst.device).cuda_stream), "b200_slot_copy")
    else:
        dst.copy_(src, non_blocking=True)


def migrate_slot(src: List[torch.Tensor], src_slot: int, dst: List[torch.Tensor], dst_slot: int) -> int:
    """Move one request's state between two workers' slot tables (possibly on different GPUs): two strided DMA copies
    (``b200_slot_copy``: the L WKV tiles, the 2L shift rows) + the elapsed counter, device to device — over NVLink between
    peers, never through host memory, no SM work.  Returns the bytes moved.  Ordered on the destination's current stream after
    the source's current stream."""
    sdev, ddev = src[0].device, dst[0].device
    if sdev != ddev and sdev.type == "cuda" and ddev.type == "cuda":
        _enable_peer(sdev, ddev)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(sdev))
        torch.cuda.current_stream(ddev).wait_event(ev)
    with torch.cuda.device(ddev) if ddev.type == "cuda" else _Null():
        _copy_rows(dst[0][:, :, dst_slot], src[0][:, :, src_slot])
        _copy_rows(dst[1][:, dst_slot], src[1][:, src_slot])
        dst[2][dst_slot:dst_slot + 1].copy_(src[2][src_slot:src_slot + 1], non_blocking=True)
    return slot_bytes(src)


def migrate_slots(src: List[torch.Tensor], src_slots, dst: List[torch.Tensor], dst_slots) -> int:
    """Several slots between two CUDA slot tables with the per-slot host work cut to ONE C call (``b200_slot_move``: both strided
    copies and the elapsed counter) and ONE cross-device ordering event for the whole batch — the Python bookkeeping of
    ``migrate_slot`` (three calls and an event pair per slot, ~100 us) is what bounded the migration rate in round 1, not the
    link.  Copies run on the destination's current stream.  Returns the bytes moved."""
    from . import _lib

    sdev, ddev = src[0].device, dst[0].device
    assert sdev.type == "cuda" and ddev.type == "cuda"
    if sdev != ddev:
        _enable_peer(sdev, ddev)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(sdev))
        torch.cuda.current_stream(ddev).wait_event(ev)
    s0, s1, d0, d1 = src[0], src[1], dst[0], dst[1]
    L = s1.shape[0]
    w1, w0 = s1[0, 0].numel() * 2, s0.shape[3] * 2
    sp1, dp1 = s1.stride(0) * 2, d1.stride(0) * 2                      # layer pitch of the WKV tiles
    sp0, dp0 = s0.stride(1) * 2, d0.stride(1) * 2                      # (layer, which) rows at a uniform pitch
    assert s0.stride(0) == 2 * s0.stride(1) and d0.stride(0) == 2 * d0.stride(1)
    sb1, db1, sb0, db0 = s1.data_ptr(), d1.data_ptr(), s0.data_ptr(), d0.data_ptr()
    se, de = src[2].data_ptr(), dst[2].data_ptr()
    ss1, ds1, ss0, ds0 = s1.stride(1) * 2, d1.stride(1) * 2, s0.stride(2) * 2, d0.stride(2) * 2
    fn = _lib.lib().b200_slot_move
    with torch.cuda.device(ddev):
        stream = torch.cuda.current_stream(ddev).cuda_stream
        for a, b in zip(src_slots, dst_slots):
            rc = fn(db1 + b * ds1, dp1, sb1 + a * ss1, sp1, w1, L, db0 + b * ds0, dp0, sb0 + a * ss0, sp0, w0, 2 * L,
                    de + 4 * b, se + 4 * a, stream)
            if rc:
                _lib.check(rc, "b200_slot_move")
    return slot_bytes(src) * len(list(src_slots))


class _Null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
