"""Operator layer: torch tensors in, C-ABI calls out.  Mirrors the reference's op surface
(``torch.ops.rwkv7_state_fwd_fp16.{forward_seq,forward_one,spmv_forward}``, ``/root/reference/Albatross/rwkv7.py:61-160``)
plus the dense linear op.  torch is used for device memory and the current stream only.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from . import _lib

HEAD = 64
EPI_NONE, EPI_TANH, EPI_SIGMOID, EPI_RELU_SQ, EPI_RESADD = 0, 1, 2, 3, 4


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _req(t: torch.Tensor, name: str, dtype=torch.float16) -> None:
    if not t.is_cuda:
        raise _lib.B200Error(f"{name}: chirrup_b200 has no CPU path; got a {t.device} tensor")
    if t.dtype != dtype:
        raise TypeError(f"{name}: expected {dtype}, got {t.dtype}")


def wkv7_forward_seq(state: torch.Tensor, r, w, k, v, a, b, elapsed_t: torch.Tensor,
                     slot_idx: Optional[torch.Tensor] = None, y: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``RWKV7_BATCH_OP`` / ``forward_seq`` (reference ``Albatross/rwkv7.py:144-160``): r..b are (B,T,C) fp16, ``state`` is
    (n_slots,H,64,64) fp16 (a B-axis view is fine) updated in place, ``elapsed_t`` int32 (B,).  Returns y (B,T,C)."""
    B, T, C = r.shape
    H = C // HEAD
    for n, t in (("state", state), ("r", r), ("w", w), ("k", k), ("v", v), ("a", a), ("b", b)):
        _req(t, n)
    _req(elapsed_t, "elapsed_t", torch.int32)
    if B == 0 or T == 0:
        return y if y is not None else torch.empty((B, T, C), dtype=torch.float16, device=r.device)
    if state.dim() != 4 or tuple(state.shape[1:]) != (H, HEAD, HEAD) or not state[0].is_contiguous():
        raise ValueError(f"state must be (n_slots,{H},64,64) with contiguous slots, got {tuple(state.shape)}")
    ins = [t if t.is_contiguous() else t.contiguous() for t in (r, w, k, v, a, b)]
    if y is None:
        y = torch.empty((B, T, C), dtype=torch.float16, device=r.device)
    et = elapsed_t.reshape(-1)
    if et.numel() != B:
        raise ValueError("elapsed_t must have B elements")
    et = et if et.is_contiguous() else et.contiguous()
    n_slots = state.shape[0]
    sp = 0
    if slot_idx is not None:
        _req(slot_idx, "slot_idx", torch.int32)
        sp = slot_idx.data_ptr()
    elif n_slots < B:
        raise ValueError("state has fewer slots than sequences")
    stride = state.stride(0) if n_slots > 1 else C * HEAD
    rc = _lib.lib().b200_wkv7_forward_seq(B, T, C, H, state.data_ptr(), stride, n_slots, *[t.data_ptr() for t in ins],
                                          y.data_ptr(), et.data_ptr(), sp, _stream(r))
    _lib.check(rc, "b200_wkv7_forward_seq")
    return y


def wkv7_forward_one(state, r, w, k, v, a, b, elapsed_t, slot_idx=None) -> torch.Tensor:
    """``RWKV7_ONE_OP`` / ``RWKV7_ONE_BATCH_OP`` (``Albatross/rwkv7.py:99-141``): inputs (B,C) or (C,)."""
    shp = r.shape
    C = shp[-1]
    H = C // HEAD
    B = 1 if r.dim() == 1 else shp[0]
    y = wkv7_forward_seq(state.view(-1, H, HEAD, HEAD), *(t.reshape(B, 1, C) for t in (r, w, k, v, a, b)),
                         elapsed_t.reshape(-1), slot_idx)
    return y.view(shp)


def spmv_forward(vec: torch.Tensor, mat: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``SPMV_OP`` / ``rwkv_mm_sparsity`` (``Albatross/rwkv7.py:61-87``): vec (D,), mat (D,C) -> (C,)."""
    _req(vec, "vec")
    _req(mat, "mat")
    D, C = mat.shape
    if out is None:
        out = torch.empty((C,), dtype=torch.float16, device=vec.device)
    rc = _lib.lib().b200_spmv_forward(D, C, vec.contiguous().data_ptr(), mat.contiguous().data_ptr(), out.data_ptr(),
                                      _stream(vec))
    _lib.check(rc, "b200_spmv_forward")
    return out


_ws = {}


def _workspace(dev: torch.device, M: int, N: int) -> torch.Tensor:
    need = _lib.lib().b200_linear_workspace_bytes(M, N)
    ws = _ws.get(dev)
    if ws is None or ws.numel() < need:
        ws = torch.zeros((max(need, 1 << 20),), dtype=torch.uint8, device=dev)
        _ws[dev] = ws
    return ws


def linear(x: torch.Tensor, W: torch.Tensor, bias: Optional[torch.Tensor] = None, epilogue: int = EPI_NONE,
           residual: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``F.linear(x, W, bias)`` with a fused epilogue: x (..., K) fp16, W (N, K) fp16 -> (..., N) fp16 on tcgen05."""
    _req(x, "x")
    _req(W, "W")
    K = x.shape[-1]
    N = W.shape[0]
    x2 = x.reshape(-1, K)
    if x2.stride(-1) != 1:
        x2 = x2.contiguous()
    if W.stride(-1) != 1:
        raise ValueError("W must be K-major (row-major (N,K))")
    M = x2.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=torch.float16, device=x.device)
    ws = _workspace(x.device, M, N)
    rc = _lib.lib().b200_linear_f16(M, N, K, x2.data_ptr(), x2.stride(0), W.data_ptr(), W.stride(0), out.data_ptr(),
                                    out.stride(0), bias.data_ptr() if bias is not None else 0,
                                    residual.data_ptr() if residual is not None else 0, epilogue, ws.data_ptr(),
                                    ws.numel(), _stream(x))
    _lib.check(rc, "b200_linear_f16")
    return out.view(*x.shape[:-1], N)


def argmax(logits: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Greedy sampling: (B, V) fp16 logits -> (B,) int32 token ids (lowest index on ties, like ``torch.argmax`` on fp32)."""
    _req(logits, "logits")
    B, V = logits.shape
    if out is None:
        out = torch.empty((B,), dtype=torch.int32, device=logits.device)
    rc = _lib.lib().b200_argmax_f16(B, V, logits.data_ptr(), logits.stride(0), out.data_ptr(), _stream(logits))
    _lib.check(rc, "b200_argmax_f16")
    return out


def tmix_core_fused(state, r, w, k, v, a, g, vgate, v_first, k_k, k_a, r_k, ln_w, ln_b, elapsed_t, slot_idx=None, out=None,
                    chunked=False):
    """The fused TMix core (``b200_tmix_core_fused``): inputs (B,T,C) fp16, per-channel vectors (C,), state (n_slots,H,64,64)
    updated in place; ``vgate`` / ``v_first`` None on layer 0.  Returns (B,T,C).  ``chunked=True``: the tensor-core chunked
    prefill form (``b200_tmix_core_chunked``), same arguments, tolerance-level agreement with the sequential kernel."""
    B, T, C = r.shape
    H = C // HEAD
    if out is None:
        out = torch.empty((B, T, C), dtype=torch.float16, device=r.device)
    for n, t in (("state", state), ("r", r), ("w", w), ("k", k), ("v", v), ("a", a), ("g", g)):
        _req(t, n)
    p = lambda t: 0 if t is None else t.contiguous().data_ptr()
    n_slots = state.shape[0]
    stride = state.stride(0) if n_slots > 1 else C * HEAD
    fn = _lib.lib().b200_tmix_core_chunked if chunked else _lib.lib().b200_tmix_core_fused
    rc = fn(B, T, C, H, state.data_ptr(), stride, n_slots, p(r), p(w), p(k), p(v), p(a), p(g),
            p(vgate), p(v_first), p(k_k), p(k_a), p(r_k), p(ln_w), p(ln_b), out.data_ptr(),
            elapsed_t.data_ptr(), p(slot_idx), _stream(r))
    _lib.check(rc, "b200_tmix_core_chunked" if chunked else "b200_tmix_core_fused")
    return out


def linear_group(xs, Ws, biases=None, epilogues=None, residuals=None, outs=None):
    """Several ``F.linear`` calls that share the number of rows M in ONE launch (``b200_linear_group_f16``): ``xs[i]`` (M, K_i),
    ``Ws[i]`` (N_i, K_i) -> list of (M, N_i).  Mirrors the r/k/v + LoRA-down group of one TMix layer (``Albatross/rwkv7.py:625-630``)."""
    n = len(xs)
    biases = biases or [None] * n
    epilogues = epilogues or [EPI_NONE] * n
    residuals = residuals or [None] * n
    M = xs[0].shape[0]
    descs = (_lib.LinearDesc * n)()
    keep = []
    res = []
    for i in range(n):
        x, W = xs[i], Ws[i]
        _req(x, "x")
        _req(W, "W")
        if x.dim() != 2 or x.shape[0] != M or x.stride(1) != 1 or W.stride(1) != 1:
            raise ValueError("linear_group: every x must be (M, K) with unit inner stride, every W K-major")
        out = outs[i] if outs is not None else torch.empty((M, W.shape[0]), dtype=torch.float16, device=x.device)
        res.append(out)
        keep.append((x, W, out))
        d = descs[i]
        d.N, d.K = W.shape[0], x.shape[1]
        d.A, d.lda, d.W, d.ldw, d.out, d.ldo = x.data_ptr(), x.stride(0), W.data_ptr(), W.stride(0), out.data_ptr(), out.stride(0)
        d.bias = biases[i].data_ptr() if biases[i] is not None else None
        d.residual = residuals[i].data_ptr() if residuals[i] is not None else None
        d.epilogue = epilogues[i]
    ws = _workspace(xs[0].device, M, max(W.shape[0] for W in Ws))
    rc = _lib.lib().b200_linear_group_f16(n, ctypes.cast(descs, ctypes.c_void_p), M, ws.data_ptr(), ws.numel(), _stream(xs[0]))
    _lib.check(rc, "b200_linear_group_f16")
    return res


def sample(logits: torch.Tensor, temperature: Optional[torch.Tensor] = None, top_p: Optional[torch.Tensor] = None,
           top_k: Optional[torch.Tensor] = None, *, slot: Optional[torch.Tensor] = None,
           occurrence: Optional[torch.Tensor] = None, presence: Optional[torch.Tensor] = None,
           penalty_decay: Optional[torch.Tensor] = None, frequency_penalty: Optional[torch.Tensor] = None,
           presence_penalty: Optional[torch.Tensor] = None, forbidden: Optional[torch.Tensor] = None,
           no_penalty: Optional[torch.Tensor] = None, seed: int = 0, step: int = 0, return_kept: bool = False,
           mode: int = 0):
    """One decode sampling step (``b200_sample_f16``): the worker's penalty block (``chirrup/worker.py:713-740``), its sampler
    (``chirrup/utils/samplers.py:98-168``) and the penalty update of the drawn token (``worker.py:527-547``) in one kernel.
    ``logits`` fp16 (rows, V) is penalised IN PLACE; per-slot parameter tensors are 1-D (n_slots) — fp32, ``top_k`` int32 —
    indexed by ``slot`` (int32 (rows), default: row index); ``occurrence`` / ``presence`` fp32 (n_slots, V) are updated in place;
    ``forbidden`` int32 (rows, F) padded with -1.  ``mode`` 0: the single-sort sampler (``samplers.py:98-168``); 1: the one the
    worker imports, ``sample_logits_rwkv_pip_compatible`` (``samplers.py:171-255``, ``worker.py:12``).  Returns int32 (rows)
    token ids (and the kept-set sizes if asked)."""
    _req(logits, "logits")
    if logits.dim() != 2 or logits.stride(1) != 1:
        raise ValueError("logits must be (rows, V) with unit inner stride")
    rows, V = logits.shape
    dev = logits.device
    P = _lib.SampleParams()

    def ptr(t, name, dtype, numel=None):
        if t is None:
            return None
        _req(t, name, dtype)
        if not t.is_contiguous():
            raise ValueError(f"{name} must be contiguous")
        if numel is not None and t.numel() < numel:
            raise ValueError(f"{name} too small")
        return t.data_ptr()

    P.slot = ptr(slot, "slot", torch.int32, rows)
    P.occurrence = ptr(occurrence, "occurrence", torch.float32)
    P.presence = ptr(presence, "presence", torch.float32)
    if occurrence is not None and (occurrence.dim() != 2 or occurrence.shape[1] != V or presence is None or presence.shape != occurrence.shape):
        raise ValueError("occurrence / presence must both be (n_slots, V)")
    P.penalty_decay = ptr(penalty_decay, "penalty_decay", torch.float32)
    P.frequency_penalty = ptr(frequency_penalty, "frequency_penalty", torch.float32)
    P.presence_penalty = ptr(presence_penalty, "presence_penalty", torch.float32)
    P.temperature = ptr(temperature, "temperature", torch.float32)
    P.top_p = ptr(top_p, "top_p", torch.float32)
    P.top_k = ptr(top_k, "top_k", torch.int32)
    P.forbidden = ptr(forbidden, "forbidden", torch.int32)
    P.max_forbidden = 0 if forbidden is None else (forbidden.shape[1] if forbidden.dim() == 2 and forbidden.shape[0] == rows else -1)
    P.no_penalty = ptr(no_penalty, "no_penalty", torch.int32)
    P.n_no_penalty = 0 if no_penalty is None else no_penalty.numel()
    P.seed, P.step = seed & (2 ** 64 - 1), step & (2 ** 64 - 1)
    P.mode = int(mode)
    kept = torch.empty((rows,), dtype=torch.int32, device=dev) if return_kept else None
    P.kept = kept.data_ptr() if kept is not None else None
    out = torch.empty((rows,), dtype=torch.int32, device=dev)
    rc = _lib.lib().b200_sample_f16(rows, V, logits.data_ptr(), logits.stride(0), ctypes.byref(P), out.data_ptr(), _stream(logits))
    _lib.check(rc, "b200_sample_f16")
    return (out, kept) if return_kept else out
