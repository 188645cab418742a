"""Registers the reference's operator names on top of the C ABI, so that the unmodified ``Albatross/rwkv7.py`` can run on the
B200 kernels (kernel-level A/B inside the reference model): ``torch.ops.rwkv7_state_fwd_fp16.{forward_seq, forward_one,
spmv_forward}`` with the signatures of ``/root/reference/Albatross/cuda/rwkv7_state_fwd_fp16.cpp:10-25``.

Import this module INSTEAD of letting rwkv7.py JIT-build its extension."""
from __future__ import annotations

import torch

from . import _lib

_registered = False


def register() -> None:
    global _registered
    if _registered:
        return
    lib = torch.library.Library("rwkv7_state_fwd_fp16", "DEF")
    lib.define("forward_seq(int B, int T, int C, int H, Tensor(a!) state, Tensor r, Tensor w, Tensor k, Tensor v, Tensor a, "
               "Tensor b, Tensor(b!) y, Tensor elapsed_t) -> ()")
    lib.define("forward_one(int B, int C, int H, Tensor(a!) state, Tensor r, Tensor w, Tensor k, Tensor v, Tensor a, "
               "Tensor b, Tensor(b!) y, Tensor elapsed_t) -> ()")
    lib.define("spmv_forward(int D, int C, Tensor vec, Tensor mat, Tensor(a!) out) -> ()")

    def _stream(t):
        return torch.cuda.current_stream(t.device).cuda_stream

    def forward_seq(B, T, C, H, state, r, w, k, v, a, b, y, elapsed_t):
        rc = _lib.lib().b200_wkv7_forward_seq(B, T, C, H, state.data_ptr(), C * 64, B, r.data_ptr(), w.data_ptr(), k.data_ptr(),
                                              v.data_ptr(), a.data_ptr(), b.data_ptr(), y.data_ptr(), elapsed_t.data_ptr(), 0,
                                              _stream(r))
        _lib.check(rc, "b200_wkv7_forward_seq")

    def forward_one(B, C, H, state, r, w, k, v, a, b, y, elapsed_t):
        forward_seq(B, 1, C, H, state, r, w, k, v, a, b, y, elapsed_t)

    def spmv_forward(D, C, vec, mat, out):
        # the reference pre-zeroes `out` and accumulates with atomics (rwkv7.py:65-66); ours overwrites, same result
        rc = _lib.lib().b200_spmv_forward(D, C, vec.data_ptr(), mat.data_ptr(), out.data_ptr(), _stream(vec))
        _lib.check(rc, "b200_spmv_forward")

    lib.impl("forward_seq", forward_seq, "CUDA")
    lib.impl("forward_one", forward_one, "CUDA")
    lib.impl("spmv_forward", spmv_forward, "CUDA")
    register._lib = lib  # keep alive
    _registered = True
