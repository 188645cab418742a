"""CPU tier: the C-ABI library builds/loads without a GPU and exports every symbol include/chirrup_b200.h declares."""
import ctypes
import os

import pytest

from chirrup_b200 import _lib


def test_library_exports_every_declared_symbol():
    names = _lib.header_symbols()
    assert "b200_wkv7_forward_seq" in names and "b200_linear_f16" in names
    L = _lib.lib()
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/chirrup_b200.h but not exported"
    assert set(names) == set(_lib.SIGNATURES), "ctypes signatures out of sync with the header"
    assert L.b200_abi_version() >= 1


def test_library_does_not_link_libcuda_or_torch():
    import subprocess

    out = subprocess.run(["ldd", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "libcuda.so" not in out and "torch" not in out and "c10" not in out


def test_ops_refuse_cpu_tensors():
    import torch

    from chirrup_b200 import ops

    z = torch.zeros((1, 1, 64), dtype=torch.float16)
    st = torch.zeros((1, 1, 64, 64), dtype=torch.float16)
    with pytest.raises(_lib.B200Error):
        ops.wkv7_forward_seq(st, z, z, z, z, z, z, torch.zeros(1, dtype=torch.int32))
    with pytest.raises(_lib.B200Error):
        ops.linear(torch.zeros((2, 64), dtype=torch.float16), torch.zeros((8, 64), dtype=torch.float16))


def test_bad_arguments_are_rejected_without_a_gpu():
    L = _lib.lib()
    # H*64 != C  (reference: assert(H*_N_ == C), cuda/rwkv7_state_fwd_fp16.cu:314)
    assert L.b200_wkv7_forward_seq(1, 1, 100, 1, 0, 0, 1, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0) == -1
    assert L.b200_linear_f16(4, 8, 7, 0, 8, 0, 8, 0, 8, 0, 0, 0, 0, 0, 0) == -1
    assert L.b200_linear_workspace_bytes(256, 4096) > 0
