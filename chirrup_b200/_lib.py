"""ctypes binding of ``libchirrup_b200.so`` (the C ABI declared in ``include/chirrup_b200.h``).

The library is the product: if it is missing it is built (nvcc) or the import fails loudly — there is no fallback path.
"""
from __future__ import annotations

import ctypes
import os
import re
from typing import Dict, List

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libchirrup_b200.so")
HEADER = os.path.join(os.path.dirname(HERE), "include", "chirrup_b200.h")

_lib = None

_c = ctypes
_P = _c.c_void_p
_I = _c.c_int
_L = _c.c_long

# name -> (restype, argtypes); must match include/chirrup_b200.h (tests/test_abi.py checks the symbol list against the header)
SIGNATURES: Dict[str, tuple] = {
    "b200_abi_version": (_I, []),
    "b200_wkv7_forward_seq": (_I, [_I, _I, _I, _I, _P, _L, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "b200_wkv7_forward_one": (_I, [_I, _I, _I, _P, _L, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "b200_tmix_core_fused": (_I, [_I, _I, _I, _I, _P, _L, _I] + [_P] * 17),
    "b200_tmix_core_chunked": (_I, [_I, _I, _I, _I, _P, _L, _I] + [_P] * 17),
    "b200_spmv_forward": (_I, [_I, _I, _P, _P, _P, _P]),
    "b200_linear_f16": (_I, [_I, _I, _I, _P, _L, _P, _L, _P, _L, _P, _P, _I, _P, _L, _P]),
    "b200_linear_workspace_bytes": (_L, [_I, _I]),
    "b200_linear_f16_static": (_I, [_I, _I, _I, _P, _L, _P, _L, _P, _L, _P, _P, _I, _P]),
    "b200_linear_static_scratch_bytes": (_L, []),
    "b200_linear_group_f16": (_I, [_I, _P, _I, _P, _L, _P]),
    "b200_layer_norm_f16": (_I, [_I, _I, _P, _P, _P, _P, _P]),
    "b200_model_create": (_I, [_I, _I, _I, _I, _I, _I, _I, _P, _P, _P]),
    "b200_model_destroy": (None, [_P]),
    "b200_model_set_ffn_value_rows": (_I, [_P, _P]),
    "b200_model_workspace_bytes": (_L, [_P, _I]),
    "b200_model_forward": (_I, [_P, _I, _I, _P, _P, _L, _L, _P, _L, _I, _P, _P, _P, _I, _P, _L, _P]),
    "b200_model_last_launch_count": (_I, [_P]),
    "b200_argmax_f16": (_I, [_I, _I, _P, _L, _P, _P]),
    "b200_sample_f16": (_I, [_I, _I, _P, _L, _P, _P, _P]),
    "b200_slot_copy": (_I, [_P, _L, _P, _L, _L, _L, _P]),
    "b200_slot_move": (_I, [_P, _L, _P, _L, _L, _L, _P, _L, _P, _L, _L, _L, _P, _P, _P]),
}


class LinearDesc(ctypes.Structure):
    """``b200_linear_desc`` of include/chirrup_b200.h."""
    _fields_ = [("N", _I), ("K", _I), ("A", _P), ("lda", _L), ("W", _P), ("ldw", _L), ("out", _P), ("ldo", _L),
                ("bias", _P), ("residual", _P), ("epilogue", _I)]


class SampleParams(ctypes.Structure):
    """``b200_sample_params`` of include/chirrup_b200.h."""
    _fields_ = [("slot", _P), ("occurrence", _P), ("presence", _P), ("penalty_decay", _P), ("frequency_penalty", _P),
                ("presence_penalty", _P), ("temperature", _P), ("top_p", _P), ("top_k", _P), ("forbidden", _P),
                ("max_forbidden", _I), ("no_penalty", _P), ("n_no_penalty", _I), ("seed", _c.c_ulonglong),
                ("step", _c.c_ulonglong), ("kept", _P), ("mode", _I)]


def header_symbols() -> List[str]:
    """Function names declared in include/chirrup_b200.h."""
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", src)))


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH) or not os.environ.get("CHIRRUP_B200_NO_BUILD"):
            from . import build as _build

            _build.build()  # no-op when the in-tree library is newer than its sources
        L = ctypes.CDLL(os.environ.get("CHIRRUP_B200_LIB") or LIB_PATH)  # override: A/B of two builds on one box
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the library does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


class B200Error(RuntimeError):
    pass


def check(rc: int, what: str) -> None:
    if rc != 0:
        kind = "cudaError" if rc > 0 else "b200 error"
        raise B200Error(f"{what} failed: {kind} {rc}")
