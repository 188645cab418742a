"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

Runs the reference's OWN model code (``Albatross/rwkv7.py``, unmodified) on the host CPU through the seam of SURVEY.md §8(c).
Used by ``tests/golden/make_golden.py`` (to pin the oracle) and by ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs
(to time the reference's CPU path: ``kind: "reference"``).  Nothing under ``chirrup_b200/`` imports it.

Where the reference comes from: ``/root/reference`` in the build container; on the GPU box (where that path does not exist)
the copy that ``oracle/stage_ref_model.sh`` staged, byte for byte, under the git-ignored ``oracle/_ref/`` together with the
pre-built torch extension the module JIT-compiles at import (``Albatross/rwkv7.py:47-58``).

The seam (everything else that runs is reference code):

* ``RWKV_x070(args, auto_load=False)`` — skips the hard-coded ``"cuda"`` of ``rwkv7.py:203``; ``torch.load`` is handed the
  in-memory synthetic checkpoint instead of a ``.pth`` file;
* the reference's own load-time transform (``load_weights_to_device``, ``:211-221``) applied on CPU, the post-load fixes of
  ``:206-209`` replayed, the whole dict assigned back (``m.z = zz``; a ScriptModule copies attribute dicts);
* hand-built CPU state tensors (``generate_zero_state`` hard-codes ``"cuda"``, ``:228-234``);
* ``oracle/wkv7_oracle.c`` registered as the CPU kernel of the reference's ``mylib::RWKV7_BATCH_OP`` (the reference has no
  CPU body for it, ``:144-160``).
"""
from __future__ import annotations

import os
import sys
import types
from typing import Dict, List, Optional

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
STAGED = os.path.join(HERE, "_ref")
# the spelling the extension was pre-built under (on the GPU box /root/repo is a symlink to the snapshot): ninja then has no work
_CANON = "/root/repo/oracle/_ref"
try:
    if os.path.isdir(_CANON) and os.path.samefile(_CANON, STAGED):
        STAGED = _CANON
except OSError:
    pass

_R = None


def ref_root() -> Optional[str]:
    for cand in (os.environ.get("REF_ROOT"), STAGED, "/root/reference"):
        if cand and os.path.exists(os.path.join(cand, "Albatross", "rwkv7.py")):
            return cand
    return None


def available() -> bool:
    return ref_root() is not None


def import_reference():
    """``import Albatross.rwkv7`` (JIT-builds / re-uses its torch extension; CPU box: the ``cuda/sm80`` fallback sources)."""
    global _R
    if _R is not None:
        return _R
    root = ref_root()
    if root is None:
        raise FileNotFoundError("reference model not found: neither /root/reference nor oracle/_ref/Albatross (run "
                                "`make -C oracle ref` where /root/reference exists)")
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0a")
    # the module picks its sources by whether a GPU is visible (rwkv7.py:40-58): one pre-built extension directory per variant
    ext_dir = os.path.join(STAGED, "torch_ext_gpu" if torch.cuda.is_available() else "torch_ext_cpu")
    try:
        os.makedirs(ext_dir, exist_ok=True)
        probe = os.path.join(ext_dir, ".w")
        open(probe, "w").close()
        os.remove(probe)
    except OSError:
        ext_dir = "/tmp/ref_torch_ext"
    os.environ.setdefault("TORCH_EXTENSIONS_DIR", ext_dir)
    if root not in sys.path:
        sys.path.insert(0, root)
    import Albatross.rwkv7 as R

    from . import rwkv7_oracle as O

    R.RWKV7_BATCH_OP.register_kernel("cpu")(
        lambda state, r, w, k, v, a, b, elapsed_t: O.wkv_cpu(state, r, w, k, v, a, b, elapsed_t))
    _R = R
    return R


def load_model(raw: Dict[str, torch.Tensor], vocab_size: int):
    """The reference ``RWKV_x070`` on CPU over the raw (pre-transform) checkpoint dict ``raw`` (consumed)."""
    R = import_reference()
    args = types.SimpleNamespace(vocab_size=vocab_size, head_size=64, MODEL_NAME="<in-memory synthetic checkpoint>")
    real_load = torch.load
    torch.load = lambda *a, **k: raw
    try:
        m = R.RWKV_x070(args, auto_load=False)
    finally:
        torch.load = real_load
    zz = dict(m.z)
    holder = types.SimpleNamespace(z=zz)
    R.RWKV_x070.load_weights_to_device(holder, list(zz.keys()), "cpu")
    zz = holder.z
    C = zz["emb.weight"].shape[1]
    zz["emb.weight"] = torch.nn.functional.layer_norm(zz["emb.weight"], (C,), weight=zz["blocks.0.ln0.weight"],
                                                      bias=zz["blocks.0.ln0.bias"])
    for s in ("0", "1", "2"):
        zz[f"blocks.0.att.v{s}"] = zz[f"blocks.0.att.a{s}"]
    m.z = zz
    return R, m


def zero_state(m, bsz: int) -> List[torch.Tensor]:
    L, C = m.n_layer, m.n_embd
    return [torch.zeros((L, 2, bsz, C), dtype=torch.half), torch.zeros((L, bsz, C // 64, 64, 64), dtype=torch.half),
            torch.zeros((bsz,), dtype=torch.int32)]
