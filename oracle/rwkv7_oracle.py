"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

CPU restatement ("oracle") of the batched RWKV-7 forward step that Chirrup drives through Albatross.
Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may
import this module; nothing under ``chirrup_b200/`` does.

What it restates (reference file:line, all under ``/root/reference``):

* checkpoint -> ``z`` transform            ``Albatross/rwkv7.py:171-221`` (+ ``:206-209`` ln0 fold, layer-0 v aliases)
* state factory                            ``Albatross/rwkv7.py:224-235``
* ``forward_seq_batch[_seperate]``         ``Albatross/rwkv7.py:351-382``, ``:503-563``
* ``RWKV_x070_TMix_seq_batch``             ``Albatross/rwkv7.py:618-649``
* ``RWKV_x070_CMix_seq_batch``             ``Albatross/rwkv7.py:673-679``
* bsz-1 ``forward_one`` / ``forward_seq``  ``Albatross/rwkv7.py:287-349``, ``:568-616``, ``:653-671``
* ``forward_batch`` (ragged)               ``Albatross/rwkv7.py:250-285``
* the WKV op body                          ``Albatross/cuda/rwkv7_state_fwd_fp16.cu:26-97`` -> ``oracle/wkv7_oracle.c``

Every float op is the same ATen op, on the same fp16 tensors and in the same order as the reference, so on CPU the
result is bit-identical to the unmodified reference model run through the CPU seam of SURVEY.md §8(c) — that is how
the oracle is pinned (``tests/golden/make_golden.py`` ran the reference here and committed the vectors; the reference
itself holds no golden vectors for this path).  The oracle is device-agnostic: on the GPU box tests may also run it
with ``device="cuda"`` and ``wkv="ref_cuda"`` — torch/cuBLAS ops plus the reference's own CUDA kernel from
``oracle/_ref`` — which is the closest thing to "the unmodified reference on the same B200" that can travel.
"""
from __future__ import annotations

import ctypes
import os
from typing import Callable, Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
HEAD = 64
DTYPE = torch.half

_T_KEYS = ("att.g1", "att.g2", "att.a1", "att.a2", "att.w1", "att.w2", "att.v1", "att.v2", "ffn.value.weight")

# --------------------------------------------------------------------------------------------------
# WKV back-ends
# --------------------------------------------------------------------------------------------------
_clib = None


def _c_oracle():
    global _clib
    if _clib is None:
        path = os.path.join(HERE, "liboracle_wkv7.so")
        if not os.path.exists(path):
            import subprocess

            subprocess.check_call(["make", "-C", HERE, "liboracle_wkv7.so"])
        lib = ctypes.CDLL(path)
        p = ctypes.c_void_p
        lib.wkv7_oracle_forward_seq.argtypes = [ctypes.c_int] * 4 + [p, ctypes.c_long] + [p] * 8
        lib.wkv7_oracle_forward_seq.restype = None
        lib.wkv7_oracle_spmv.argtypes = [ctypes.c_int, ctypes.c_int, p, p, p]
        lib.wkv7_oracle_spmv.restype = None
        lib.wkv7_oracle_decay.argtypes = [ctypes.c_int, p, ctypes.c_int, p]
        lib.wkv7_oracle_decay.restype = None
        _clib = lib
    return _clib


def wkv_cpu(state: torch.Tensor, r, w, k, v, a, b, elapsed_t: torch.Tensor) -> torch.Tensor:
    """(B,T,C) fp16 inputs, state (B,H,64,64) fp16 updated in place (may be a B-axis view); returns y (B,T,C).

    Restates ``forward_seq`` (``Albatross/cuda/rwkv7_state_fwd_fp16.cu:313-316`` -> ``:26-97``)."""
    B, T, C = r.shape
    H = C // HEAD
    assert state.dtype == DTYPE and state.device.type == "cpu"
    assert state.shape[-3:] == (H, HEAD, HEAD) and state[0].is_contiguous()
    stride_b = state.stride(0) if state.dim() == 4 else C * HEAD
    ins = [t.contiguous() for t in (r, w, k, v, a, b)]
    y = torch.empty((B, T, C), dtype=DTYPE)
    et = elapsed_t.to(torch.int32).reshape(-1).contiguous()
    assert et.numel() == B
    _c_oracle().wkv7_oracle_forward_seq(
        B, T, C, H, state.data_ptr(), stride_b, *[t.data_ptr() for t in ins], y.data_ptr(), et.data_ptr()
    )
    return y


def spmv_cpu(vec: torch.Tensor, mat: torch.Tensor) -> torch.Tensor:
    """``rwkv_mm_sparsity(k, V_)`` = ``SPMV_OP`` (``Albatross/rwkv7.py:61-76``, ``.cu:222-310``): exact product, one rounding."""
    D, C = mat.shape
    out = torch.empty((C,), dtype=DTYPE)
    vec = vec.contiguous()
    mat = mat.contiguous()
    _c_oracle().wkv7_oracle_spmv(D, C, vec.data_ptr(), mat.data_ptr(), out.data_ptr())
    return out


_reflib = None


def ref_cuda_lib():
    """The reference's own CUDA kernels built for sm_100a by oracle/build_ref.sh (GPU box only)."""
    global _reflib
    if _reflib is None:
        path = os.path.join(HERE, "_ref", "libref_wkv7.so")
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing — run `make -C oracle ref` where /root/reference exists")
        lib = ctypes.CDLL(path)
        p = ctypes.c_void_p
        seq = getattr(lib, "_Z16cuda_forward_seqiiiiPN3c104HalfES1_S1_S1_S1_S1_S1_S1_Pi")
        seq.argtypes = [ctypes.c_int] * 4 + [p] * 9
        seq.restype = None
        one = getattr(lib, "_Z16cuda_forward_oneiiiPN3c104HalfES1_S1_S1_S1_S1_S1_S1_Pi")
        one.argtypes = [ctypes.c_int] * 3 + [p] * 9
        one.restype = None
        spmv = getattr(lib, "_Z17cuda_spmv_forwardiiPN3c104HalfES1_S1_")
        spmv.argtypes = [ctypes.c_int] * 2 + [p] * 3
        spmv.restype = None
        lib.ref_seq, lib.ref_one, lib.ref_spmv = seq, one, spmv
        _reflib = lib
    return _reflib


def wkv_ref_cuda(state, r, w, k, v, a, b, elapsed_t) -> torch.Tensor:
    """Calls the reference's ``cuda_forward_seq`` launcher (legacy default stream, ``.cu:313-316``)."""
    B, T, C = r.shape
    H = C // HEAD
    assert state.is_cuda and state.is_contiguous() and state.dtype == DTYPE
    ins = [t.contiguous() for t in (r, w, k, v, a, b)]
    y = torch.empty((B, T, C), dtype=DTYPE, device=r.device)
    et = elapsed_t.to(torch.int32).reshape(-1).contiguous()
    torch.cuda.synchronize()
    ref_cuda_lib().ref_seq(B, T, C, H, state.data_ptr(), *[t.data_ptr() for t in ins], y.data_ptr(), et.data_ptr())
    torch.cuda.synchronize()
    return y


def wkv_one_ref_cuda(state, r, w, k, v, a, b, elapsed_t) -> torch.Tensor:
    """Calls the reference's ``cuda_forward_one`` launcher (``.cu:318-322``); inputs (B,C)."""
    B, C = r.shape
    H = C // HEAD
    ins = [t.contiguous() for t in (r, w, k, v, a, b)]
    y = torch.empty((B, C), dtype=DTYPE, device=r.device)
    et = elapsed_t.to(torch.int32).reshape(-1).contiguous()
    torch.cuda.synchronize()
    ref_cuda_lib().ref_one(B, C, H, state.data_ptr(), *[t.data_ptr() for t in ins], y.data_ptr(), et.data_ptr())
    torch.cuda.synchronize()
    return y


def spmv_ref_cuda(vec, mat) -> torch.Tensor:
    D, C = mat.shape
    out = torch.zeros((C,), dtype=DTYPE, device=vec.device)
    torch.cuda.synchronize()
    ref_cuda_lib().ref_spmv(D, C, vec.contiguous().data_ptr(), mat.contiguous().data_ptr(), out.data_ptr())
    torch.cuda.synchronize()
    return out


def _wkv_backend(name: str) -> Callable:
    return {"cpu": wkv_cpu, "ref_cuda": wkv_ref_cuda}[name]


# --------------------------------------------------------------------------------------------------
# weights + state
# --------------------------------------------------------------------------------------------------
def transform_checkpoint(raw: Dict[str, torch.Tensor], device="cpu") -> Dict[str, torch.Tensor]:
    """``RWKV_x070.load_weights_to_device`` + the post-load fixes (``Albatross/rwkv7.py:211-221``, ``:206-209``)."""
    z = {}
    for key, t in raw.items():
        if any(s in key for s in _T_KEYS):
            t = t.t()
        t = t.squeeze().to(dtype=DTYPE, device=device)
        if key.endswith("att.r_k"):
            t = t.flatten()
        z[key] = t.contiguous()
    C = z["emb.weight"].shape[1]
    z["emb.weight"] = F.layer_norm(z["emb.weight"], (C,), weight=z["blocks.0.ln0.weight"], bias=z["blocks.0.ln0.bias"])
    for s in ("0", "1", "2"):
        z[f"blocks.0.att.v{s}"] = z[f"blocks.0.att.a{s}"]
    return z


class Geometry:
    def __init__(self, z: Dict[str, torch.Tensor], raw_r_k_shape=None):
        self.n_layer = 1 + max(int(k.split(".")[1]) for k in z if k.startswith("blocks."))
        self.n_embd = z["emb.weight"].shape[1]
        self.head_size = HEAD
        self.n_head = self.n_embd // HEAD
        self.vocab_size = z["head.weight"].shape[0]


def zero_state(geom: Geometry, bsz: int, device="cpu") -> List[torch.Tensor]:
    """``generate_zero_state`` (``Albatross/rwkv7.py:224-235``); bsz == 0 gives the unbatched layout."""
    L, C, H = geom.n_layer, geom.n_embd, geom.n_head
    if bsz >= 1:
        return [
            torch.zeros((L, 2, bsz, C), dtype=DTYPE, device=device),
            torch.zeros((L, bsz, H, HEAD, HEAD), dtype=DTYPE, device=device),
            torch.zeros((bsz,), dtype=torch.int32, device=device),
        ]
    return [
        torch.zeros((L, 2, C), dtype=DTYPE, device=device),
        torch.zeros((L, H, HEAD, HEAD), dtype=DTYPE, device=device),
        torch.zeros((), dtype=torch.int32, device=device),
    ]


# --------------------------------------------------------------------------------------------------
# the step
# --------------------------------------------------------------------------------------------------
def _linear(x, W, bias=None, fast=False):
    if not fast:
        return F.linear(x, W, bias)
    # Same contract (fp32 accumulate, one rounding to fp16) but through sgemm: for the big CPU cases.
    y = x.float() @ W.float().t()
    if bias is not None:
        y = y + bias.float()
    return y.to(DTYPE)


def tmix_seq_batch(layer_id, H, x, x_prev, v_first, wkv_state, p, elapsed_t, wkv, fast=False, taps=None):
    """``RWKV_x070_TMix_seq_batch`` (``Albatross/rwkv7.py:618-649``). ``p`` maps short names to the layer's tensors."""
    B, T, C = x.shape
    N = HEAD
    lin = lambda a, W, bias=None: _linear(a, W, bias, fast)
    xx = torch.cat((x_prev[0].unsqueeze(1), x[:, :-1, :]), dim=1) - x           # :621
    x_prev[0] = x[:, -1, :]                                                       # :622
    xr, xw, xk, xv, xa, xg = (x + xx * p[n] for n in ("x_r", "x_w", "x_k", "x_v", "x_a", "x_g"))   # :623
    r = lin(xr, p["R"])                                                           # :625
    w = lin(torch.tanh(lin(xw, p["w1"])), p["w2"], p["w0"])                       # :626
    k = lin(xk, p["K"])                                                           # :627
    v = lin(xv, p["V"])                                                           # :628
    a = torch.sigmoid(lin(lin(xa, p["a1"]), p["a2"], p["a0"]))                    # :629
    g = lin(torch.sigmoid(lin(xg, p["g1"])), p["g2"])                             # :630
    kk = F.normalize((k * p["k_k"]).view(B, T, H, N), dim=-1, p=2.0).view(B, T, H * N)   # :632
    k = k * (1 + (a - 1) * p["k_a"])                                              # :633
    kka = kk * a                                                                  # :634
    if layer_id == 0:
        v_first = v                                                               # :636
    else:
        v = v + (v_first - v) * torch.sigmoid(lin(lin(xv, p["v1"]), p["v2"], p["v0"]))   # :637
    if taps is not None:
        taps.append(dict(r=r.clone(), w=w.clone(), k=k.clone(), v=v.clone(), a=(-kk).clone(), b=kka.clone(), g=g.clone()))
    y = wkv(wkv_state, r, w, k, v, -kk, kka, elapsed_t).view(B * T, H * N)        # :645
    y = F.group_norm(y, num_groups=H, weight=p["ln_x_w"], bias=p["ln_x_b"], eps=64e-5).view(B, T, H * N)   # :647
    y = y + ((r * k * p["r_k"]).view(B, T, H, N).sum(dim=-1, keepdim=True) * v.view(B, T, H, N)).view(B, T, H * N)  # :648
    return lin(y * g, p["O"]), v_first                                            # :649


def cmix_seq_batch(x, x_prev, x_k, K_, V_, fast=False):
    """``RWKV_x070_CMix_seq_batch`` (``Albatross/rwkv7.py:673-679``)."""
    xx = torch.cat((x_prev[1].unsqueeze(1), x[:, :-1, :]), dim=1) - x
    x_prev[1] = x[:, -1, :]
    k = x + xx * x_k
    k = torch.relu(_linear(k, K_, None, fast)) ** 2
    if fast:
        return (k.float() @ V_.float()).to(DTYPE)
    return k @ V_


def layer_params(z, i):
    att, ffn = f"blocks.{i}.att.", f"blocks.{i}.ffn."
    p = {n: z[att + n] for n in ("x_r", "x_w", "x_k", "x_v", "x_a", "x_g", "w0", "w1", "w2", "a0", "a1", "a2",
                                 "v0", "v1", "v2", "g1", "g2", "k_k", "k_a", "r_k")}
    p.update(R=z[att + "receptance.weight"], K=z[att + "key.weight"], V=z[att + "value.weight"],
             O=z[att + "output.weight"], ln_x_w=z[att + "ln_x.weight"], ln_x_b=z[att + "ln_x.bias"],
             ffn_x_k=z[ffn + "x_k"], ffn_K=z[ffn + "key.weight"], ffn_V=z[ffn + "value.weight"],
             ln1_w=z[f"blocks.{i}.ln1.weight"], ln1_b=z[f"blocks.{i}.ln1.bias"],
             ln2_w=z[f"blocks.{i}.ln2.weight"], ln2_b=z[f"blocks.{i}.ln2.bias"])
    return p


def forward_seq_batch(z, geom: Geometry, idxs: Sequence[Sequence[int]], state: List[torch.Tensor],
                      full_output: bool = False, wkv: str = "cpu", fast: bool = False,
                      layer_range: Optional[tuple] = None, taps: Optional[list] = None,
                      skip_head: bool = False) -> torch.Tensor:
    """``forward_seq_batch_seperate`` (``Albatross/rwkv7.py:555-563`` -> ``:503-553``); state mutated in place."""
    wkv_fn = _wkv_backend(wkv)
    C, H = geom.n_embd, geom.n_head
    dev = z["emb.weight"].device
    x = z["emb.weight"][torch.tensor(idxs, device=dev)]                           # :507
    v_first = torch.empty_like(x)
    lo, hi = layer_range if layer_range is not None else (0, geom.n_layer)
    for i in range(lo, hi):
        p = layer_params(z, i)
        xx = F.layer_norm(x, (C,), weight=p["ln1_w"], bias=p["ln1_b"])            # :523
        xx, v_first = tmix_seq_batch(i, H, xx, state[0][i], v_first, state[1][i], p, state[2], wkv_fn, fast, taps)
        x = x + xx                                                                # :531
        xx = F.layer_norm(x, (C,), weight=p["ln2_w"], bias=p["ln2_b"])            # :533
        xx = cmix_seq_batch(xx, state[0][i], p["ffn_x_k"], p["ffn_K"], p["ffn_V"], fast)
        x = x + xx                                                                # :536
    if not full_output:
        x = x[:, -1, :]                                                           # :547
    if skip_head:
        state[2] += len(idxs[0])
        return x
    x = F.layer_norm(x, (C,), weight=z["ln_out.weight"], bias=z["ln_out.bias"])  # :548
    x = _linear(x, z["head.weight"], None, fast)                                  # :551
    state[2] += len(idxs[0])                                                      # :552
    return x


def forward_batch(z, geom, tokens: List[List[int]], state, full_output=False, wkv="cpu", fast=False):
    """``forward_batch`` for ragged lengths (``Albatross/rwkv7.py:250-280``): common-prefix stepping with
    gather/scatter of state rows."""
    lengths = [len(t) for t in tokens]
    if len(set(lengths)) == 1 and not full_output:
        return forward_seq_batch(z, geom, tokens, state, full_output, wkv, fast)
    bsz = len(tokens)
    pos = [0] * bsz
    dev = z["emb.weight"].device
    if not full_output:
        out = torch.empty((bsz, geom.vocab_size), dtype=DTYPE, device=dev)
    else:
        out = [torch.empty((0, geom.vocab_size), dtype=DTYPE, device=dev) for _ in range(bsz)]
    while True:
        active = [i for i in range(bsz) if pos[i] < lengths[i]]
        if not active:
            break
        step = min(lengths[i] - pos[i] for i in active)
        chunk = [tokens[i][pos[i]:pos[i] + step] for i in active]
        sub = [state[0][:, :, active], state[1][:, active], state[2][active]]
        new = forward_seq_batch(z, geom, chunk, sub, full_output, wkv, fast)
        for k_, i in enumerate(active):
            if not full_output:
                out[i] = new[k_]
            else:
                out[i] = torch.cat([out[i], new[k_]], dim=0)
            state[0][:, :, i] = sub[0][:, :, k_]
            state[1][:, i] = sub[1][:, k_]
            state[2][i] = sub[2][k_]
            pos[i] += step
    return out


# ---- bsz-1 path ------------------------------------------------------------------------------------
def forward_one(z, geom, token_or_x, state, wkv: str = "cpu", spmv: Optional[Callable] = None, fast: bool = False):
    """``forward`` -> ``forward_one`` with ``RWKV_x070_TMix_one`` / ``RWKV_x070_CMix_one``
    (``Albatross/rwkv7.py:237-248``, ``:287-316``, ``:568-590``, ``:653-662``). Unbatched state layout."""
    C, H, N = geom.n_embd, geom.n_head, HEAD
    wkv_fn = _wkv_backend(wkv)
    spmv = spmv or (spmv_cpu if z["emb.weight"].device.type == "cpu" else spmv_ref_cuda)
    if fast:   # same contract (fp32 accumulate, one rounding) through sgemm: measures the reference's GEMM-order noise (tests)
        class _L:
            linear = staticmethod(lambda a, W, bias=None: _linear(a, W, bias, True))
            layer_norm, group_norm, normalize = F.layer_norm, F.group_norm, F.normalize
        return _forward_one_impl(z, geom, token_or_x, state, wkv_fn, spmv, _L)
    return _forward_one_impl(z, geom, token_or_x, state, wkv_fn, spmv, F)


def _forward_one_impl(z, geom, token_or_x, state, wkv_fn, spmv, F):
    C, H, N = geom.n_embd, geom.n_head, HEAD
    x = z["emb.weight"][token_or_x] if not torch.is_tensor(token_or_x) else token_or_x
    v_first = torch.empty_like(x)
    for i in range(geom.n_layer):
        p = layer_params(z, i)
        xx = F.layer_norm(x, (C,), weight=p["ln1_w"], bias=p["ln1_b"])
        x_prev = state[0][i]
        dx = x_prev[0] - xx                                                       # :569
        x_prev[0] = xx
        xr, xw, xk, xv, xa, xg = (xx + dx * p[n] for n in ("x_r", "x_w", "x_k", "x_v", "x_a", "x_g"))
        r = F.linear(xr, p["R"])
        w = F.linear(torch.tanh(F.linear(xw, p["w1"])), p["w2"], bias=p["w0"])
        k = F.linear(xk, p["K"])
        v = F.linear(xv, p["V"])
        a = torch.sigmoid(F.linear(F.linear(xa, p["a1"]), p["a2"], bias=p["a0"]))
        g = F.linear(torch.sigmoid(F.linear(xg, p["g1"])), p["g2"])
        kk = F.normalize((k * p["k_k"]).view(H, N), dim=-1, p=2.0).view(H * N)
        k = k * (1 + (a - 1) * p["k_a"])
        kka = kk * a
        if i == 0:
            v_first = v
        else:
            v = v + (v_first - v) * torch.sigmoid(F.linear(F.linear(xv, p["v1"]), p["v2"], bias=p["v0"]))
        s = state[1][i]
        y = wkv_fn(s.view(1, H, N, N), *(t.view(1, 1, C) for t in (r, w, k, v, -kk, kka)), state[2].view(1)).view(C)
        y = F.group_norm(y.view(1, C), num_groups=H, weight=p["ln_x_w"], bias=p["ln_x_b"], eps=64e-5).view(C)
        y = y + ((r * k * p["r_k"]).view(H, N).sum(dim=-1, keepdim=True) * v.view(H, N)).view(C)
        x = x + F.linear(y * g, p["O"])
        xx = F.layer_norm(x, (C,), weight=p["ln2_w"], bias=p["ln2_b"])
        dx = x_prev[1] - xx                                                       # :655
        x_prev[1] = xx
        kf = xx + dx * p["ffn_x_k"]
        kf = torch.relu(F.linear(kf, p["ffn_K"])) ** 2
        x = x + spmv(kf, p["ffn_V"])                                              # :659
    x = F.layer_norm(x, (C,), weight=z["ln_out.weight"], bias=z["ln_out.bias"])
    x = F.linear(x, z["head.weight"])
    state[2] += 1
    return x


def greedy_decode(z, geom, prompt: Sequence[Sequence[int]], n_new: int, state=None, wkv="cpu", fast=False):
    """Prefill + greedy decode; returns (tokens (B, n_new) int64, last logits, state)."""
    B = len(prompt)
    dev = z["emb.weight"].device
    state = state or zero_state(geom, B, dev)
    logits = forward_seq_batch(z, geom, prompt, state, wkv=wkv, fast=fast)
    out = []
    for _ in range(n_new):
        nxt = logits.float().argmax(dim=-1)
        out.append(nxt.cpu())
        logits = forward_seq_batch(z, geom, [[int(t)] for t in nxt], state, wkv=wkv, fast=fast)
    return torch.stack(out, dim=1), logits, state
