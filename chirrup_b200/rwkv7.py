"""Host-side mirror of Albatross's model/state API over the C-ABI library.

Drop-in for ``from Albatross.rwkv7 import RWKV_x070`` (``/root/reference/Albatross/rwkv7.py:164-563``): same constructor
arguments, attributes, state layout (``generate_zero_state``, ``:224-235``) and ``forward*`` family with the same argument
meaning, in-place state mutation and fp16 logits.  All arithmetic runs in ``libchirrup_b200.so``; torch only owns device
memory and the stream.  There is no CPU path: constructing the model without a CUDA device raises.
"""
from __future__ import annotations

import ctypes
import types
import os
from collections import OrderedDict
from typing import Dict, List, Optional, Sequence

import torch

from . import _lib

HEAD_SIZE = 64
DTYPE = torch.half

# keys transposed at load by the reference (rwkv7.py:217-218).  ffn.value.weight is NOT transposed here: the checkpoint
# layout (C, 4C) is already the K-major (N, K) layout the tensor-core kernel wants; z[...] exposes the reference view.
_T_KEYS = ("att.g1", "att.g2", "att.a1", "att.a2", "att.w1", "att.w2", "att.v1", "att.v2")

_LW_ORDER = [
    "ln1.weight", "ln1.bias", "ln2.weight", "ln2.bias",
    "att.x_r", "att.x_w", "att.x_k", "att.x_v", "att.x_a", "att.x_g",
    "att.w0", "att.w1", "att.w2", "att.a0", "att.a1", "att.a2", "att.v0", "att.v1", "att.v2", "att.g1", "att.g2",
    "att.k_k", "att.k_a", "att.r_k",
    "att.receptance.weight", "att.key.weight", "att.value.weight", "att.output.weight",
    "att.ln_x.weight", "att.ln_x.bias",
    "ffn.x_k", "ffn.key.weight", "ffn.value.weight",
]


_STEP_ENV = ("B200_UNFUSED_TMIX", "B200_NO_PDL", "B200_NO_PAIR", "B200_NO_DEFER", "B200_WKV_GENERIC", "B200_CHAIN", "B200_WKV_SEQUENTIAL", "B200_NO_SPMV")


def _stream(dev) -> int:
    return torch.cuda.current_stream(dev).cuda_stream


class RWKV_x070:
    """B200-native RWKV-7 ("x070") inference model with Albatross's interface."""

    def __init__(self, args, auto_load: bool = True, state_dict: Optional[Dict[str, torch.Tensor]] = None,
                 device: Optional[torch.device] = None):
        if not torch.cuda.is_available():
            raise _lib.B200Error("chirrup_b200.RWKV_x070 needs a CUDA device (sm_100a); there is no CPU fallback")
        self.args = args
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self._handle = None
        self._ws = None
        self._ws_tokens = 0
        # Decode steps (T = 1) are replayed from CUDA graphs keyed by every address and shape of the call: the same ~260 launches,
        # without the per-launch driver latency (eager 6.8 ms vs replay 6.5 ms per 7.2B step at bsz 256).  A key is captured once it
        # has been seen `graph_after` times (a batch composition that lasts: a capture costs a few ms); least-recently-used
        # graphs beyond `graph_cache_size` are dropped.
        self.graph_decode = os.environ.get("B200_NO_STEP_GRAPH") is None and os.environ.get("B200_DEBUG_SYNC") is None
        self.graph_cache_size = 8
        self.graph_after = 12
        self._graphs: "OrderedDict[tuple, torch.cuda.CUDAGraph]" = OrderedDict()
        self._graph_seen: Dict[tuple, int] = {}
        self._capture_stream = None
        self._gio = None
        self._ffn_rows = None
        self.graph_hits = self.graph_misses = 0
        raw = state_dict if state_dict is not None else torch.load(args.MODEL_NAME + ".pth", map_location="cpu")
        self.n_head, self.head_size = raw["blocks.0.att.r_k"].shape          # rwkv7.py:173
        assert self.head_size == HEAD_SIZE
        args.head_size = self.head_size
        args.vocab_size, args.n_embd = raw["emb.weight"].shape                # rwkv7.py:174-175
        args.n_layer = 1 + max(int(k.split(".")[1]) for k in raw if k.startswith("blocks."))
        self.n_layer, self.n_embd, self.vocab_size = args.n_layer, args.n_embd, args.vocab_size
        self._raw = raw
        self.z: Dict[str, torch.Tensor] = {}
        if auto_load:
            self.load_weights()

    # ---------------------------------------------------------------- weights
    def load_weights(self) -> None:
        """``load_weights_to_device`` + the post-load fixes (rwkv7.py:199-221)."""
        L = _lib.lib()
        dev, C = self.device, self.n_embd
        w: Dict[str, torch.Tensor] = {}
        for key, t in self._raw.items():
            if any(s in key for s in _T_KEYS):
                t = t.t()
            t = t.squeeze().to(device=dev, dtype=DTYPE)
            if key.endswith("att.r_k"):
                t = t.flatten()
            w[key] = t.contiguous()
        self._raw = None
        with torch.cuda.device(dev):
            # emb <- LayerNorm(emb, ln0)  (rwkv7.py:206), on our own kernel, in place, in row blocks
            emb = w["emb.weight"]
            rc = L.b200_layer_norm_f16(emb.shape[0], C, emb.data_ptr(), w["blocks.0.ln0.weight"].data_ptr(),
                                       w["blocks.0.ln0.bias"].data_ptr(), emb.data_ptr(), _stream(dev))
            _lib.check(rc, "b200_layer_norm_f16")
        dims = {n: w[f"blocks.0.att.{n}1"].shape[0] for n in ("w", "a", "v", "g")}   # LoRA ranks, before the aliasing below
        for s in ("0", "1", "2"):                                             # rwkv7.py:207-209
            w[f"blocks.0.att.v{s}"] = w[f"blocks.0.att.a{s}"]
        self._w = w
        # reference-layout view of the dict (ffn.value.weight as (4C, C))
        self.z = {k: (v.t() if k.endswith("ffn.value.weight") else v) for k, v in w.items()}
        g = [w["emb.weight"], w["ln_out.weight"], w["ln_out.bias"], w["head.weight"]]
        lw = [w[f"blocks.{i}.{name}"] for i in range(self.n_layer) for name in _LW_ORDER]
        self._keep = g + lw
        GP = (ctypes.c_void_p * 4)(*[t.data_ptr() for t in g])
        LP = (ctypes.c_void_p * len(lw))(*[t.data_ptr() for t in lw])
        d = lambda n: dims[n]
        h = ctypes.c_void_p()
        rc = L.b200_model_create(self.n_layer, C, self.vocab_size, d("w"), d("a"), d("v"), d("g"), GP, LP, ctypes.byref(h))
        _lib.check(rc, "b200_model_create")
        self._handle = h
        torch.cuda.synchronize(dev)

    def __del__(self):
        try:
            if getattr(self, "_handle", None):
                _lib.lib().b200_model_destroy(self._handle)
                self._handle = None
        except Exception:
            pass

    def _workspace(self, n_tokens: int) -> torch.Tensor:
        if self._ws is None or n_tokens > self._ws_tokens:
            cap = max(n_tokens, 64)
            need = _lib.lib().b200_model_workspace_bytes(self._handle, cap)
            # captured step graphs hold the OLD workspace address: they die with it
            self._graphs.clear()
            self._graph_seen.clear()
            self._ws = None
            self._ws = torch.zeros((need,), dtype=torch.uint8, device=self.device)
            self._ws_tokens = cap
        return self._ws

    def _graph_io(self, B: int):
        """Model-owned staging buffers of the graphed decode step (token ids, slot indices, logits): a captured graph then
        depends on the caller's STATE tensors only, not on where this call's token / slot-index / logits tensors happen to
        live (a fresh ``torch.tensor(list)`` every step, a new index tensor whenever rows join or leave the batch)."""
        if self._gio is None or self._gio[0].numel() < B:
            cap = max(B, 64)
            self._graphs.clear()
            self._graph_seen.clear()
            self._gio = (torch.zeros((cap,), dtype=torch.int32, device=self.device),
                         torch.zeros((cap,), dtype=torch.int32, device=self.device),
                         torch.empty((cap, self.vocab_size), dtype=DTYPE, device=self.device))
        return self._gio

    # ---------------------------------------------------------------- state
    def generate_zero_state(self, bsz: int) -> List[torch.Tensor]:
        """rwkv7.py:224-235: bsz >= 1 -> batched layout, 0 -> unbatched (bsz-1 ``forward`` path)."""
        L, C, H, N, dev = self.n_layer, self.n_embd, self.n_head, self.head_size, self.device
        if bsz >= 1:
            return [torch.zeros((L, 2, bsz, C), dtype=DTYPE, device=dev),
                    torch.zeros((L, bsz, H, N, N), dtype=DTYPE, device=dev),
                    torch.zeros((bsz,), dtype=torch.int32, device=dev)]
        return [torch.zeros((L, 2, C), dtype=DTYPE, device=dev),
                torch.zeros((L, H, N, N), dtype=DTYPE, device=dev),
                torch.zeros((), dtype=torch.int32, device=dev)]

    # ---------------------------------------------------------------- the step
    def _step(self, tokens: torch.Tensor, B: int, T: int, state: List[torch.Tensor], full_output: bool,
              want_logits: bool = True, slot_idx: Optional[torch.Tensor] = None,
              out: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
        """tokens: int32 device tensor (B*T).  state: batched layout (views allowed)."""
        s0, s1, s2 = state
        C, H, V = self.n_embd, self.n_head, self.vocab_size
        if s0.dtype != DTYPE or s1.dtype != DTYPE or s2.dtype != torch.int32:
            raise TypeError("state must be [fp16, fp16, int32]")
        if s0.dim() != 4 or s0.stride(3) != 1 or s0.stride(2) != C:
            raise ValueError("state[0] must be a B-axis view of an (L,2,B,C) tensor")
        if s1.dim() != 5 or s1.stride(1) != C * HEAD_SIZE or not s1[0, 0].is_contiguous():
            raise ValueError("state[1] must be a B-axis view of an (L,B,H,64,64) tensor")
        if s2.dim() != 1 or (s2.numel() > 1 and s2.stride(0) != 1) or s2.numel() != B:
            raise ValueError("state[2] must be a contiguous int32 (B,) tensor")
        n_slots = s1.shape[1]
        sp = 0
        if slot_idx is not None:
            sp = slot_idx.data_ptr()
        elif n_slots < B:
            raise ValueError("state has fewer rows than sequences")
        graphed = T == 1 and self.graph_decode and not torch.cuda.is_current_stream_capturing()
        ws = self._workspace(B * T)
        logits, staged = None, None
        tok_ptr = tokens.data_ptr()
        if graphed:
            # stage the per-call operands in model-owned buffers so that the graph key is (B, state tensors, switches) only
            tok_buf, slot_buf, lg_buf = self._graph_io(B)
            tok_buf[:B].copy_(tokens.view(-1), non_blocking=True)
            tok_ptr = tok_buf.data_ptr()
            if slot_idx is not None:
                slot_buf[:B].copy_(slot_idx, non_blocking=True)
                sp = slot_buf.data_ptr()
            if want_logits:
                if out is not None:
                    logits = out                       # a caller-owned fixed buffer (the worker's): part of the key
                else:
                    logits = staged = lg_buf[:B]
        elif want_logits:
            rows = B * T if full_output else B
            logits = out if out is not None else torch.empty((rows, V), dtype=DTYPE, device=self.device)
        cargs = (self._handle, B, T, tok_ptr, s0.data_ptr(), s0.stride(0), s0.stride(1), s1.data_ptr(), s1.stride(0),
                 n_slots, s2.data_ptr(), sp, logits.data_ptr() if logits is not None else 0, 1 if full_output else 0,
                 ws.data_ptr(), ws.numel())

        def launch():
            _lib.check(_lib.lib().b200_model_forward(*cargs, _stream(self.device)), "b200_model_forward")

        if graphed:
            # the A/B switches model.cu reads per call are part of what a captured graph froze
            self._step_graphed(cargs[1:] + tuple(os.environ.get(k) for k in _STEP_ENV), launch)
        else:
            launch()
        if logits is None:
            return None
        if staged is not None:
            logits = staged.clone()                    # the reference returns a fresh tensor every call
        return logits.view(B, T, V) if full_output else logits

    def _step_graphed(self, key: tuple, launch) -> None:
        g = self._graphs.get(key)
        if g is not None:
            self._graphs.move_to_end(key)
            self.graph_hits += 1
            g.replay()
            return
        self.graph_misses += 1
        seen = self._graph_seen.get(key, 0) + 1
        if seen < self.graph_after:
            if len(self._graph_seen) > 256:
                self._graph_seen.clear()
            self._graph_seen[key] = seen
            launch()
            return
        self._graph_seen.pop(key, None)
        cur = torch.cuda.current_stream(self.device)
        if self._capture_stream is None:
            self._capture_stream = torch.cuda.Stream(self.device)
        side = self._capture_stream
        g = torch.cuda.CUDAGraph()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            g.capture_begin(capture_error_mode="thread_local")   # other workers' threads keep using CUDA during this capture
            try:
                launch()
            finally:
                g.capture_end()
        cur.wait_stream(side)
        self._graphs[key] = g
        while len(self._graphs) > self.graph_cache_size:
            self._graphs.popitem(last=False)
        g.replay()

    def _tokens(self, idxs) -> torch.Tensor:
        if torch.is_tensor(idxs):
            return idxs.to(device=self.device, dtype=torch.int32, non_blocking=True).reshape(-1).contiguous()
        return torch.tensor(idxs, dtype=torch.int32).reshape(-1).to(self.device, non_blocking=True)

    # Long prompts are processed in chunks of at most this many tokens per call (B * T_chunk): the WKV kernel carries each
    # sequence's state from chunk to chunk through HBM, so the result is bit-identical to one long call, while the activation
    # workspace stays bounded (23 buffers of tokens x C fp16).  The reference instead caps T at 100 in the worker
    # (chirrup/worker.py:177) and runs whatever T it is given in one go.
    max_tokens_per_step = 16384

    def forward_seq_batch(self, idxs: Sequence[Sequence[int]], state: List[torch.Tensor], full_output: bool = False):
        """rwkv7.py:351-382 — rectangular (B,T) token lists, batched state, logits (B,V) or (B,T,V)."""
        B = len(idxs)
        T = len(idxs[0]) if B else 0
        if B == 0:
            return torch.empty((0, self.vocab_size), dtype=DTYPE, device=self.device)
        if full_output or B * T <= self.max_tokens_per_step:
            return self._step(self._tokens(idxs), B, T, state, full_output)
        return self.prefill(idxs, state)

    def prefill(self, idxs, state: List[torch.Tensor], chunk_len: Optional[int] = None, want_logits: bool = True,
                slot_idx: Optional[torch.Tensor] = None):
        """Chunked prefill of B sequences x T tokens (config 4: 4k-token prompts, bsz 32): T is cut into chunks of
        ``chunk_len`` tokens (default: as many as fit ``max_tokens_per_step``); only the last chunk computes ln_out + head.
        Returns the last-token logits (B,V) (or None)."""
        tok = idxs if torch.is_tensor(idxs) else torch.tensor(idxs, dtype=torch.int32)
        tok = tok.to(device=self.device, dtype=torch.int32, non_blocking=True)
        B, T = tok.shape
        if chunk_len is None:
            chunk_len = max(1, self.max_tokens_per_step // max(B, 1))
        out = None
        for t0 in range(0, T, chunk_len):
            t1 = min(T, t0 + chunk_len)
            last = t1 == T
            out = self._step(tok[:, t0:t1].contiguous().view(-1), B, t1 - t0, state, False, want_logits=want_logits and last,
                             slot_idx=slot_idx)
        return out

    def forward_seq_batch_seperate(self, idxs, state, full_output: bool = False):
        """rwkv7.py:555-563 (the entry point Chirrup's worker calls, chirrup/worker.py:704, :775)."""
        return self.forward_seq_batch(idxs, state, full_output)

    def forward_batch_same_length(self, tokens, state, full_output: bool = False):
        """rwkv7.py:282-285."""
        assert len(set(len(x) for x in tokens)) == 1, "here all sequences must have the same length"
        return self.forward_seq_batch(tokens, state, full_output)

    def forward_batch(self, tokens: List[List[int]], state: List[torch.Tensor], full_output: bool = False):
        """rwkv7.py:250-280 — ragged lengths: step the common prefix of the still-active sequences.  The reference gathers and
        scatters state rows by fancy indexing; here the kernels address the rows through a slot-index vector instead."""
        assert isinstance(tokens, list)
        lengths = [len(x) for x in tokens]
        if len(set(lengths)) == 1 and not full_output:
            return self.forward_batch_same_length(tokens, state, full_output)
        bsz = len(tokens)
        pos = [0] * bsz
        V = self.vocab_size
        out = torch.empty((bsz, V), dtype=DTYPE, device=self.device) if not full_output else \
            [torch.empty((0, V), dtype=DTYPE, device=self.device) for _ in range(bsz)]
        while True:
            active = [i for i in range(bsz) if pos[i] < lengths[i]]
            if not active:
                break
            step = min(lengths[i] - pos[i] for i in active)
            chunk = [tokens[i][pos[i]:pos[i] + step] for i in active]
            idx = torch.tensor(active, dtype=torch.int32, device=self.device)
            elapsed = state[2][idx.long()].contiguous()
            new = self._step(self._tokens(chunk), len(active), step, [state[0], state[1], elapsed], full_output, slot_idx=idx)
            state[2][idx.long()] = elapsed
            for k_, i in enumerate(active):
                if not full_output:
                    out[i] = new[k_]
                else:
                    out[i] = torch.cat([out[i], new[k_]], dim=0)
                pos[i] += step
        return out

    # ---------------------------------------------------------------- bsz-1 API (unbatched state layout)
    def _batched_view(self, state):
        return [state[0].unsqueeze(2), state[1].unsqueeze(1), state[2].view(1)]

    def forward(self, idx, state, full_output: bool = False):
        """rwkv7.py:237-248."""
        if isinstance(idx, list):
            if len(idx) > 1:
                return self.forward_seq(idx, state, full_output)
            return self.forward_one(idx[0], state)
        return self.forward_one(idx, state)

    def _ensure_ffn_value_rows(self) -> None:
        """bsz-1 path only: ffn.value.weight once more in the reference's transformed (4C, C) layout (rwkv7.py:217-218), so
        that ``RWKV_x070_CMix_one``'s ReLU^2-sparse product (rwkv7.py:653-662, ``rwkv_mm_sparsity`` / ``SPMV_OP``) can skip
        the rows whose key activation is zero.  The batched path keeps the checkpoint's (C, 4C) = K-major layout for the
        tensor-core GEMM; the second copy (4.3 GB at 7.2B) exists only once ``forward`` / ``forward_one`` has been called."""
        if self._ffn_rows is not None:
            return
        rows = [self._w[f"blocks.{i}.ffn.value.weight"].t().contiguous() for i in range(self.n_layer)]
        ptrs = (ctypes.c_void_p * len(rows))(*[t.data_ptr() for t in rows])
        _lib.check(_lib.lib().b200_model_set_ffn_value_rows(self._handle, ptrs), "b200_model_set_ffn_value_rows")
        self._ffn_rows = rows
        for i, t in enumerate(rows):
            self.z[f"blocks.{i}.ffn.value.weight"] = t
        self._graphs.clear()
        self._graph_seen.clear()

    def forward_one(self, idx: int, state):
        """rwkv7.py:287-316 — one token, unbatched state; returns logits (V,).  The reference's ``forward_one(x, state)`` takes
        the EMBEDDING VECTOR of the token (``forward`` looks it up, rwkv7.py:237-248); here the embedding gather is part of the
        step, so this takes the token id — a (C,) tensor is rejected rather than misread."""
        if torch.is_tensor(idx) and idx.numel() != 1:
            raise TypeError("forward_one takes a token id here (the embedding lookup is inside the step); the reference's "
                            "forward_one(x: (C,) embedding, state) form is not supported — call forward(token_id, state)")
        self._ensure_ffn_value_rows()
        return self._step(self._tokens([int(idx)]), 1, 1, self._batched_view(state), False)[0]

    def forward_seq(self, idx: List[int], state, full_output: bool = False):
        """rwkv7.py:318-349 — one sequence, unbatched state; logits (V,) or (T,V)."""
        out = self._step(self._tokens(list(idx)), 1, len(idx), self._batched_view(state), full_output)
        return out[0]


def make_args(model_name: str = "", vocab_size: int = 65536, head_size: int = 64):
    return types.SimpleNamespace(MODEL_NAME=model_name, vocab_size=vocab_size, head_size=head_size)
