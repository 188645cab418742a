"""Continuous-batching driver over the B200 model: Chirrup's ``AsyncEngineCore`` / ``Worker`` contract
(``/root/reference/chirrup/engine_core.py:60-332``, ``chirrup/worker.py:107-889``, ``chirrup/interface.py:31-142``,
``chirrup/core_structure.py``) kept at the API level, with the state handling redesigned for the C ABI:

* the worker owns one ``batch_state`` (``generate_zero_state(batch_size)``, worker.py:260) and a slot table with the
  reference's ``StateCategory`` classes; but instead of physically swapping state rows so that each class is contiguous
  (``_organize_batch`` / ``_switch_batch``, worker.py:304-383 — three 16.5 MiB copies per swap at 7.2B) the kernels take a
  slot-index vector, so grouping is integer-only (``group_slots``; offsets identical to ``min_swaps_to_target_fast``);
* decode step = ``forward`` over the DECODE + ONE_PREFILL slots with T = 1 (worker.py:671-742); prefill = up to
  ``max_forward_seq_len_per_forward`` = 100 tokens of every SEQ slot, every ``decode_prefill_ratio`` = 5 decode steps
  (worker.py:173-179, :744-776, :854-856); admission of at most ``max_prefill_count`` new tasks per loop (worker.py:143, :570-669);
* sampling is ONE kernel and ONE batched D2H per step instead of a full sort, two (B,V) fp32 temporaries and one ``.item()`` per
  slot (worker.py:713-740): the arg-max kernel when every row is greedy without penalties, else the fused sampling kernel
  ``b200_sample_f16`` (forbidden mask, decay + presence / frequency penalties, temperature, top-k, top-p, draw, penalty update;
  samplers.py:98-168, worker.py:527-547) over device-resident per-slot parameters and penalty rows.

One worker per GPU, all pulling from one task queue (engine_core.py:71, :135-153): request-level data parallelism, no collective.
"""
from __future__ import annotations

import asyncio
import enum
import queue
from collections import deque
import threading
import time
import uuid
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch

DEFAULT_STOP_TOKENS = [0, 261, 24281]                      # core_structure.py:15
DEFAULT_SAMPLING_CONFIG = {"temperature": 1.0, "top_p": 0.3, "top_k": 0, "presence_penalty": 0.5,
                           "frequency_penalty": 0.5, "penalty_decay": 0.996, "max_tokens": 8192}   # core_structure.py:28-36


class FinishReason(enum.IntEnum):                           # core_structure.py:39-58
    STOP = 0
    LENGTH = 1
    ABORT = 2

    def __str__(self):
        return ("stop", "length", "abort")[self.value]


class RequestStatus(enum.IntEnum):                          # core_structure.py:62-90
    WAITING = enum.auto()
    RUNNING = enum.auto()
    FINISHED_STOPPED = enum.auto()
    FINISHED_LENGTH_CAPPED = enum.auto()
    FINISHED_ABORTED = enum.auto()

    @staticmethod
    def is_finished(status: "RequestStatus") -> bool:
        return status > RequestStatus.RUNNING


class StateCategory(enum.IntEnum):                          # worker.py:86-92 (same order: it defines the batch layout)
    FORWARD_ONE_DECODE = enum.auto()
    FORWARD_ONE_PREFILL = enum.auto()
    FORWARD_ONE_SUSPENDED = enum.auto()
    FORWARD_SEQ = enum.auto()
    FINISHED = enum.auto()
    EMPTY = enum.auto()


def group_slots(categories: Sequence[StateCategory], elements: Optional[Sequence[StateCategory]] = None):
    """Integer-only replacement of ``min_swaps_to_target_fast`` (worker.py:43-78).  Returns ``(groups, offsets)``:
    ``groups[c]`` = slot indices currently in category c (ascending), ``offsets`` = the ``[(start, end), ...]`` ranges the
    reference would obtain after physically sorting — identical values, but here no state row moves: the kernels are handed
    ``groups[c]`` as their slot-index vector."""
    elements = list(elements) if elements is not None else list(sorted(StateCategory))
    groups: Dict[StateCategory, List[int]] = {c: [] for c in elements}
    for i, c in enumerate(categories):
        groups.setdefault(c, []).append(i)
    offsets: List[Tuple[int, int]] = []
    off = 0
    for c in elements:
        n = len(groups.get(c, []))
        offsets.append((off, off + n))
        off += n
    return groups, offsets


@dataclass
class Task:                                                 # core_structure.py:93-180 (same field names)
    prompt_str: str = ""
    prefill_tokens: List[int] = field(default_factory=list)
    state: Optional[List[torch.Tensor]] = None
    task_id: Optional[str] = None
    priority: int = 0
    temperature: float = DEFAULT_SAMPLING_CONFIG["temperature"]
    top_p: float = DEFAULT_SAMPLING_CONFIG["top_p"]
    top_k: int = DEFAULT_SAMPLING_CONFIG["top_k"]
    presence_penalty: float = DEFAULT_SAMPLING_CONFIG["presence_penalty"]
    frequency_penalty: float = DEFAULT_SAMPLING_CONFIG["frequency_penalty"]
    penalty_decay: float = DEFAULT_SAMPLING_CONFIG["penalty_decay"]
    max_tokens: Optional[int] = DEFAULT_SAMPLING_CONFIG["max_tokens"]
    stop_tokens: List[int] = field(default_factory=lambda: list(DEFAULT_STOP_TOKENS))
    forbidden_tokens: List[int] = field(default_factory=list)
    cache_prefill: bool = False
    cache_prefill_padding: int = 0
    return_logits: bool = False
    state_ready: Any = None                                 # optional CUDA event: `state` (device tensors) is complete after it
    output_queue: Any = None                                # thread-safe channel back to the asyncio side
    task_event_queue: "queue.Queue" = field(default_factory=queue.Queue)
    request_status: RequestStatus = field(init=False, default=RequestStatus.WAITING)
    generated_tokens: List[int] = field(init=False, default_factory=list)

    def __post_init__(self):
        if self.task_id is None:
            self.task_id = str(uuid.uuid4())

    def is_finished(self) -> bool:
        return RequestStatus.is_finished(self.request_status)

    def is_greedy(self) -> bool:
        """Rows the sampler resolves to the arg-max: top_k == 1 or top_p == 0 (the nucleus cutoff is then the top probability).
        temperature <= 0 is NOT greedy here: the worker replaces it by 1.0 at admission and still samples with the request's
        top_p / top_k (chirrup/worker.py:611-613)."""
        return self.top_k == 1 or self.top_p == 0


@dataclass
class ModelLoadConfig:                                      # core_structure.py:182-232 (fields the worker reads)
    model_path: str = ""
    vocab_size: int = 65536
    head_size: int = 64
    state_dict: Optional[Dict[str, torch.Tensor]] = None    # synthetic checkpoints (no .pth on disk)
    vocab_path: Optional[str] = None
    synthetic: Optional[Dict[str, Any]] = None              # {"geom": "7.2B", "seed": 0, "recipe": "trained_like"}: each worker
    #                                                         generates the seeded random-init checkpoint on its own GPU


@dataclass
class _Slot:
    task: Optional[Task] = None
    category: StateCategory = StateCategory.EMPTY
    pending: List[int] = field(default_factory=list)        # prompt tokens not yet consumed
    next_token: int = 0
    prompt: Tuple[int, ...] = ()                            # the task's prompt as admitted (key material of cache_prefill)
    prefill_cached: bool = False


_MODEL_LOAD_LOCK = threading.Lock()                          # worker.py:82-83


class IdTokenizer:
    """Stand-in when no vocabulary file is available (synthetic runs): prompts are token-id lists, text is empty."""

    def encode(self, s: str) -> List[int]:
        return [int(t) for t in s.split()] if s.strip() else [0]

    def decode(self, toks: Sequence[int]) -> str:
        return ""


class Worker:
    """One continuous-batching worker bound to one GPU (worker.py:107-889)."""

    def __init__(self, worker_id: str, gpu_id: Sequence[int], model_config: ModelLoadConfig, task_queue: "queue.Queue",
                 master_event_queue: Optional["queue.Queue"] = None, worker_event_queue: Optional["queue.Queue"] = None,
                 batch_size: int = 32, tokenizer=None, model=None):
        self.worker_id, self.gpu_id, self.model_config = worker_id, list(gpu_id), model_config
        self.task_queue, self.master_event_queue = task_queue, master_event_queue
        self.worker_event_queue = worker_event_queue or queue.Queue()
        self.batch_size = batch_size
        self.max_prefill_count = max(batch_size // 8, 1)      # worker.py:143
        self.min_forward_seq_len = 10                         # worker.py:173-179
        self.max_forward_seq_len_per_forward = 100
        self.decode_prefill_ratio = 5
        self.tokenizer = tokenizer or IdTokenizer()
        self.model = model
        self.slots: List[_Slot] = [_Slot() for _ in range(batch_size)]
        # full_steps / full_time_s: consecutive loops in which EVERY slot took a T = 1 step (the steady state of a loaded worker)
        self.steps = {"decode": 0, "prefill": 0, "decode_tokens": 0, "prefill_tokens": 0, "full_steps": 0, "full_time_s": 0.0}
        self._last_full = None
        self._occurrence = None                               # device penalty state, allocated with the first non-plain request
        self._occ_host = None
        self._par_dirty = False
        self._inflight: Optional[Dict[str, Any]] = None       # decode step launched, host side not yet done
        self._rows_cache = None
        self._step_bufs = None
        self.pipeline = True
        self.sample_seed = 0x5EED0000 + sum(map(ord, str(worker_id)))
        self.perf_queue: Optional["queue.Queue"] = None       # worker -> engine telemetry ("worker_performance", worker.py:858-884)
        self._loop_times: "deque[float]" = deque(maxlen=10)   # worker.py:184
        self.no_penalty_token_ids = (33, 10, 49, 50, 51, 52, 53, 54, 55, 56, 57, 58)   # worker.py:173: "!", "\n", digits
        self._no_penalty_dev = None
        self._stop = False

    # ------------------------------------------------------------------ init
    def _init_worker(self):
        if self.model is not None and getattr(self.model, "device", None) is not None \
                and torch.device(self.model.device).type != "cuda":
            self.device = torch.device(self.model.device)     # CPU test double of the model (host-logic tests)
        else:
            torch.cuda.set_device(self.gpu_id[0])             # worker.py:255-256
            self.device = torch.device("cuda", self.gpu_id[0])
        if self.model is None:
            from .rwkv7 import RWKV_x070, make_args

            with _MODEL_LOAD_LOCK:                            # worker.py:209
                args = make_args(self.model_config.model_path, self.model_config.vocab_size, self.model_config.head_size)
                sd = self.model_config.state_dict
                if sd is None and self.model_config.synthetic:
                    from .synth import GEOMETRIES, make_checkpoint

                    syn = self.model_config.synthetic
                    sd = make_checkpoint(GEOMETRIES[syn["geom"]], seed=syn.get("seed", 0), recipe=syn.get("recipe", "trained_like"),
                                         device=self.device, dtype=torch.float16, n_layer=syn.get("n_layer"))
                self.model = RWKV_x070(args, state_dict=sd, device=self.device)
                del sd
        self.batch_state = self.model.generate_zero_state(self.batch_size)      # worker.py:260
        self._out_host = torch.empty((2, self.batch_size), dtype=torch.int32)    # double-buffered: one per step in flight
        if self.device.type == "cuda":
            self._out_host = self._out_host.pin_memory()
        if self.master_event_queue is not None:
            self.master_event_queue.put(("worker_loaded", self.worker_id))      # worker.py:186-202

    # ------------------------------------------------------------------ slot table
    def categories(self) -> List[StateCategory]:
        return [s.category for s in self.slots]

    @staticmethod
    def _keep(task: Task) -> int:
        """Prompt tokens a sequence forward must leave unconsumed beyond the next input token: max(padding - 1, 0)
        (worker.py:649, :746-750) — the state is wanted for caching at exactly that point."""
        return max(int(task.cache_prefill_padding) - 1, 0)

    def _classify(self, slot: _Slot, admission: bool = False) -> None:
        """DECODE / ONE_PREFILL / SEQ by remaining prompt length vs min_forward_seq_len.  `pending` holds the next input token
        plus what the reference calls task.prefill_tokens, so its `len(prefill_tokens)` is n - 1 here.  At admission the
        padding term counts (worker.py:646-654); after a sequence forward it does not (worker.py:437-446), but the slot leaves
        the SEQ group when only the padding is left (worker.py:411-420)."""
        n = len(slot.pending)
        rem, keep = n - 1, self._keep(slot.task) if slot.task is not None else 0
        if n == 0:
            slot.category = StateCategory.FORWARD_ONE_DECODE
        elif rem - keep <= 0 or (rem - keep if admission else rem) < self.min_forward_seq_len:
            slot.category = StateCategory.FORWARD_ONE_PREFILL
        else:
            slot.category = StateCategory.FORWARD_SEQ

    def _maybe_cache_prefill(self, pos: int) -> None:
        """("cache_prefill", {"state", "prefilled_tokens"}) when exactly max(padding - 1, 0) prompt tokens remain behind the next
        input token (worker.py:411-435 after a sequence forward, :457-476 on the one-token path).  The reference copies the
        slot to the host; here the copy stays on the GPU (`state_move.extract_slot`, stream-ordered behind the forward that
        produced it) and carries a CUDA event under "ready" for consumers on other streams / devices."""
        from . import state_move

        s = self.slots[pos]
        t = s.task
        if t is None or not t.cache_prefill or s.prefill_cached or not s.pending:
            return
        if len(s.pending) - 1 != self._keep(t):
            return
        s.prefill_cached = True
        key = s.prompt[:len(s.prompt) - len(s.pending)]
        if not key or t.output_queue is None:
            return
        st = state_move.extract_slot(self.batch_state, pos)
        ready = None
        if self.device.type == "cuda":
            ready = torch.cuda.Event()
            ready.record()
        t.output_queue.put_nowait(("cache_prefill", {"state": st, "prefilled_tokens": tuple(key), "ready": ready}))

    def _fill_task_pool(self) -> int:
        """Admit waiting tasks into EMPTY slots, at most max_prefill_count per loop (worker.py:570-669)."""
        from . import state_move

        groups, _ = group_slots(self.categories())
        free = groups[StateCategory.EMPTY]
        budget = self.max_prefill_count - len(groups[StateCategory.FORWARD_SEQ])
        admitted = 0
        while free and budget > 0:
            try:
                task: Task = self.task_queue.get_nowait()
            except queue.Empty:
                break
            pos = free.pop(0)
            slot = self.slots[pos]
            if task.state is None:
                state_move.clear_slot(self.batch_state, pos)                     # generate_zero_state(1), worker.py:586-589
            else:
                ready = task.state_ready if task.state_ready is not None else getattr(task.state, "ready", None)
                if ready is not None and self.device.type == "cuda":
                    torch.cuda.current_stream().wait_event(ready)                # the cached copy may still be in flight
                state_move.inject_slot(self.batch_state, pos, task.state)        # task.state[i].cuda(), worker.py:591-597
            self._set_slot_params(pos, task)
            if self._occ_host is not None:
                self._occ_host[pos].zero_()
            toks = list(task.prefill_tokens) or [0]
            slot.task, slot.pending, slot.prompt, slot.prefill_cached = task, toks, tuple(toks), False
            task.request_status = RequestStatus.RUNNING
            self._classify(slot, admission=True)
            admitted += 1
            budget -= 1
        return admitted

    # ------------------------------------------------------------------ forwards
    def _state_for(self, idx: torch.Tensor):
        el = self.batch_state[2][idx.long()].contiguous()
        return [self.batch_state[0], self.batch_state[1], el], el

    def _run_forward_seq(self) -> int:
        """One prefill chunk over the SEQ group (worker.py:744-776): every slot contributes exactly T tokens, logits discarded."""
        groups, _ = group_slots(self.categories())
        seq = groups[StateCategory.FORWARD_SEQ]
        if not seq:
            return 0
        T = min(self.max_forward_seq_len_per_forward,
                min(len(self.slots[p].pending) - 1 - self._keep(self.slots[p].task) for p in seq))      # worker.py:746-754
        assert T > 0
        toks = [self.slots[p].pending[:T] for p in seq]
        idx = torch.tensor(seq, dtype=torch.int32, device=self.device)
        st, el = self._state_for(idx)
        self.model._step(self.model._tokens(toks), len(seq), T, st, False, want_logits=False, slot_idx=idx)
        self.batch_state[2][idx.long()] = el
        for p in seq:
            s = self.slots[p]
            del s.pending[:T]
            self._classify(s)
            self._maybe_cache_prefill(p)
        self.steps["prefill"] += 1
        self.steps["prefill_tokens"] += T * len(seq)
        return T * len(seq)

    # ------------------------------------------------------------------ sampling
    def _sampling_state(self, V: int) -> None:
        """Penalty state and per-slot sampler parameters on the device (worker.py:262-300: occurrence, alpha_presence_vector,
        temperature / top_p / top_k / penalty tensors, one row per slot) plus pinned host mirrors of the parameters."""
        if self._occurrence is not None:
            return
        dev, n = self.device, self.batch_size
        self._occurrence = torch.zeros((n, V), dtype=torch.float32, device=dev)
        self._presence = torch.zeros((n, V), dtype=torch.float32, device=dev)
        self._par_host = torch.zeros((6, n), dtype=torch.float32).pin_memory()   # temperature, top_p, decay, freq, presence, top_k
        self._par_dev = torch.zeros((6, n), dtype=torch.float32, device=dev)
        self._topk_dev = torch.zeros((n,), dtype=torch.int32, device=dev)
        self._par_dirty = True
        for pos, sl in enumerate(self.slots):
            if sl.task is not None:
                self._set_slot_params(pos, sl.task)

    def _set_slot_params(self, pos: int, t: Task) -> None:
        """Per-slot sampler parameters of a newly admitted task (worker.py:599-640) and a clean penalty row."""
        if self._occurrence is None:
            return
        greedy = t.is_greedy()
        h = self._par_host
        f16 = lambda v: float(torch.tensor(v, dtype=torch.float16))   # the worker keeps these as fp16 tensors (worker.py:611-635)
        h[0, pos] = 1.0 if greedy else f16(t.temperature if t.temperature > 0 else 1.0)   # worker.py:611-613
        h[1, pos] = 1.0 if greedy else f16(t.top_p)
        h[2, pos], h[3, pos], h[4, pos] = f16(t.penalty_decay), f16(t.frequency_penalty), t.presence_penalty   # presence: fp32 (:636-640)
        h[5, pos] = 1.0 if greedy else float(max(t.top_k, 0))
        self._occurrence[pos].zero_()
        self._presence[pos].zero_()
        self._par_dirty = True

    def _sample_launch(self, rows: List[int], logits: torch.Tensor) -> torch.Tensor:
        """Next token for each row of `logits` (row i belongs to slot rows[i]) as an int32 tensor on the model's device, without
        any host synchronisation.  On the GPU one C-ABI call per step: the arg-max kernel when every row is greedy with zero
        penalties, else the fused sampling kernel
        (``b200_sample_f16``: forbidden mask + decay / presence / frequency penalties on every decode row, temperature, top-k,
        top-p, draw, penalty update — worker.py:713-740, samplers.py:98-168, worker.py:527-547)."""
        from . import ops

        n = len(rows)
        tasks = [self.slots[p].task for p in rows]
        if not logits.is_cuda:                                                   # CPU test double of the model: host logic only
            return torch.tensor(self._sample_host(rows, logits), dtype=torch.int32)
        plain = all(t is None or (t.is_greedy() and t.presence_penalty == 0 and t.frequency_penalty == 0
                                  and not t.forbidden_tokens) for t in tasks)
        if plain:
            ids = ops.argmax(logits)
        else:
            V = logits.shape[1]
            self._sampling_state(V)
            if self._par_dirty:
                self._par_dev.copy_(self._par_host, non_blocking=True)
                self._topk_dev.copy_(self._par_dev[5].to(torch.int32))
                self._par_dirty = False
            slot = torch.tensor(rows, dtype=torch.int32, device=self.device)
            forb = None
            nf = max((len(t.forbidden_tokens) for t in tasks if t is not None), default=0)
            if nf:
                fh = torch.full((n, nf), -1, dtype=torch.int32)
                for i, t in enumerate(tasks):
                    if t is not None and t.forbidden_tokens:
                        fh[i, :len(t.forbidden_tokens)] = torch.tensor(t.forbidden_tokens, dtype=torch.int32)
                forb = fh.to(self.device)
            if not logits.is_contiguous():
                logits = logits.contiguous()
            pd = self._par_dev
            if self._no_penalty_dev is None:
                self._no_penalty_dev = torch.tensor(self.no_penalty_token_ids, dtype=torch.int32, device=self.device)
            # mode 1: the sampler the reference worker imports (worker.py:12 -> samplers.py:171-255); tokens in
            # no_penalty_token_ids do not accumulate a frequency penalty (worker.py:508)
            ids = ops.sample(logits, pd[0], pd[1], self._topk_dev, slot=slot, occurrence=self._occurrence, presence=self._presence,
                             penalty_decay=pd[2], frequency_penalty=pd[3], presence_penalty=pd[4], forbidden=forb,
                             no_penalty=self._no_penalty_dev, seed=self.sample_seed, step=self.steps["decode"], mode=1)
        return ids

    def _sample_host(self, rows: List[int], logits: torch.Tensor) -> List[int]:
        """torch restatement of the same step for the CPU test double (never runs on the product path)."""
        out = logits.float().argmax(-1).tolist()
        for i, p in enumerate(rows):
            t = self.slots[p].task
            if t is None or t.is_greedy():
                continue
            temperature = t.temperature if t.temperature > 0 else 1.0
            if self._occ_host is None:
                self._occ_host = torch.zeros((self.batch_size, logits.shape[1]), dtype=torch.float32)
            occ = self._occ_host[p]
            lg = logits[i].float()
            if t.forbidden_tokens:
                lg[torch.tensor(t.forbidden_tokens)] -= 1e10
            lg = lg - t.presence_penalty * (occ > 0).float() - t.frequency_penalty * occ
            probs = torch.softmax(lg / temperature, dim=-1)
            sp, si = torch.sort(probs, descending=True)
            keep = torch.cumsum(sp, 0) <= t.top_p
            if t.top_k > 0:
                keep[t.top_k:] = False
            keep[0] = True
            sp = sp * keep
            tok = int(si[torch.multinomial(sp / sp.sum(), 1)])
            occ.mul_(t.penalty_decay)
            occ[tok] += 1.0
            out[i] = tok
        return out

    def _finish(self, pos: int, status: RequestStatus) -> None:
        slot = self.slots[pos]
        t = slot.task
        t.request_status = status
        if t.output_queue is not None:
            reason = {RequestStatus.FINISHED_STOPPED: FinishReason.STOP, RequestStatus.FINISHED_LENGTH_CAPPED: FinishReason.LENGTH,
                      RequestStatus.FINISHED_ABORTED: FinishReason.ABORT}[status]
            t.output_queue.put_nowait(("finish", reason, t))
        self.slots[pos] = _Slot()                                                # -> EMPTY

    def _launch_decode(self, rows: List[int], feed: Optional[torch.Tensor]) -> Dict[str, Any]:
        """Enqueue one T = 1 step over `rows` (worker.py:671-742): forward, sampling and the D2H of the drawn ids into a pinned
        buffer — NO host synchronisation.  `feed` = the previous step's device-resident ids when they are exactly the input
        tokens of this step (every row decoding, same rows), else None: tokens come from the host-side slot table."""
        n = len(rows)
        for p in rows:                                                           # worker.py:457-476
            if self.slots[p].pending and self.slots[p].task.cache_prefill:
                self._maybe_cache_prefill(p)
        if feed is None:
            toks = []
            for p in rows:
                s = self.slots[p]
                toks.append([s.pending.pop(0) if s.pending else s.next_token])
            tok_dev = self.model._tokens(toks)
        else:
            ids_prev, pos_in_prev = feed
            if ids_prev.numel() == n and len(pos_in_prev) == n and all(pos_in_prev.get(p) == i for i, p in enumerate(rows)):
                tok_dev = ids_prev                                              # same rows, all decoding: no host data at all
            else:                                                               # rows left / joined: gather on the device
                src, host = [], []
                for p in rows:
                    s = self.slots[p]
                    if s.pending:
                        host.append(s.pending.pop(0))
                        src.append(-1)
                    else:
                        host.append(0)
                        src.append(pos_in_prev[p])
                both = torch.tensor([src, host], dtype=torch.int32).to(self.device, non_blocking=True)
                tok_dev = torch.where(both[0] >= 0, ids_prev[both[0].clamp(min=0).long()], both[1]).contiguous()
        if self._rows_cache is None or self._rows_cache[0] != rows:
            idx = torch.tensor(rows, dtype=torch.int32, device=self.device)
            self._rows_cache = (list(rows), idx, idx.long())
        idx, idx_long = self._rows_cache[1], self._rows_cache[2]
        # fixed input / output buffers: the same addresses every step, so the model replays the step from a CUDA graph
        if self._step_bufs is None:
            cap, V = len(self.slots), self.model.vocab_size
            self._step_bufs = (torch.empty((cap,), dtype=torch.int32, device=self.device),
                               torch.empty((cap,), dtype=torch.int32, device=self.device),
                               torch.empty((cap, V), dtype=torch.float16, device=self.device))
        tok_in, el, out = self._step_bufs[0][:n], self._step_bufs[1][:n], self._step_bufs[2][:n]
        tok_in.copy_(tok_dev.view(-1))
        torch.index_select(self.batch_state[2], 0, idx_long, out=el)
        logits = self.model._step(tok_in, n, 1, [self.batch_state[0], self.batch_state[1], el], False, slot_idx=idx, out=out)
        self.batch_state[2][idx_long] = el
        sample_rows = [i for i, p in enumerate(rows) if not self.slots[p].pending]    # prompt fully consumed: logits count
        rec: Dict[str, Any] = {"rows": list(rows), "tasks": [self.slots[p].task for p in rows], "sample_rows": sample_rows,
                               "ids": None, "host": None, "event": None, "logits": None}
        if sample_rows:
            sel = logits if len(sample_rows) == n else logits[torch.tensor(sample_rows, device=logits.device)]
            if any(self.slots[rows[i]].task.return_logits for i in sample_rows):
                rec["logits"] = sel.detach().clone()                                 # before the in-place penalties
            ids = self._sample_launch([rows[i] for i in sample_rows], sel)
            rec["ids"] = ids
            if ids.is_cuda:
                buf = self._out_host[self.steps["decode"] & 1]
                buf[:len(sample_rows)].copy_(ids, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record()
                rec["host"], rec["event"] = buf, ev
            else:
                rec["host"] = ids
        self.steps["decode"] += 1
        return rec

    def _finish_decode(self, rec: Dict[str, Any]) -> int:
        """Host side of a launched step: wait for ITS ids only, deliver tokens, stop / length checks (worker.py:478-510, :738-742).
        Rows whose slot no longer holds the task the step was launched for (finished or aborted meanwhile) are dropped."""
        rows, sample_rows = rec["rows"], rec["sample_rows"]
        if not sample_rows:
            return 0
        if rec["event"] is not None:
            rec["event"].synchronize()
        new = rec["host"][:len(sample_rows)].tolist()
        delivered = 0
        for j, (i, tok) in enumerate(zip(sample_rows, new)):
            p = rows[i]
            s = self.slots[p]
            t = s.task
            if t is None or t is not rec["tasks"][i]:
                continue
            delivered += 1
            if tok in t.stop_tokens:                                             # worker.py:488-490: a stop token ends the request
                self._finish(p, RequestStatus.FINISHED_STOPPED)                  # and is neither recorded nor sent to the client
                continue
            s.next_token = tok
            s.category = StateCategory.FORWARD_ONE_DECODE
            t.generated_tokens.append(tok)
            if t.output_queue is not None:
                t.output_queue.put_nowait(("token", tok, self.tokenizer.decode([tok]),
                                           rec["logits"][j] if (t.return_logits and rec["logits"] is not None) else None))
            if t.max_tokens is not None and len(t.generated_tokens) >= t.max_tokens:
                self._finish(p, RequestStatus.FINISHED_LENGTH_CAPPED)
        self.steps["decode_tokens"] += delivered
        return delivered

    def _run_forward_one(self) -> int:
        """One T = 1 step over the DECODE and ONE_PREFILL groups (worker.py:671-742), software-pipelined one step deep: when the
        step in flight covers exactly the rows that decode next, the NEXT step is enqueued first — its input tokens are the
        in-flight step's ids, still on the device — and only then is the in-flight step's host work done, under the GPU time
        of the next one.  A request that finishes is noticed one step late; the extra token computed for it is dropped."""
        groups, _ = group_slots(self.categories())
        dec, pre = groups[StateCategory.FORWARD_ONE_DECODE], groups[StateCategory.FORWARD_ONE_PREFILL]
        rows = dec + pre                                                         # decode rows first, like the reference's layout
        prev, self._inflight = self._inflight, None
        if prev is not None and self.pipeline and prev["ids"] is not None:
            prow, ptask = prev["rows"], prev["tasks"]
            pos_in_prev = {}                                                     # slot -> position of its id in the in-flight step
            for j, i in enumerate(prev["sample_rows"]):
                p = prow[i]
                if self.slots[p].task is ptask[i] and not self.slots[p].pending:
                    pos_in_prev[p] = j

            def capped(p):                                                       # the in-flight token is this request's last one
                t = self.slots[p].task
                return p in pos_in_prev and t.max_tokens is not None and len(t.generated_tokens) + 1 >= t.max_tokens

            nxt = [p for p in rows if not capped(p)]
            # every row's input token must be known without waiting: on the device (in-flight id) or in its prompt
            if nxt and all(p in pos_in_prev or self.slots[p].pending for p in nxt):
                self._inflight = self._launch_decode(nxt, (prev["ids"], {p: pos_in_prev[p] for p in nxt if p in pos_in_prev}))
                self._finish_decode(prev)
                return len(nxt)
        if prev is not None:
            self._finish_decode(prev)
            groups, _ = group_slots(self.categories())                           # finished slots left the batch
            dec, pre = groups[StateCategory.FORWARD_ONE_DECODE], groups[StateCategory.FORWARD_ONE_PREFILL]
            rows = dec + pre
        if not rows:
            return len(prev["rows"]) if prev is not None else 0
        self._inflight = self._launch_decode(rows, None)
        if not self.pipeline:
            rec, self._inflight = self._inflight, None
            self._finish_decode(rec)
        return len(rows)

    def _process_events(self) -> None:
        """Shutdown / abort signals (worker.py:385-405, :512-525)."""
        try:
            while True:
                ev = self.worker_event_queue.get_nowait()
                if ev == "shutdown" or (isinstance(ev, tuple) and ev and ev[0] == "shutdown"):
                    self._stop = True
        except queue.Empty:
            pass
        for pos, s in enumerate(self.slots):
            if s.task is not None and not s.task.task_event_queue.empty():
                try:
                    if s.task.task_event_queue.get_nowait() == "abort":
                        self._finish(pos, RequestStatus.FINISHED_ABORTED)
                except queue.Empty:
                    pass

    def step(self) -> int:
        """One iteration of the main loop (worker.py:793-889): events, admission, decode step, periodic prefill chunk."""
        t_loop = time.perf_counter()
        self._process_events()
        self._fill_task_pool()
        layout = group_slots(self.categories()) if self.perf_queue is not None else None   # the batch layout this loop runs
        n = self._run_forward_one()
        if n == self.batch_size:
            now = time.perf_counter()
            if self._last_full is not None:
                self.steps["full_steps"] += 1
                self.steps["full_time_s"] += now - self._last_full
            self._last_full = now
        else:
            self._last_full = None
        if self.steps["decode"] % self.decode_prefill_ratio == 0 or n == 0:      # worker.py:854-856
            n += self._run_forward_seq()
        if n:
            self._loop_times.append(time.perf_counter() - t_loop)
            if layout is not None:
                self._emit_performance(*layout)
        return n

    def _emit_performance(self, groups, offsets) -> None:
        """(worker_id, "worker_performance", {...}) with the reference's payload keys (worker.py:858-884), consumed by
        ``AsyncEngineCore.iter_worker_performance`` (engine_core.py:290-325).  Dropped when the consumer lags (bounded queue)."""
        names = ("decode_offset", "one_prefill_offset", "decode_suspended_offset", "seq_perfill_offset", "accomplished_offset")
        mem = torch.cuda.max_memory_allocated(self.device) / 1024 ** 3 if self.device.type == "cuda" else 0.0
        info = (self.worker_id, "worker_performance", {
            "avg_loop_time": sum(self._loop_times) / len(self._loop_times),
            "state_size": self.batch_size,
            "state_offset_details": dict(zip(names, offsets[:5])),
            "task_details": {"decode_count": len(groups[StateCategory.FORWARD_ONE_DECODE]),
                             "one_prefill_count": len(groups[StateCategory.FORWARD_ONE_PREFILL]),
                             "seq_prefill_count": len(groups[StateCategory.FORWARD_SEQ])},
            "max_allocated_memory_GB": mem})
        try:
            self.perf_queue.put_nowait(info)
        except queue.Full:
            pass

    def start(self) -> None:
        self._init_worker()
        while not self._stop:
            if self.step() == 0:
                time.sleep(0.001)

    def busy(self) -> bool:
        return self._inflight is not None or any(s.category != StateCategory.EMPTY for s in self.slots)


# ------------------------------------------------------------------------------------------------ asyncio side
class ThreadSafeAsyncQueue:
    """Worker threads put, the event loop gets (engine_core.py:30-57)."""

    def __init__(self, event_loop: asyncio.AbstractEventLoop):
        self.loop = event_loop
        self.queue: asyncio.Queue = asyncio.Queue()

    def put_nowait(self, item) -> None:
        self.loop.call_soon_threadsafe(self.queue.put_nowait, item)


class AsyncEngineCompletion:
    """Async iterator over one request's events (interface.py:31-142): yields ``("token", id, text, logits|None)`` and ends
    after ``("finish", FinishReason, task)``."""

    def __init__(self, task: Task, task_queue: "queue.Queue", channel: ThreadSafeAsyncQueue):
        self.task, self._task_queue, self._channel = task, task_queue, channel
        self.finish_reason: Optional[FinishReason] = None
        self._started = False

    def start(self) -> "AsyncEngineCompletion":
        if not self._started:
            self._started = True
            self._task_queue.put(self.task)                                      # interface.py:95-98
        return self

    def __aiter__(self):
        return self.start()

    async def __anext__(self):
        if self.finish_reason is not None:
            raise StopAsyncIteration
        ev = await self._channel.queue.get()
        if ev[0] == "finish":
            self.finish_reason = ev[1]
            raise StopAsyncIteration
        return ev

    async def get_full_completion(self) -> List[int]:
        toks = []
        async for ev in self:
            if ev[0] == "token":
                toks.append(ev[1])
        return toks

    def abort(self) -> None:
        self.task.task_event_queue.put("abort")                                  # interface.py:140-142


class CachedState(list):
    """A cached single-request state ([(L,2,1,C), (L,1,H,64,64), (1,)] device tensors) with the CUDA event after which the
    copy is complete; `Worker._fill_task_pool` waits on it before injecting (possibly on another GPU's stream)."""
    ready: Any = None


class CachedCompletion:
    """What ``chirrup/web_service/app.py`` does around ``engine_core.completion`` when ``use_state_cache`` is set, as one
    object (the HTTP layer itself is out of scope): ``check_and_wait_prefill`` before submitting (``app.py:187-190``), the
    request runs from the cached state on the uncached remainder (``:202-214``), and the worker's ``("cache_prefill", ...)``
    event feeds the cache and wakes identical requests parked on it (``:293-298``, ``:367-371``).  One deliberate deviation:
    the new entry is keyed by the FULL consumed prompt (cached prefix + the tokens this request prefilled); the reference keys
    it by the latter alone (``app.py:295-297``), which is a different string whenever the request itself started from a hit."""

    def __init__(self, engine: "AsyncEngineCore", cache, prefill_tokens: Sequence[int], cache_prefill_padding: int = 0,
                 use_state_cache: bool = True, cache_prefill: bool = True, **kw):
        self.engine, self.cache, self.tokens = engine, cache, list(prefill_tokens)
        self.padding, self.use_state_cache, self.cache_prefill, self.kw = cache_prefill_padding, use_state_cache, cache_prefill, kw
        self.inner: Optional[AsyncEngineCompletion] = None
        self.cached_token_len = 0

    async def _submit(self):
        rest, state, n = self.tokens, None, 0
        if self.use_state_cache:
            rest, state, n = await self.cache.check_and_wait_prefill(self.tokens, self.padding)
        self.cached_token_len = n
        self.inner = self.engine.completion(prefill_tokens=list(rest), state=state, cache_prefill=self.cache_prefill,
                                            cache_prefill_padding=self.padding, **self.kw).start()

    def __aiter__(self):
        return self

    async def __anext__(self):
        if self.inner is None:
            await self._submit()
        while True:
            ev = await self.inner.__anext__()
            if ev[0] != "cache_prefill":
                return ev
            st = CachedState(ev[1]["state"])
            st.ready = ev[1].get("ready")
            key = self.cache.cache(tuple(self.tokens[:self.cached_token_len]) + tuple(ev[1]["prefilled_tokens"]), st)
            await self.cache.awake_hang_up_prefills(key)

    @property
    def finish_reason(self):
        return self.inner.finish_reason if self.inner is not None else None

    async def get_full_completion(self) -> List[int]:
        return [ev[1] async for ev in self if ev[0] == "token"]

    def abort(self) -> None:
        if self.inner is not None:
            self.inner.abort()


class AsyncEngineCore:
    """engine_core.py:60-332: N worker threads (worker k on GPU k) pulling from one shared task queue."""

    def __init__(self):
        self.task_queue: "queue.Queue" = queue.Queue()
        self.master_event_queue: "queue.Queue" = queue.Queue()
        self.workers: List[Worker] = []
        self.threads: List[threading.Thread] = []
        self.is_initialized = False
        self.is_shutdown = False
        self.tokenizer = IdTokenizer()
        self.event_loop: Optional[asyncio.AbstractEventLoop] = None
        self.perf_queue: "queue.Queue" = queue.Queue(maxsize=800)

    async def iter_worker_performance(self, timeout: float = 1.0):
        """engine_core.py:290-325: async generator over the workers' ("worker_performance") telemetry, as dicts with the
        reference's WorkerPerformanceInfo fields."""
        while not self.is_shutdown:
            try:
                worker_id, kind, payload = self.perf_queue.get_nowait()
            except queue.Empty:
                await asyncio.sleep(min(timeout, 0.05))
                continue
            if kind == "worker_performance":
                yield dict(worker_id=worker_id, **payload)

    def init(self, worker_num: int, model_config: ModelLoadConfig, batch_size: int = 32, tokenizer=None,
             load_timeout_s: float = 300.0) -> "asyncio.Task":
        self.event_loop = asyncio.get_event_loop()
        if tokenizer is not None:
            self.tokenizer = tokenizer
        for k in range(worker_num):                                              # engine_core.py:135-153
            w = Worker(f"worker_{k}", [k], model_config, self.task_queue, self.master_event_queue, batch_size=batch_size,
                       tokenizer=self.tokenizer)
            w.perf_queue = self.perf_queue
            th = threading.Thread(target=w.start, daemon=True, name=w.worker_id)
            self.workers.append(w)
            self.threads.append(th)
            th.start()

        async def wait_for_workers_loaded():                                     # engine_core.py:156-173
            loaded, t0 = 0, time.time()
            while loaded < worker_num:
                try:
                    ev = self.master_event_queue.get_nowait()
                    if ev[0] == "worker_loaded":
                        loaded += 1
                except queue.Empty:
                    if time.time() - t0 > load_timeout_s:
                        raise TimeoutError("workers did not load in time")
                    if any(not th.is_alive() for th in self.threads):
                        raise RuntimeError("a worker thread died while loading")   # the reference would wait out the timeout
                    await asyncio.sleep(0.05)
            self.is_initialized = True

        return asyncio.ensure_future(wait_for_workers_loaded())

    def completion(self, prompt_str: str = "", prefill_tokens: Optional[List[int]] = None, state=None, **kw) -> AsyncEngineCompletion:
        """engine_core.py:181-267 (same keyword names as ``Task``)."""
        assert not (state is not None and prefill_tokens is None), "prefill_tokens cannot be None when state is not None"
        if not self.is_initialized:
            raise RuntimeError("Engine not initialized")
        if self.is_shutdown:
            raise RuntimeError("Engine has been shutdown")
        if not prefill_tokens:
            prefill_tokens = self.tokenizer.encode(prompt_str)
        channel = ThreadSafeAsyncQueue(self.event_loop)
        task = Task(prompt_str=prompt_str, prefill_tokens=list(prefill_tokens), state=state, output_queue=channel, **kw)
        return AsyncEngineCompletion(task, self.task_queue, channel)

    def shutdown(self) -> None:
        """One shutdown event PER worker (the reference enqueues a single one, so only one of N workers exits cleanly —
        engine_core.py:279-283)."""
        self.is_shutdown = True
        for w in self.workers:
            w.worker_event_queue.put("shutdown")
        for th in self.threads:
            th.join(timeout=10)
