"""``AsyncEngineCore`` with ONE PROCESS PER GPU behind the same API (SURVEY §7 "Host loop scaling to 8 GPUs", §8(f)-4).

The reference runs its N workers as threads of one process on one ``queue.Queue`` (``/root/reference/chirrup/engine_core.py:71``,
``:135-153``) and needs free-threaded CPython 3.14 for that to scale (``README.md:52``, ``:93-95``).  On a GIL interpreter the
per-slot Python of N worker loops serialises (measured here: 1.84x on 2 GPUs with threads).  Same contract, different plumbing:

* one shared task queue (``multiprocessing`` queue): whichever worker has a free slot takes the next request — request-level
  data parallelism, no collective, exactly the reference's partitioning;
* each worker process runs the unchanged ``engine.Worker`` loop on its own GPU with its own interpreter;
* tokens come back in ONE message per worker step (every event of that step), and the parent delivers a whole batch to the
  per-request asyncio queues with a single ``call_soon_threadsafe`` — the reference pays one per token per request
  (``engine_core.py:30-57``, ``worker.py:492-501``);
* abort / shutdown travel on a per-worker control queue (one shutdown event per worker: SURVEY §5 notes the reference's
  single event stops only one of N).

``MPEngineCore.completion`` takes the same keywords as ``AsyncEngineCore.completion``; ``state`` tensors and the states of
``("cache_prefill", ...)`` events cross the process boundary through host memory (the thread engine keeps them on the GPU).
"""
from __future__ import annotations

import asyncio
import queue
import threading
import time
from typing import Any, Callable, Dict, List, Optional, Sequence

import torch
import torch.multiprocessing as mp

from .engine import (AsyncEngineCompletion, FinishReason, IdTokenizer, ModelLoadConfig, RequestStatus, Task, Worker)

_TASK_FIELDS = ("prompt_str", "prefill_tokens", "priority", "temperature", "top_p", "top_k", "presence_penalty",
                "frequency_penalty", "penalty_decay", "max_tokens", "stop_tokens", "forbidden_tokens", "cache_prefill",
                "cache_prefill_padding", "return_logits")


class _StepSink:
    """output_queue of every task of one worker process: collects the events of the current step.  Plain tokens (the bulk of
    the traffic) travel as two parallel integer lists; everything else as tuples."""

    def __init__(self):
        self.events: List[tuple] = []
        self.tok_ids: List[int] = []
        self.toks: List[int] = []

    def bind(self, task_id: int) -> "_Bound":
        return _Bound(self, task_id)

    def take(self):
        if not (self.events or self.tok_ids):
            return None
        batch = (self.tok_ids, self.toks, self.events)
        self.tok_ids, self.toks, self.events = [], [], []
        return batch


class _Bound:
    def __init__(self, sink: _StepSink, task_id: int):
        self.sink, self.task_id = sink, task_id

    def put_nowait(self, ev) -> None:
        kind = ev[0]
        if kind == "token":
            if ev[2] or (len(ev) > 3 and ev[3] is not None):          # decoded text or raw logits ride along
                lg = ev[3].float().cpu() if len(ev) > 3 and ev[3] is not None else None
                self.sink.events.append((self.task_id, "token", ev[1], ev[2], lg))
            else:
                self.sink.tok_ids.append(self.task_id)
                self.sink.toks.append(ev[1])
        elif kind == "finish":
            self.sink.events.append((self.task_id, "finish", int(ev[1]), list(ev[2].generated_tokens)))
        elif kind == "cache_prefill":
            p = ev[1]
            if p.get("ready") is not None:
                p["ready"].synchronize()
            self.sink.events.append((self.task_id, "cache_prefill",
                                     {"state": [t.cpu() for t in p["state"]], "prefilled_tokens": p["prefilled_tokens"]}))
        else:
            self.sink.events.append((self.task_id,) + tuple(ev))


class _TaskAdapter:
    """The worker's ``task_queue``: pulls request dicts off the shared multiprocessing queue and rebuilds ``Task`` objects."""

    def __init__(self, mpq, sink: _StepSink, by_id: Dict[int, Task]):
        self.mpq, self.sink, self.by_id = mpq, sink, by_id

    def get_nowait(self) -> Task:
        d = self.mpq.get_nowait()                     # raises queue.Empty
        t = Task(output_queue=self.sink.bind(d["task_id"]), state=d.get("state"), task_id=str(d["task_id"]),
                 **{k: d[k] for k in _TASK_FIELDS if k in d})
        self.by_id[d["task_id"]] = t
        return t

    def empty(self) -> bool:
        return self.mpq.empty()

    def put(self, t) -> None:                          # not used by the worker; kept for interface completeness
        raise NotImplementedError


def _worker_main(worker_id: str, gpu_id: int, model_config: ModelLoadConfig, batch_size: int, task_q, result_q, ctrl_q,
                 model_factory: Optional[Callable], options: Dict[str, Any]) -> None:
    try:
        sink = _StepSink()
        by_id: Dict[int, Task] = {}
        model = model_factory() if model_factory is not None else None
        w = Worker(worker_id, [gpu_id], model_config, _TaskAdapter(task_q, sink, by_id), batch_size=batch_size, model=model)
        for k, v in options.items():
            setattr(w, k, v)
        perf = queue.Queue(maxsize=4)
        w.perf_queue = perf
        w._init_worker()
        result_q.put(("worker_loaded", worker_id))
        last_perf = 0.0
        while True:
            try:
                while True:
                    msg = ctrl_q.get_nowait()
                    if msg[0] == "shutdown":
                        stats = dict(w.steps)
                        stats["graph_hits"] = getattr(w.model, "graph_hits", 0)      # decode steps replayed from a step graph
                        stats["graph_misses"] = getattr(w.model, "graph_misses", 0)  # ... launched eagerly (or captured)
                        result_q.put(("worker_exit", worker_id, stats))
                        return
                    if msg[0] == "abort" and msg[1] in by_id:
                        by_id[msg[1]].task_event_queue.put("abort")
            except queue.Empty:
                pass
            n = w.step()
            batch = sink.take()
            if batch is not None:
                for ev in batch[2]:
                    if ev[1] == "finish":
                        by_id.pop(ev[0], None)
                result_q.put(("events", worker_id, batch))
            latest = None
            try:
                while True:
                    latest = perf.get_nowait()
            except queue.Empty:
                pass
            now = time.perf_counter()
            if latest is not None and now - last_perf > 0.5:        # telemetry at 2 Hz, not one message per loop
                result_q.put(("perf", latest))
                last_perf = now
            if n == 0:
                time.sleep(0.0005)
    except BaseException as e:  # noqa: BLE001 — surface the failure to the parent instead of dying silently
        import traceback

        result_q.put(("worker_error", worker_id, f"{e!r}\n{traceback.format_exc()}"))


class MPCompletion(AsyncEngineCompletion):
    """``AsyncEngineCompletion`` (interface.py:31-142) over the process boundary."""

    def __init__(self, engine: "MPEngineCore", task_id: int, payload: Dict[str, Any], channel: "_Channel"):
        self.engine, self.task_id, self.payload = engine, task_id, payload
        self._channel = channel
        self._queue = channel.queue
        self.finish_reason: Optional[FinishReason] = None
        self.generated_tokens: List[int] = []
        self._started = False

    async def get_full_completion(self) -> List[int]:
        """Bulk form: tokens are collected on the delivery side and the caller is woken once, by the finish event — no
        per-token task switch (the per-token async iteration costs a few microseconds of event-loop time per token, which
        at 8 GPUs x 37 k tok/s is the whole budget of one interpreter)."""
        if not self._started:
            self._channel.stream = False
        toks = [ev[1] async for ev in self if ev[0] == "token"]
        return self._channel.tokens + toks if not self._channel.stream else toks

    def start(self) -> "MPCompletion":
        if not self._started:
            self._started = True
            self.engine.task_queue.put(self.payload)
        return self

    async def __anext__(self):
        if self.finish_reason is not None:
            raise StopAsyncIteration
        ev = await self._queue.get()
        if ev[0] == "finish":
            self.finish_reason = FinishReason(ev[1])
            self.generated_tokens = ev[2]
            self.engine._channels.pop(self.task_id, None)
            raise StopAsyncIteration
        if ev[0] == "error":
            raise RuntimeError(ev[1])
        return ev

    def abort(self) -> None:
        for q in self.engine._ctrl:
            q.put(("abort", self.task_id))


class _Channel:
    __slots__ = ("queue", "tokens", "stream")

    def __init__(self):
        self.queue: "asyncio.Queue" = asyncio.Queue()
        self.tokens: List[int] = []
        self.stream = True


class MPEngineCore:
    """Process-per-GPU engine: worker k in its own process on GPU ``gpu_ids[k]``, one shared task queue."""

    def __init__(self, start_method: str = "spawn"):
        self._ctx = mp.get_context(start_method)
        self.task_queue = self._ctx.Queue()
        self._results = self._ctx.Queue()
        self._ctrl: List[Any] = []
        self.procs: List[Any] = []
        self._channels: Dict[int, _Channel] = {}
        self._next_id = 0
        self.is_initialized = False
        self.is_shutdown = False
        self.tokenizer = IdTokenizer()
        self.event_loop: Optional[asyncio.AbstractEventLoop] = None
        self._loaded = 0
        self._errors: List[str] = []
        self._perf: "queue.Queue" = queue.Queue(maxsize=800)
        self.worker_steps: Dict[str, Dict[str, int]] = {}
        self._dispatcher: Optional[threading.Thread] = None
        self.tokens_delivered = 0

    # ---------------------------------------------------------------- parent-side delivery
    def _deliver(self, batch) -> None:
        """Runs on the event loop: every event of one worker step, one call."""
        tok_ids, toks, events = batch
        chans = self._channels
        for tid, tok in zip(tok_ids, toks):
            ch = chans.get(tid)
            if ch is not None:
                if ch.stream:
                    ch.queue.put_nowait(("token", tok, "", None))
                else:
                    ch.tokens.append(tok)
        self.tokens_delivered += len(toks)
        for ev in events:
            ch = chans.get(ev[0])
            if ch is not None:
                ch.queue.put_nowait(ev[1:])
            if ev[1] == "token":
                self.tokens_delivered += 1

    def _dispatch_loop(self) -> None:
        while True:
            msg = self._results.get()
            kind = msg[0]
            if kind == "events":
                self.event_loop.call_soon_threadsafe(self._deliver, msg[2])
            elif kind == "worker_loaded":
                self._loaded += 1
            elif kind == "perf":
                try:
                    self._perf.put_nowait(msg[1])
                except queue.Full:
                    pass
            elif kind == "worker_error":
                self._errors.append(f"{msg[1]}: {msg[2]}")
                err = [(tid, "error", msg[2]) for tid in list(self._channels)]
                self.event_loop.call_soon_threadsafe(self._deliver, ([], [], err))
            elif kind == "worker_exit":
                self.worker_steps[msg[1]] = msg[2]
            elif kind == "stop":
                return

    # ---------------------------------------------------------------- AsyncEngineCore surface
    def init(self, worker_num: int, model_config: ModelLoadConfig, batch_size: int = 32, tokenizer=None,
             load_timeout_s: float = 600.0, gpu_ids: Optional[Sequence[int]] = None, model_factory: Optional[Callable] = None,
             worker_options: Optional[Dict[str, Any]] = None) -> "asyncio.Task":
        self.event_loop = asyncio.get_event_loop()
        if tokenizer is not None:
            self.tokenizer = tokenizer
        gpu_ids = list(gpu_ids) if gpu_ids is not None else list(range(worker_num))
        for k in range(worker_num):
            ctrl = self._ctx.Queue()
            p = self._ctx.Process(target=_worker_main, name=f"worker_{k}", daemon=True,
                                  args=(f"worker_{k}", gpu_ids[k], model_config, batch_size, self.task_queue, self._results, ctrl,
                                        model_factory, dict(worker_options or {})))
            p.start()
            self._ctrl.append(ctrl)
            self.procs.append(p)
        self._dispatcher = threading.Thread(target=self._dispatch_loop, daemon=True, name="mp-engine-dispatch")
        self._dispatcher.start()

        async def wait_for_workers_loaded():
            t0 = time.time()
            while self._loaded < worker_num:
                if self._errors:
                    raise RuntimeError("worker failed while loading: " + self._errors[0])
                if time.time() - t0 > load_timeout_s:
                    raise TimeoutError("workers did not load in time")
                if any(not p.is_alive() for p in self.procs):
                    raise RuntimeError("a worker process died while loading")
                await asyncio.sleep(0.05)
            self.is_initialized = True

        return asyncio.ensure_future(wait_for_workers_loaded())

    def completion(self, prompt_str: str = "", prefill_tokens: Optional[List[int]] = None, state=None, **kw) -> MPCompletion:
        assert not (state is not None and prefill_tokens is None), "prefill_tokens cannot be None when state is not None"
        if not self.is_initialized:
            raise RuntimeError("Engine not initialized")
        if self.is_shutdown:
            raise RuntimeError("Engine has been shutdown")
        if not prefill_tokens:
            prefill_tokens = self.tokenizer.encode(prompt_str)
        unknown = set(kw) - set(_TASK_FIELDS)
        if unknown:
            raise TypeError(f"unknown completion arguments: {sorted(unknown)}")
        tid = self._next_id
        self._next_id += 1
        payload = dict(kw, task_id=tid, prompt_str=prompt_str, prefill_tokens=list(prefill_tokens))
        if state is not None:
            payload["state"] = [t.detach().cpu() for t in state]
        ch = _Channel()
        self._channels[tid] = ch
        return MPCompletion(self, tid, payload, ch)

    async def iter_worker_performance(self, timeout: float = 1.0):
        while not self.is_shutdown:
            try:
                worker_id, kind, payload = self._perf.get_nowait()
            except queue.Empty:
                await asyncio.sleep(min(timeout, 0.05))
                continue
            if kind == "worker_performance":
                yield dict(worker_id=worker_id, **payload)

    def shutdown(self, timeout_s: float = 15.0) -> None:
        if self.is_shutdown:
            return
        self.is_shutdown = True
        for q in self._ctrl:
            q.put(("shutdown",))
        t0 = time.time()
        for p in self.procs:
            p.join(timeout=max(0.1, timeout_s - (time.time() - t0)))
            if p.is_alive():
                p.terminate()
        time.sleep(0.05)                              # let the dispatcher book the worker_exit records
        self._results.put(("stop",))
        if self._dispatcher is not None:
            self._dispatcher.join(timeout=5)
