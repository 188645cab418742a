"""CPU tier: host logic of the continuous-batching driver (chirrup_b200/engine.py) with a CPU test double of the model —
the approach of the reference's tests/test_worker_state_category.py (model and tokenizer patched out), whose integer fixtures
(StateCategory order, offsets) are restated here and must match exactly."""
import asyncio
import queue

import pytest
import torch

from chirrup_b200.engine import (AsyncEngineCore, FinishReason, ModelLoadConfig, RequestStatus, StateCategory, Task, Worker,
                                 group_slots)

SC = StateCategory


def test_state_category_order():
    # reference tests/test_worker_state_category.py:28-47
    assert SC.FORWARD_ONE_DECODE < SC.FORWARD_ONE_PREFILL < SC.FORWARD_ONE_SUSPENDED < SC.FORWARD_SEQ < SC.FINISHED < SC.EMPTY
    assert list(sorted(SC)) == [SC.FORWARD_ONE_DECODE, SC.FORWARD_ONE_PREFILL, SC.FORWARD_ONE_SUSPENDED, SC.FORWARD_SEQ,
                                SC.FINISHED, SC.EMPTY]


def test_offsets_match_reference_fixture():
    # reference tests/test_worker_state_category.py:88-106: offsets order DECODE, PREFILL, SUSPENDED, SEQ, FINISHED, EMPTY
    lst = [SC.FORWARD_ONE_DECODE, SC.FORWARD_ONE_PREFILL, SC.FORWARD_ONE_PREFILL, SC.FORWARD_SEQ, SC.EMPTY, SC.EMPTY]
    groups, offsets = group_slots(lst)
    assert offsets == [(0, 1), (1, 3), (3, 3), (3, 4), (4, 4), (4, 6)]
    assert groups[SC.FORWARD_ONE_PREFILL] == [1, 2] and groups[SC.EMPTY] == [4, 5]


def test_grouping_is_the_sort_without_moving_rows():
    # reference :64-86: after its swaps decode rows come first, then prefill rows; here the same rows, by index
    lst = [SC.FORWARD_ONE_PREFILL, SC.FORWARD_ONE_DECODE, SC.FORWARD_ONE_PREFILL, SC.FORWARD_ONE_DECODE, SC.EMPTY]
    groups, offsets = group_slots(lst)
    order = [i for c in sorted(SC) for i in groups[c]]
    assert [lst[i] for i in order] == sorted(lst)
    assert offsets[0] == (0, 2) and offsets[1] == (2, 4) and offsets[5] == (4, 5)
    g = torch.Generator().manual_seed(0)
    for _ in range(50):
        cats = [SC(int(x)) for x in torch.randint(1, 7, (17,), generator=g)]
        groups, offsets = group_slots(cats)
        flat = [i for c in sorted(SC) for i in groups[c]]
        assert sorted(flat) == list(range(17)) and [cats[i] for i in flat] == sorted(cats)
        for c, (a, b) in zip(sorted(SC), offsets):
            assert b - a == cats.count(c)


class ToyModel:
    """CPU stand-in with the model's calling convention: next token = (last input + 1) mod V, elapsed += T; records every call."""

    def __init__(self, V=50, L=2, C=8):
        self.device = torch.device("cpu")
        self.V, self.L, self.C = V, L, C
        self.vocab_size = V
        self.calls = []

    def generate_zero_state(self, bsz):
        return [torch.zeros((self.L, 2, bsz, self.C), dtype=torch.float16), torch.zeros((self.L, bsz, 1, 64, 64), dtype=torch.float16),
                torch.zeros((bsz,), dtype=torch.int32)]

    def _tokens(self, idxs):
        return torch.tensor(idxs, dtype=torch.int32).reshape(-1)

    def _step(self, tokens, B, T, state, full_output, want_logits=True, slot_idx=None, out=None):
        self.calls.append((B, T, slot_idx.tolist() if slot_idx is not None else None))
        tk = tokens.view(B, T).long()
        state[2] += T
        rows = slot_idx.long()
        state[0][0, 0, rows, 0] += tk.sum(1).to(torch.float16)          # a visible trace of what each slot consumed
        if not want_logits:
            return None
        lg = torch.zeros((B, self.V), dtype=torch.float16)
        lg[torch.arange(B), (tk[:, -1] + 1) % self.V] = 5.0
        return lg


def make_worker(bsz=8, V=50):
    w = Worker("w0", [0], ModelLoadConfig(), queue.Queue(), batch_size=bsz, model=ToyModel(V))
    w._init_worker()
    return w


class Sink:
    def __init__(self):
        self.items = []

    def put_nowait(self, x):
        self.items.append(x)


def test_decode_prefill_schedule_and_outputs():
    w = make_worker(bsz=8)
    sinks = []
    prompts = [[3], [4, 5, 6], list(range(10, 45)), list(range(1, 13))]        # len 1, 3 (one-prefill), 35 and 12 (seq)
    for p in prompts:
        s = Sink()
        sinks.append(s)
        w.task_queue.put(Task(prefill_tokens=p, output_queue=s, top_k=1, max_tokens=4, stop_tokens=[]))
    for _ in range(200):
        w.step()
        if not w.busy() and w.task_queue.empty():
            break
    assert not w.busy()
    for p, s in zip(prompts, sinks):
        toks = [e[1] for e in s.items if e[0] == "token"]
        want = [(p[-1] + 1 + i) % 50 for i in range(4)]
        assert toks == want, (p, toks)
        assert s.items[-1][0] == "finish" and s.items[-1][1] == FinishReason.LENGTH
    # prefill chunks: never more than 100 tokens, every SEQ row the same T, and a slot drops to one-token prefill at <= 10 left
    seq_calls = [c for c in w.model.calls if c[1] > 1]
    assert seq_calls and all(c[1] <= 100 for c in seq_calls)
    assert w.steps["prefill_tokens"] > 0 and w.steps["decode_tokens"] == 16


def test_admission_limit_stop_tokens_abort_and_state_injection():
    w = make_worker(bsz=16)
    assert w.max_prefill_count == 2
    for i in range(6):
        w.task_queue.put(Task(prefill_tokens=[i + 1], output_queue=Sink(), top_k=1, max_tokens=50, stop_tokens=[9]))
    w.step()
    assert sum(c != SC.EMPTY for c in w.categories()) == 2                    # at most batch/8 admissions per loop
    for _ in range(100):
        w.step()
    assert not w.busy()                                                      # all hit the stop token 9 (or wrapped to it)
    # abort
    s = Sink()
    t = Task(prefill_tokens=[20], output_queue=s, top_k=1, max_tokens=10 ** 6, stop_tokens=[])
    w.task_queue.put(t)
    for _ in range(3):
        w.step()
    t.task_event_queue.put("abort")
    w.step()
    assert t.request_status == RequestStatus.FINISHED_ABORTED and s.items[-1][1] == FinishReason.ABORT
    # state injection: the injected elapsed counter continues
    st = [torch.zeros((2, 2, 1, 8), dtype=torch.float16), torch.zeros((2, 1, 1, 64, 64), dtype=torch.float16),
          torch.tensor([1000], dtype=torch.int32)]
    t2 = Task(prefill_tokens=[7], state=st, output_queue=Sink(), top_k=1, max_tokens=1, stop_tokens=[])
    w.task_queue.put(t2)
    w.step()
    assert 1001 in w.batch_state[2].tolist()


def test_async_engine_roundtrip_with_toy_model():
    async def main():
        eng = AsyncEngineCore()
        eng.event_loop = asyncio.get_event_loop()
        w = Worker("w0", [0], ModelLoadConfig(), eng.task_queue, eng.master_event_queue, batch_size=4, model=ToyModel())
        import threading

        th = threading.Thread(target=w.start, daemon=True)
        eng.workers.append(w)
        eng.threads.append(th)
        th.start()
        eng.is_initialized = True
        comps = [eng.completion(prefill_tokens=[i + 1], top_k=1, max_tokens=3, stop_tokens=[]) for i in range(6)]
        outs = await asyncio.gather(*[c.get_full_completion() for c in comps])
        eng.shutdown()
        return outs, [c.finish_reason for c in comps]

    outs, reasons = asyncio.run(main())
    assert outs == [[(i + 2 + k) % 50 for k in range(3)] for i in range(6)]
    assert all(r == FinishReason.LENGTH for r in reasons)


def test_stop_token_is_neither_recorded_nor_streamed():
    """chirrup/worker.py:488-490: on a stop token the request finishes WITHOUT appending it to generated_tokens or emitting a
    token event; it does not count toward max_tokens either."""
    w = make_worker(bsz=4)
    s = Sink()
    t = Task(prefill_tokens=[5], output_queue=s, top_k=1, max_tokens=4, stop_tokens=[9])
    w.task_queue.put(t)
    for _ in range(50):
        w.step()
        if not w.busy():
            break
    assert t.generated_tokens == [6, 7, 8] and t.request_status == RequestStatus.FINISHED_STOPPED
    assert [e[1] for e in s.items if e[0] == "token"] == [6, 7, 8]
    assert s.items[-1][0] == "finish" and s.items[-1][1] == FinishReason.STOP
    # the cap is on delivered tokens: with max_tokens = 3 the request ends by LENGTH before the stop token is ever sampled
    s2 = Sink()
    t2 = Task(prefill_tokens=[5], output_queue=s2, top_k=1, max_tokens=3, stop_tokens=[9])
    w.task_queue.put(t2)
    for _ in range(50):
        w.step()
        if not w.busy():
            break
    assert t2.generated_tokens == [6, 7, 8] and s2.items[-1][1] == FinishReason.LENGTH


def test_greedy_classification_follows_the_reference_worker():
    assert Task(top_k=1).is_greedy() and Task(top_p=0.0).is_greedy()
    assert not Task(temperature=0.0).is_greedy()          # worker.py:611-613: temperature <= 0 becomes 1.0 and still samples
