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


from toy_model import ToyModel  # noqa: E402  (CPU stand-in with the model's calling convention)


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


def _run_until_idle(w, n=400):
    for _ in range(n):
        w.step()
        if not w.busy() and w.task_queue.empty():
            return
    raise AssertionError("worker did not drain")


@pytest.mark.parametrize("plen,pad", [(40, 0), (40, 3), (7, 2), (131, 7), (12, 1), (3, 0)])
def test_cache_prefill_event_fires_at_the_reference_trigger_point(plen, pad):
    """chirrup/worker.py:411-435, :457-476: exactly one ("cache_prefill", ...) event, carrying the slot state after
    prompt[:len - max(padding, 1)] and that token tuple; sequence forwards never run into the padding (worker.py:746-754)."""
    w = make_worker(bsz=4)
    s = Sink()
    prompt = [(7 * i) % 40 + 1 for i in range(plen)]
    w.task_queue.put(Task(prefill_tokens=list(prompt), output_queue=s, top_k=1, max_tokens=2, stop_tokens=[],
                          cache_prefill=True, cache_prefill_padding=pad))
    _run_until_idle(w)
    evs = [e for e in s.items if e[0] == "cache_prefill"]
    n_key = plen - max(pad, 1)
    if n_key <= 0:
        assert not evs
        return
    assert len(evs) == 1
    payload = evs[0][1]
    assert payload["prefilled_tokens"] == tuple(prompt[:n_key])
    # the ToyModel leaves a trace of what the slot consumed: elapsed == tokens consumed, state0[0,0,.,0] == their sum
    assert int(payload["state"][2][0]) == n_key
    assert float(payload["state"][0][0, 0, 0, 0]) == float(sum(prompt[:n_key]))
    toks = [e[1] for e in s.items if e[0] == "token"]
    assert toks == [(prompt[-1] + 1) % 50, (prompt[-1] + 2) % 50]          # the request itself is unaffected


def test_cached_completion_prefix_hit_and_dedup_through_the_engine():
    """app.py:187-214, :293-298 at the engine level: request 1 prefills and caches; an identical concurrent request waits and is
    then served from the cache; a later request sharing the prefix starts from the cached state — all produce the tokens of a
    cold run, and the worker consumed only the uncached remainder."""
    from chirrup_b200.engine import CachedCompletion
    from chirrup_b200.state_cache import PrefixStateCache

    async def main():
        eng = AsyncEngineCore()
        eng.event_loop = asyncio.get_event_loop()
        model = ToyModel()
        w = Worker("w0", [0], ModelLoadConfig(), eng.task_queue, eng.master_event_queue, batch_size=4, model=model)
        import threading

        th = threading.Thread(target=w.start, daemon=True)
        eng.workers.append(w)
        eng.threads.append(th)
        th.start()
        eng.is_initialized = True
        cache = PrefixStateCache(8)
        prompt = [(3 * i) % 40 + 1 for i in range(30)]
        kw = dict(top_k=1, max_tokens=3, stop_tokens=[])
        a = CachedCompletion(eng, cache, prompt, cache_prefill_padding=2, **kw)
        b = CachedCompletion(eng, cache, prompt, cache_prefill_padding=2, **kw)       # identical, concurrent: must wait for a
        ta, tb = await asyncio.gather(a.get_full_completion(), b.get_full_completion())
        assert ta == tb == [(prompt[-1] + 1 + i) % 50 for i in range(3)]
        assert a.cached_token_len == 0 and b.cached_token_len == len(prompt) - 2
        assert tuple(prompt[:-2]) in cache
        c = CachedCompletion(eng, cache, prompt + [9, 9, 9, 9], cache_prefill_padding=2, **kw)   # shares the cached prefix
        tc = await c.get_full_completion()
        assert tc == [(9 + 1 + i) % 50 for i in range(3)] and c.cached_token_len == len(prompt) - 2
        assert tuple((prompt + [9, 9, 9, 9])[:-2]) in cache                            # keyed by the FULL consumed prompt
        eng.shutdown()
        return model.calls

    calls = asyncio.run(main())
    consumed = sum(B * T for B, T, _ in calls)
    # cold: 30 prompt tokens + 2 decode steps after the first sampled token; b: 2 + 2; c: 6 + 2  (vs 30 + 34 + ... uncached)
    assert consumed <= (30 + 2) + (2 + 2) + (6 + 2) + 6


def test_worker_performance_telemetry():
    """chirrup/worker.py:858-884 -> engine_core.py:290-325: one ("worker_performance") record per busy loop with the
    reference's payload keys."""
    w = make_worker(bsz=8)
    w.perf_queue = queue.Queue(maxsize=100)
    w.task_queue.put(Task(prefill_tokens=list(range(1, 30)), output_queue=Sink(), top_k=1, max_tokens=3, stop_tokens=[]))
    w.task_queue.put(Task(prefill_tokens=[4], output_queue=Sink(), top_k=1, max_tokens=3, stop_tokens=[]))
    _run_until_idle(w)
    recs = []
    while not w.perf_queue.empty():
        recs.append(w.perf_queue.get_nowait())
    assert recs and all(r[0] == "w0" and r[1] == "worker_performance" for r in recs)
    p = recs[0][2]
    assert set(p) == {"avg_loop_time", "state_size", "state_offset_details", "task_details", "max_allocated_memory_GB"}
    assert set(p["task_details"]) == {"decode_count", "one_prefill_count", "seq_prefill_count"} and p["state_size"] == 8
    assert p["state_offset_details"]["decode_offset"][0] == 0 and p["avg_loop_time"] > 0
    assert any(r[2]["task_details"]["seq_prefill_count"] == 1 for r in recs)
