"""CPU tier: the process-per-worker engine (chirrup_b200/engine_mp.py) with the CPU model double — N worker PROCESSES on one
shared task queue behind the AsyncEngineCore API (the reference's deployment shape, chirrup/engine_core.py:71, :135-153, with
processes instead of free-threaded threads)."""
import asyncio

import pytest

from chirrup_b200.engine import FinishReason, ModelLoadConfig
from chirrup_b200.engine_mp import MPEngineCore
from toy_model import make_toy_model


def test_two_worker_processes_share_one_queue():
    async def main():
        eng = MPEngineCore()
        await eng.init(2, ModelLoadConfig(), batch_size=4, model_factory=make_toy_model, load_timeout_s=120)
        comps = [eng.completion(prefill_tokens=[i + 1] + list(range(2, 2 + (i % 5) * 7)), top_k=1, max_tokens=5, stop_tokens=[])
                 for i in range(24)]
        outs = await asyncio.gather(*[c.get_full_completion() for c in comps])
        reasons = [c.finish_reason for c in comps]
        # abort path
        long = eng.completion(prefill_tokens=[3], top_k=1, max_tokens=10 ** 9, stop_tokens=[]).start()
        got = []
        async for ev in long:
            got.append(ev[1])
            if len(got) == 5:
                long.abort()
        # telemetry
        perf = None
        c2 = [eng.completion(prefill_tokens=[5], top_k=1, max_tokens=4000, stop_tokens=[]) for _ in range(4)]
        run = asyncio.gather(*[c.get_full_completion() for c in c2])
        async for rec in eng.iter_worker_performance():
            perf = rec
            break
        await run
        eng.shutdown()
        return outs, reasons, got, long.finish_reason, perf, eng.worker_steps

    outs, reasons, got, abort_reason, perf, steps = asyncio.run(main())
    for i, toks in enumerate(outs):
        last = ([i + 1] + list(range(2, 2 + (i % 5) * 7)))[-1]
        assert toks == [(last + 1 + k) % 50 for k in range(5)]
    assert all(r == FinishReason.LENGTH for r in reasons)
    assert got[:5] == [4, 5, 6, 7, 8] and abort_reason == FinishReason.ABORT
    assert perf is not None and perf["worker_id"] in ("worker_0", "worker_1") and "task_details" in perf
    assert set(steps) == {"worker_0", "worker_1"} and all(v["decode_tokens"] > 0 for v in steps.values())   # both took work


def test_stop_token_and_unknown_argument():
    async def main():
        eng = MPEngineCore()
        await eng.init(1, ModelLoadConfig(), batch_size=2, model_factory=make_toy_model, load_timeout_s=120)
        with pytest.raises(TypeError):
            eng.completion(prefill_tokens=[1], no_such_option=1)
        c = eng.completion(prefill_tokens=[5], top_k=1, max_tokens=50, stop_tokens=[9])
        toks = await c.get_full_completion()
        eng.shutdown()
        return toks, c.finish_reason, c.generated_tokens

    toks, reason, gen = asyncio.run(main())
    assert toks == [6, 7, 8] == gen and reason == FinishReason.STOP
