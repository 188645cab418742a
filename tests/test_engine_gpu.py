"""GPU: the continuous-batching driver on the real model; slot movement; the reference op names on our kernels."""
import asyncio
import queue

import pytest
import torch

from chirrup_b200.synth import GEOMETRIES, make_checkpoint

pytestmark = pytest.mark.gpu


def _model(geom_name="small", seed=8):
    from chirrup_b200.rwkv7 import RWKV_x070, make_args

    geom = GEOMETRIES[geom_name]
    raw = make_checkpoint(geom, seed=seed, recipe="trained_like")
    return geom, RWKV_x070(make_args(), state_dict=raw)


def _greedy_alone(m, prompt, n):
    from chirrup_b200 import ops

    st = m.generate_zero_state(1)
    lg = m.forward_seq_batch([prompt], st)
    out = []
    for _ in range(n):
        t = int(ops.argmax(lg)[0])
        out.append(t)
        lg = m.forward_seq_batch([[t]], st)
    return out


def test_worker_continuous_batching_matches_isolated_greedy():
    """Requests of mixed prompt lengths (1 .. 230 tokens: one-token prefill, chunked prefill <= 100, decode) admitted over time
    into 6 slots produce exactly the tokens each request produces alone (slot indirection, no physical swaps)."""
    from chirrup_b200.engine import ModelLoadConfig, Task, Worker

    geom, m = _model()
    w = Worker("w0", [0], ModelLoadConfig(), queue.Queue(), batch_size=6, model=m)
    w._init_worker()
    g = torch.Generator().manual_seed(0)
    lens = [1, 3, 9, 11, 40, 150, 230, 12, 2, 64]
    prompts = [torch.randint(1, geom.vocab_size, (n,), generator=g).tolist() for n in lens]

    class Sink:
        def __init__(self):
            self.items = []

        def put_nowait(self, x):
            self.items.append(x)

    sinks = [Sink() for _ in prompts]
    for p, s in zip(prompts, sinks):
        w.task_queue.put(Task(prefill_tokens=p, output_queue=s, top_k=1, max_tokens=12, stop_tokens=[], presence_penalty=0.0,
                              frequency_penalty=0.0))
    for _ in range(2000):
        w.step()
        if not w.busy() and w.task_queue.empty():
            break
    assert not w.busy()
    for p, s in zip(prompts, sinks):
        got = [e[1] for e in s.items if e[0] == "token"]
        assert got == _greedy_alone(m, p, 12), f"prompt len {len(p)}"
    assert w.steps["prefill"] > 0 and max(c for c in [w.max_forward_seq_len_per_forward]) == 100


def test_async_engine_core_end_to_end():
    from chirrup_b200.engine import AsyncEngineCore, FinishReason, ModelLoadConfig

    geom = GEOMETRIES["tiny"]
    raw = make_checkpoint(geom, seed=1, recipe="trained_like")

    async def main():
        eng = AsyncEngineCore()
        await eng.init(worker_num=1, model_config=ModelLoadConfig(state_dict=raw, vocab_size=geom.vocab_size), batch_size=8)
        comps = [eng.completion(prefill_tokens=[5 + i, 9, 11], top_k=1, max_tokens=6, stop_tokens=[], presence_penalty=0.0,
                                frequency_penalty=0.0) for i in range(12)]
        outs = await asyncio.gather(*[c.get_full_completion() for c in comps])
        eng.shutdown()
        return outs, [c.finish_reason for c in comps], eng.workers[0].model

    outs, reasons, m = asyncio.run(main())
    assert all(len(o) == 6 for o in outs) and all(r == FinishReason.LENGTH for r in reasons)
    for i, o in enumerate(outs):
        assert o == _greedy_alone(m, [5 + i, 9, 11], 6)


def test_slot_extract_inject_migrate_bit_exact():
    from chirrup_b200 import state_move

    geom, m = _model("tiny", 3)
    a = m.generate_zero_state(5)
    m.forward_seq_batch(torch.randint(1, geom.vocab_size, (5, 7), generator=torch.Generator().manual_seed(1)).tolist(), a)
    packed = state_move.extract_slot(a, 3)
    assert packed[0].shape[2] == 1 and packed[1].shape[1] == 1 and int(packed[2][0]) == 7
    b = m.generate_zero_state(4)
    state_move.inject_slot(b, 1, packed)
    moved = state_move.migrate_slot(a, 2, b, 0)
    torch.cuda.synchronize()
    assert moved == state_move.slot_bytes(a)
    assert torch.equal(b[0][:, :, 1], a[0][:, :, 3]) and torch.equal(b[1][:, 1], a[1][:, 3]) and int(b[2][1]) == 7
    assert torch.equal(b[0][:, :, 0], a[0][:, :, 2]) and torch.equal(b[1][:, 0], a[1][:, 2])
    # a request continued in its new slot gives the same logits as in the old one
    la = m._step(m._tokens([[11]]), 1, 1, [a[0], a[1], a[2][3:4].clone()], False, slot_idx=torch.tensor([3], dtype=torch.int32, device="cuda"))
    lb = m._step(m._tokens([[11]]), 1, 1, [b[0], b[1], b[2][1:2].clone()], False, slot_idx=torch.tensor([1], dtype=torch.int32, device="cuda"))
    assert torch.equal(la, lb)
    # host round trip (what Chirrup's state cache stores)
    cpu = state_move.extract_slot(a, 3, device="cpu", non_blocking=False)
    c = m.generate_zero_state(2)
    state_move.inject_slot(c, 0, cpu, non_blocking=False)
    assert torch.equal(c[1][:, 0], a[1][:, 3])


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_slot_migration_between_gpus_p2p():
    from chirrup_b200 import state_move

    geom, m = _model("tiny", 3)
    a = m.generate_zero_state(3)
    m.forward_seq_batch([[1, 2, 3], [4, 5, 6], [7, 8, 9]], a)
    b = [t.to("cuda:1").zero_() for t in m.generate_zero_state(2)]
    state_move.migrate_slot(a, 1, b, 0)
    torch.cuda.synchronize(0)
    torch.cuda.synchronize(1)
    assert torch.equal(b[1][:, 0].cpu(), a[1][:, 1].cpu()) and int(b[2][0]) == 3


def test_reference_op_names_run_on_our_kernels():
    """torch.ops.rwkv7_state_fwd_fp16.{forward_seq,forward_one,spmv_forward} (reference .cpp:21-25) registered by
    chirrup_b200.torch_ops give the same bits as the direct C-ABI path."""
    from chirrup_b200 import ops, torch_ops
    from wkv_inputs import make_wkv_inputs

    torch_ops.register()
    B, T, H = 3, 4, 2
    C = H * 64
    st, r, w, k, v, a, b = [t.cuda() for t in make_wkv_inputs(B, T, H, seed=2)]
    et = torch.tensor([0, 5, 9], dtype=torch.int32, device="cuda")
    s1, s2 = st.clone(), st.clone()
    y1 = ops.wkv7_forward_seq(s1, r, w, k, v, a, b, et)
    y2 = torch.empty_like(r)
    torch.ops.rwkv7_state_fwd_fp16.forward_seq(B, T, C, H, s2, r, w, k, v, a, b, y2, et)
    assert torch.equal(y1, y2) and torch.equal(s1, s2)
    vec = torch.relu(torch.randn(512, device="cuda") - 0.5).half()
    mat = (torch.randn(512, 256, device="cuda") * 0.1).half()
    out = torch.zeros(256, dtype=torch.float16, device="cuda")
    torch.ops.rwkv7_state_fwd_fp16.spmv_forward(512, 256, vec, mat, out)
    assert torch.equal(out, ops.spmv_forward(vec, mat))


def _run_worker(m, tasks, batch_size=4):
    from chirrup_b200.engine import ModelLoadConfig, Worker

    class Sink:
        def __init__(self):
            self.items = []

        def put_nowait(self, x):
            self.items.append(x)

    w = Worker("w0", [0], ModelLoadConfig(), queue.Queue(), batch_size=batch_size, model=m)
    w._init_worker()
    sinks = [Sink() for _ in tasks]
    for t, s in zip(tasks, sinks):
        t.output_queue = s
        w.task_queue.put(t)
    for _ in range(4000):
        w.step()
        if not w.busy() and w.task_queue.empty():
            break
    assert not w.busy()
    return [[e[1] for e in s.items if e[0] == "token"] for s in sinks], w


def test_worker_penalised_greedy_matches_reference_penalty_formula():
    """top_k = 1 requests WITH presence / frequency penalties, decay and forbidden tokens through the fused sampling kernel
    (device-resident per-slot penalty rows, slot indirection) against a per-request loop over the same model with the oracle's
    restatement of the reference's penalty block (worker.py:719-728) and update (worker.py:527-547): token-identical."""
    from chirrup_b200.engine import Task
    from oracle import sampler_oracle as SO

    geom, m = _model("small", seed=3)
    V = geom.vocab_size
    g = torch.Generator().manual_seed(1)
    specs = [(5, 0.5, 0.5, 0.996, []), (17, 2.0, 0.3, 0.9, [3, 9]), (2, 0.0, 1.5, 1.0, []), (130, 0.7, 0.0, 0.5, [11]),
             (40, 0.4, 0.6, 0.996, []), (9, 1.0, 1.0, 0.95, [])]
    prompts = [torch.randint(1, V, (n,), generator=g).tolist() for n, *_ in specs]
    n_new = 24
    tasks = [Task(prefill_tokens=p, top_k=1, max_tokens=n_new, stop_tokens=[], presence_penalty=pp, frequency_penalty=fp,
                  penalty_decay=dc, forbidden_tokens=fb) for p, (_, pp, fp, dc, fb) in zip(prompts, specs)]
    got, w = _run_worker(m, tasks, batch_size=4)
    for p, (_, pp, fp, dc, fb), toks in zip(prompts, specs, got):
        st = m.generate_zero_state(1)
        lg = m.forward_seq_batch([p], st)
        occ, alpha = torch.zeros(1, V), torch.zeros(1, V)
        want = []
        for _ in range(n_new):
            row = lg.cpu().clone()
            SO.penalise(row, occ, alpha, torch.tensor([[dc]]), torch.tensor([[fp]]), [fb])
            t = int(row.float().argmax(-1)[0])
            want.append(t)
            SO.update_penalty(occ, alpha, torch.tensor([[pp]]), [0], [t])
            lg = m.forward_seq_batch([[t]], st)
        assert toks == want, (len(p), toks[:8], want[:8])
        assert not set(fb) & set(toks)


def test_worker_sampled_requests_are_reproducible_and_respect_top_k():
    """Non-greedy requests: same worker seed -> same tokens (the draw is a pure function of seed, step and row); with top_k = 3
    and no penalties every drawn token is one of the three most likely ones of a replay of the same sequence."""
    from chirrup_b200.engine import Task

    geom, m = _model("small", seed=4)
    V = geom.vocab_size
    g = torch.Generator().manual_seed(2)
    prompts = [torch.randint(1, V, (n,), generator=g).tolist() for n in (4, 21, 7)]
    mk = lambda: [Task(prefill_tokens=p, temperature=1.3, top_p=0.95, top_k=3, max_tokens=16, stop_tokens=[], presence_penalty=0.0,
                       frequency_penalty=0.0) for p in prompts]
    a, _ = _run_worker(m, mk())
    b, _ = _run_worker(m, mk())
    assert a == b
    for p, toks in zip(prompts, a):
        st = m.generate_zero_state(1)
        lg = m.forward_seq_batch([p], st)
        for t in toks:
            assert t in torch.topk(lg.float()[0], 3).indices.tolist()
            lg = m.forward_seq_batch([[t]], st)
    assert len({tuple(x) for x in a}) > 1


def test_state_cache_hit_inside_the_engine_matches_cold_run():
    """SURVEY §8(f)-3 through the engine (chirrup/web_service/app.py:187-214, :293-298; chirrup/worker.py:411-435, :457-476):
    request 1 prefills a 180-token prompt and the worker emits its ("cache_prefill") state — kept on the GPU; an identical
    concurrent request waits for it; request 3 shares the prefix and extends it.  Every request produces exactly the tokens of
    a cold isolated run, and the cached requests consume only their uncached remainder."""
    from chirrup_b200.engine import AsyncEngineCore, CachedCompletion, ModelLoadConfig
    from chirrup_b200.state_cache import PrefixStateCache

    geom = GEOMETRIES["small"]
    raw = make_checkpoint(geom, seed=8, recipe="trained_like")
    _, m_ref = _model()                                       # same seed: the isolated reference runs
    g = torch.Generator().manual_seed(3)
    prompt = torch.randint(1, geom.vocab_size, (180,), generator=g).tolist()
    ext = prompt + torch.randint(1, geom.vocab_size, (25,), generator=g).tolist()
    pad = 3

    async def main():
        eng = AsyncEngineCore()
        await eng.init(worker_num=1, model_config=ModelLoadConfig(state_dict=raw, vocab_size=geom.vocab_size), batch_size=8)
        cache = PrefixStateCache(8)
        kw = dict(top_k=1, max_tokens=10, stop_tokens=[], presence_penalty=0.0, frequency_penalty=0.0)
        a = CachedCompletion(eng, cache, prompt, cache_prefill_padding=pad, **kw)
        b = CachedCompletion(eng, cache, prompt, cache_prefill_padding=pad, **kw)
        ta, tb = await asyncio.gather(a.get_full_completion(), b.get_full_completion())
        c = CachedCompletion(eng, cache, ext, cache_prefill_padding=pad, **kw)
        tc = await c.get_full_completion()
        w = eng.workers[0]
        stats = dict(w.steps)
        entry = cache.entries[tuple(prompt[:-pad])]
        eng.shutdown()
        return ta, tb, tc, (a.cached_token_len, b.cached_token_len, c.cached_token_len), stats, entry

    ta, tb, tc, hits, stats, entry = asyncio.run(main())
    assert hits == (0, len(prompt) - pad, len(prompt) - pad)
    assert ta == tb == _greedy_alone(m_ref, prompt, 10)
    assert tc == _greedy_alone(m_ref, ext, 10)
    assert entry[0].is_cuda and int(entry[2][0]) == len(prompt) - pad          # GPU-resident tier; elapsed = tokens consumed
    # prefill work: 180 (cold) + 3 (b: padding only) + 28 (c: padding + extension), none of the cached 177 tokens twice
    assert stats["prefill_tokens"] <= 180 + 3 + 28


def test_batched_slot_migration_equals_slot_by_slot():
    """state_move.migrate_slots (one b200_slot_move call per slot, one ordering event per batch) moves exactly the bytes
    migrate_slot moves — on one GPU here; between peers when two are visible."""
    from chirrup_b200 import state_move

    geom, m = _model()
    devs = ["cuda:0", "cuda:1" if torch.cuda.device_count() > 1 else "cuda:0"]
    a = m.generate_zero_state(6)
    for t in a[:2]:
        t.normal_(0, 1)
    a[2].copy_(torch.arange(6, dtype=torch.int32) * 7 + 3)
    b1 = [torch.zeros_like(t, device=devs[1]) for t in a]
    b2 = [torch.zeros_like(t, device=devs[1]) for t in a]
    src, dst = [4, 0, 2], [1, 5, 3]
    moved = state_move.migrate_slots(a, src, b1, dst)
    for s_, d_ in zip(src, dst):
        state_move.migrate_slot(a, s_, b2, d_)
    for dev in set(devs):
        torch.cuda.synchronize(dev)
    assert moved == 3 * state_move.slot_bytes(a)
    for x, y in zip(b1, b2):
        assert torch.equal(x.cpu(), y.cpu())
    for s_, d_ in zip(src, dst):
        assert torch.equal(b1[1][:, d_].cpu(), a[1][:, s_].cpu()) and int(b1[2][d_]) == int(a[2][s_])
    assert torch.count_nonzero(b1[1][:, 0]) == 0 and torch.count_nonzero(b1[0][:, :, 2]) == 0        # other slots untouched


def test_process_per_gpu_engine_matches_isolated_greedy():
    """MPEngineCore (one worker PROCESS per GPU behind the AsyncEngineCore API): the worker builds the seeded synthetic model on
    its own GPU; every request's tokens equal an isolated greedy run of the same model in this process; stop tokens end a request
    without being delivered; both workers of a 2-GPU box take work."""
    from chirrup_b200.engine import FinishReason, ModelLoadConfig
    from chirrup_b200.engine_mp import MPEngineCore
    from chirrup_b200.rwkv7 import RWKV_x070, make_args

    geom = GEOMETRIES["small"]
    syn = {"geom": "small", "seed": 8, "recipe": "trained_like"}
    m_ref = RWKV_x070(make_args(), state_dict=make_checkpoint(geom, seed=8, recipe="trained_like", device="cuda", dtype=torch.float16))
    g = torch.Generator().manual_seed(2)
    prompts = [torch.randint(1, geom.vocab_size, (n,), generator=g).tolist() for n in (1, 5, 12, 47, 130, 3, 64, 9)]
    n_workers = min(2, torch.cuda.device_count())

    async def main():
        eng = MPEngineCore()
        await eng.init(n_workers, ModelLoadConfig(vocab_size=geom.vocab_size, synthetic=syn), batch_size=4, load_timeout_s=300)
        kw = dict(top_k=1, max_tokens=9, stop_tokens=[], presence_penalty=0.0, frequency_penalty=0.0)
        comps = [eng.completion(prefill_tokens=p, **kw) for p in prompts]
        outs = await asyncio.gather(*[c.get_full_completion() for c in comps])
        streamed = [ev[1] async for ev in eng.completion(prefill_tokens=prompts[3], **kw)]        # per-token iteration form
        want = _greedy_alone(m_ref, prompts[2], 9)
        stopper = eng.completion(prefill_tokens=prompts[2], **dict(kw, stop_tokens=[want[4]]))
        stopped = await stopper.get_full_completion()
        eng.shutdown()
        return outs, streamed, stopped, stopper.finish_reason, want, eng.worker_steps

    outs, streamed, stopped, reason, want, steps = asyncio.run(main())
    for p, o in zip(prompts, outs):
        assert o == _greedy_alone(m_ref, p, 9), f"prompt len {len(p)}"
    assert streamed == outs[3]
    assert stopped == want[:want.index(want[4])] and reason == FinishReason.STOP
    assert len(steps) == n_workers and all(v["decode_tokens"] > 0 for v in steps.values())
