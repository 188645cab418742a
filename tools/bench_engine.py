"""BASELINE.json configs[2] through the continuous-batching WORKER loop (chirrup_b200/engine.py Worker.step, the counterpart of
chirrup/worker.py:793-889): 7.2B geometry, 256 decode slots (the reference's batch_size 257 is 256 + one scratch row for its
physical row swaps, worker.py:141-142; here rows are addressed through a slot-index vector, so no scratch row),
synthetic prompts of 16-64 seeded tokens, max_tokens 256, greedy with zero penalties (as scripts/test_mmlu/test_mmlu.py:59-69),
stop_tokens = [].  Reports decode tok/s of the whole job (wall clock, admission + prefill + decode + host loop) and of the
steady-state window in which all 256 slots decode.
Usage (GPU box): python tools/bench_engine.py [--geom 7.2B] [--slots 256] [--requests 512] [--max-tokens 256] [--sampled]"""
import argparse, json, os, queue, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chirrup_b200.engine import ModelLoadConfig, StateCategory, Task, Worker
from chirrup_b200.rwkv7 import RWKV_x070, make_args
from chirrup_b200.synth import GEOMETRIES, make_checkpoint


class Sink:
    def __init__(self):
        self.n = 0

    def put_nowait(self, x):
        self.n += 1


class CountSink:
    def __init__(self):
        import threading
        self.n, self.done, self.lock = 0, 0, threading.Lock()

    def put_nowait(self, x):
        with self.lock:
            if x[0] == "token":
                self.n += 1
            else:
                self.done += 1


def multi(a, geom):
    """SURVEY §8(d) config 3 across GPUs: one Worker thread per GPU in ONE process, one shared task queue, no collective."""
    import threading
    tq = queue.Queue()
    workers, threads = [], []
    for k in range(a.workers):
        dev = torch.device("cuda", k)
        with torch.cuda.device(dev):
            m = RWKV_x070(make_args(), state_dict=make_checkpoint(geom, seed=0, recipe="trained_like", device=dev, dtype=torch.float16),
                          device=dev)
        w = Worker(f"w{k}", [k], ModelLoadConfig(vocab_size=geom.vocab_size), tq, batch_size=a.slots, model=m)
        w.max_prefill_count = a.slots
        w.pipeline = not a.no_pipeline
        workers.append(w)
    g = torch.Generator().manual_seed(0)
    sink = CountSink()
    kw = {} if a.sampled else dict(top_k=1, presence_penalty=0.0, frequency_penalty=0.0)
    total = a.requests * a.workers
    tasks = []
    for i in range(total):
        n = int(torch.randint(16, 65, (1,), generator=g))
        tasks.append(Task(prefill_tokens=torch.randint(1, geom.vocab_size, (n,), generator=g).tolist(), output_queue=sink,
                          max_tokens=a.max_tokens, stop_tokens=[], **kw))
    for w in workers:
        th = threading.Thread(target=w.start, daemon=True, name=w.worker_id)
        threads.append(th)
        th.start()
    time.sleep(2.0)                       # workers initialised and idle
    t0 = time.perf_counter()
    for t in tasks:
        tq.put(t)
    while sink.done < total:
        time.sleep(0.002)
    dt = time.perf_counter() - t0
    for w in workers:
        w._stop = True
    for th in threads:
        th.join(timeout=10)
    per = [w.steps["decode_tokens"] for w in workers]
    print(json.dumps({"config": f"{a.geom} worker loop x{a.workers} GPUs (threads of one process, one task queue), {a.slots} slots each, "
                                f"{total} requests x {a.max_tokens} new tokens, "
                                + ("sampled (T 1.0, top_p 0.3, penalties 0.5)" if a.sampled else "greedy, penalties 0"),
                      "decode_tok_per_s_whole_job_aggregate": sum(per) / dt, "wall_s": dt, "decode_tokens_per_worker": per,
                      "decode_steps_per_worker": [w.steps["decode"] for w in workers], "tokens_delivered": sink.n}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--geom", default="7.2B"); ap.add_argument("--slots", type=int, default=256)
    ap.add_argument("--requests", type=int, default=512); ap.add_argument("--max-tokens", type=int, default=256)
    ap.add_argument("--sampled", action="store_true", help="temperature 1.0, top_p 0.3, penalties 0.5 (the reference's defaults)")
    ap.add_argument("--profile", action="store_true", help="host-side wall time of the pieces of a step (no device sync added)")
    ap.add_argument("--no-pipeline", action="store_true")
    ap.add_argument("--workers", type=int, default=1, help="> 1: that many Worker threads, one per GPU, pulling from ONE task queue "
                    "(chirrup/engine_core.py:71, :135-153); --requests is per worker")
    a = ap.parse_args()
    geom = GEOMETRIES[a.geom]
    if a.workers > 1:
        return multi(a, geom)
    dev = torch.device("cuda")
    m = RWKV_x070(make_args(), state_dict=make_checkpoint(geom, seed=0, recipe="trained_like", device=dev, dtype=torch.float16))
    w = Worker("w0", [0], ModelLoadConfig(vocab_size=geom.vocab_size), queue.Queue(), batch_size=a.slots, model=m)   # no scratch row: slots are addressed by index, never swapped
    w.max_prefill_count = a.slots     # the reference admits batch_size // 8 per loop; fill the batch at once for the steady state
    w._init_worker()
    w.pipeline = not a.no_pipeline
    host = {}
    if a.profile:
        def wrap(obj, name):
            fn = getattr(obj, name)
            def timed(*x, **k):
                t = time.perf_counter()
                r = fn(*x, **k)
                host[name] = host.get(name, 0.0) + time.perf_counter() - t
                return r
            setattr(obj, name, timed)
        orig_launch, orig_finish = w._launch_decode, w._finish_decode
        def launch(rows, feed):
            e0 = torch.cuda.Event(enable_timing=True); e0.record()
            rec = orig_launch(rows, feed)
            e1 = torch.cuda.Event(enable_timing=True); e1.record()
            rec["t"] = (e0, e1)
            return rec
        def finish(rec):
            t = time.perf_counter()
            if rec["event"] is not None:
                rec["event"].synchronize()
            host["wait_for_ids"] = host.get("wait_for_ids", 0.0) + time.perf_counter() - t
            r = orig_finish(rec)
            rec["t"][1].synchronize()
            host["gpu_step_ms_events"] = host.get("gpu_step_ms_events", 0.0) + rec["t"][0].elapsed_time(rec["t"][1]) * 1e-3
            return r
        w._launch_decode, w._finish_decode = launch, finish
        for nm in ("_step", "_tokens"):
            wrap(w.model, nm)
        for nm in ("_sample_launch", "_finish_decode", "_launch_decode", "_process_events", "_fill_task_pool", "_state_for",
                   "_run_forward_seq"):
            wrap(w, nm)
    g = torch.Generator().manual_seed(0)
    sink = Sink()
    kw = {} if a.sampled else dict(top_k=1, presence_penalty=0.0, frequency_penalty=0.0)
    for i in range(a.requests):
        n = int(torch.randint(16, 65, (1,), generator=g))
        w.task_queue.put(Task(prefill_tokens=torch.randint(1, geom.vocab_size, (n,), generator=g).tolist(), output_queue=sink,
                              max_tokens=a.max_tokens, stop_tokens=[], **kw))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    steady_t, steady_tok, steps = 0.0, 0, 0
    while True:
        full = sum(s.category == StateCategory.FORWARD_ONE_DECODE for s in w.slots) >= a.slots
        d0 = w.steps["decode_tokens"]
        ts = time.perf_counter()
        n = w.step()
        if full:
            steady_t += time.perf_counter() - ts
            steady_tok += w.steps["decode_tokens"] - d0
        steps += 1
        if n == 0 and not w.busy() and w.task_queue.empty():
            break
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(json.dumps({"config": f"{a.geom} worker loop, {a.slots} slots, {a.requests} requests x {a.max_tokens} new tokens, "
                                + ("sampled (T 1.0, top_p 0.3, penalties 0.5)" if a.sampled else "greedy, penalties 0"),
                      "decode_tok_per_s_whole_job": w.steps["decode_tokens"] / dt, "wall_s": dt,
                      "decode_tok_per_s_steady_state": steady_tok / max(steady_t, 1e-9),
                      "ms_per_step_steady_state": 1e3 * steady_t / max(1, steady_tok / a.slots),
                      "decode_steps": w.steps["decode"], "prefill_steps": w.steps["prefill"], "tokens_delivered": sink.n,
                      "host_ms_per_decode_step": {k: round(1e3 * v / max(1, w.steps["decode"]), 3) for k, v in host.items()}}))


if __name__ == "__main__":
    main()
