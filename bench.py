#!/usr/bin/env python
"""Benchmark of the hot path: RWKV-7 batched decode step (BASELINE.json metric
"RWKV-7 7.2B decode tok/s/GPU @bsz256; WKV step HBM GB/s vs peak").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config decode|prefill|engine|migration]
                    [--geom 7.2B] [--bsz 256]

--config decode (default, BASELINE configs[2]): one "step" = one decode forward (T = 1) of `bsz` sequences per GPU through every
layer + head + greedy sampling.  N > 1: launched by torchrun, one replica per GPU, requests sharded by rank, no data-path
collective ("weak" scaling); NCCL is used only for the barrier and the max-over-ranks of the device time.
Printed JSON keys (contract in the task statement): value = whole-job tok/s with inputs resident in HBM (CUDA-graph replay of
the step, CUDA events); e2e = the same through the public Python API with HOST token lists (List[List[int]] as Chirrup's worker
passes them: H2D inside the API, D2H of the sampled ids, host-synchronised every step); roofline = the fused WKV kernel against
the measured HBM peak, alone and inside the step; roofline_gemm = the rest of the step against the tensor peak; engine = the
same configuration through the continuous-batching engine (one worker process per GPU, one task queue); cpu_baseline = the
reference's own model code on the host cores (a child process running the reference arm, one full-depth step).
--config prefill / engine / migration: BASELINE configs[3], configs[2] through the worker loop only, configs[4] (2 GPUs).
--impl reference: the reference's own Albatross/rwkv7.py on the host cores (oracle/ref_cpu.py), this arm's configuration.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch

METRIC = "RWKV-7 7.2B decode tok/s/GPU @bsz256; WKV step HBM GB/s vs peak"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--geom", default="7.2B")
    ap.add_argument("--bsz", type=int, default=256)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--config", default="decode", choices=["decode", "prefill", "engine", "migration"],
                    help="decode: BASELINE configs[2] (the headline, default); prefill: configs[3] (7.2B, 32 x 4096-token prompts, "
                         "chunked WKV, state-cache reuse); engine: configs[2] through the AsyncEngineCore worker loop, one worker "
                         "PROCESS per GPU on one task queue; migration: configs[4] (13.3B, 512 slots, P2P slot migration, 2 GPUs)")
    ap.add_argument("--no-engine", action="store_true", help="decode config: leave out the engine-level (worker loop) phase")
    ap.add_argument("--engine-requests", type=int, default=512, help="requests per worker in the engine phase")
    ap.add_argument("--skip-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-budget-s", type=float, default=100.0, help="reference arm: stop timing steps after this many seconds")
    ap.add_argument("--skip-configs0", action="store_true", help="reference arm: leave out the BASELINE configs[0] figure")
    return ap.parse_args()


def peaks():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(d["hbm_gbs"]), float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "measured"
    except Exception:
        return 6650.0, 1400.0, "fallback"


class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz, self.power_w, self.power_limit_w = [], set(), None, [], None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            try:
                self.power_limit_w = pynvml.nvmlDeviceGetEnforcedPowerLimit(self.h) / 1000.0
            except Exception:
                pass
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    self.power_w.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
                except Exception:
                    pass
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.02)

    def __enter__(self):
        if self.nv is not None:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t is not None:
            self._t.join()

    def summary(self):
        s = sorted(self.samples)
        p = sorted(self.power_w)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s), "power_w": p[len(p) // 2] if p else None, "power_limit_w": self.power_limit_w}


# ----------------------------------------------------------------------------------------------- CPU baseline / reference arm
def host_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


class _StdoutToStderr:
    """The reference model prints its parameter table to stdout at construction; the bench contract wants ONE JSON line there."""

    def __enter__(self):
        sys.stdout.flush()
        self._saved = os.dup(1)
        os.dup2(2, 1)

    def __exit__(self, *a):
        sys.stdout.flush()
        os.dup2(self._saved, 1)
        os.close(self._saved)


def cpu_reference_run(geom_name: str, bsz: int, prompt_len: int, steps: int, warmup: int, budget_s: float = 100.0,
                      n_layer=None):
    """Times the reference's OWN model code (Albatross/rwkv7.py, unmodified, staged under oracle/_ref by oracle/stage_ref_model.sh;
    seam: oracle/ref_cpu.py) on the host cores: `warmup` untimed + up to `steps` timed greedy decode steps (T = 1) of `bsz`
    sequences through EVERY layer + head + arg-max — measured, nothing extrapolated.  Timed steps stop early once `budget_s`
    is spent; the count actually timed is returned.  If the staged reference is missing the oracle port runs instead
    (kind "port")."""
    from chirrup_b200.synth import GEOMETRIES, make_checkpoint

    torch.set_num_threads(host_threads())      # torchrun exports OMP_NUM_THREADS=1: set the pool explicitly
    geom = GEOMETRIES[geom_name]
    gen_dev = "cuda" if torch.cuda.is_available() else "cpu"   # the checkpoint is only GENERATED on the GPU (76 s on CPU at 7.2B)
    raw = make_checkpoint(geom, seed=0, recipe="trained_like", device=gen_dev, dtype=torch.float16, n_layer=n_layer)
    raw = {k: v.cpu() for k, v in raw.items()}
    if gen_dev == "cuda":
        torch.cuda.empty_cache()
    g = torch.Generator().manual_seed(0)
    prompt = torch.randint(1, geom.vocab_size, (bsz, max(prompt_len, 1)), generator=g).tolist()
    kind = "reference"
    try:
        from oracle import ref_cpu

        with _StdoutToStderr():
            R, m = ref_cpu.load_model(raw, geom.vocab_size)
        st = ref_cpu.zero_state(m, bsz)
        fwd = lambda toks: m.forward_seq_batch_seperate(toks, st)
        what = "the reference's own Albatross/rwkv7.py (TorchScript, fp16 ATen ops) + C body for its WKV custom op"
    except Exception as e:  # noqa: BLE001 — staged reference absent: fall back to the pinned port, and say so
        sys.stderr.write(f"[bench] staged reference unavailable ({e!r}); timing the oracle port instead\n")
        from oracle import rwkv7_oracle as O

        kind = "port"
        z = O.transform_checkpoint(raw)
        og = O.Geometry(z)
        st = O.zero_state(og, bsz)
        fwd = lambda toks: O.forward_seq_batch(z, og, toks, st)
        what = "oracle port of Albatross/rwkv7.py (same fp16 ATen ops, same order) + C WKV body"
    del raw
    with _StdoutToStderr():
        lg = fwd(prompt)
        times = []
        t_begin = time.perf_counter()
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            nxt = lg.float().argmax(-1)
            lg = fwd([[t] for t in nxt.tolist()])
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt)
                if time.perf_counter() - t_begin > budget_s:
                    break
    cores = torch.get_num_threads()
    step_s = sum(times) / len(times)
    L = geom.n_layer if n_layer is None else n_layer
    sample = (f"{what}; {geom_name} geometry, all {L} layers + head, bsz {bsz}, {len(prompt[0])}-token prompt then T=1 greedy: "
              f"{len(times)} timed step(s) after {warmup} warm-up, measured (not extrapolated); {cores} threads of "
              f"{os.cpu_count()} cpus")
    return {"value": bsz / step_s, "unit": "tok/s", "cores": cores, "kind": kind, "sample": sample,
            "steps_timed": len(times), "warmup": warmup, "ms_per_step": 1e3 * step_s, "extrapolated": False}


def cpu_baseline_subprocess(geom_name: str, bsz: int):
    """cpu_baseline of the GPU arm: the reference arm of this file in a child process (keeps the reference's torch extension and
    its op registrations out of the measured process), bounded to 1 warm-up + 1 timed full-depth step."""
    import subprocess

    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT",
                                                             "OMP_NUM_THREADS", "TORCHELASTIC_RUN_ID")}
    r = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "1", "--warmup", "1",
                        "--geom", geom_name, "--bsz", str(bsz), "--skip-configs0"], capture_output=True, text=True, env=env,
                       timeout=900)
    for line in reversed(r.stdout.strip().splitlines()):
        if line.startswith("{"):
            return json.loads(line)["cpu_baseline"]
    raise RuntimeError(f"cpu baseline child failed rc={r.returncode}: {r.stderr[-2000:]}")


# ----------------------------------------------------------------------------------------------- our arm
def build_model(geom_name: str, device):
    from chirrup_b200.rwkv7 import RWKV_x070, make_args
    from chirrup_b200.synth import GEOMETRIES, make_checkpoint

    geom = GEOMETRIES[geom_name]
    raw = make_checkpoint(geom, seed=0, recipe="trained_like", device=device, dtype=torch.float16)
    m = RWKV_x070(make_args(), state_dict=raw, device=device)
    del raw
    torch.cuda.empty_cache()
    return geom, m


def time_wkv_kernel(m, geom, state, bsz, reps=5):
    """Average device time of one launch (T = 1) of the step's HBM-roofline kernel — the fused TMix core (prep + WKV-7 +
    GroupNorm/bonus/gate, b200_tmix_core_fused) — and of the plain WKV kernel (b200_wkv7_forward_seq), each over the model's real
    per-layer state (L x bsz x 8 KiB tiles: 4 GiB at 7.2B/bsz 256 >> L2), replayed from a CUDA graph (so that host launch
    overhead does not pollute the per-launch device time) and timed with CUDA events on the launching stream."""
    from chirrup_b200 import ops

    C, H, L = geom.n_embd, geom.n_head, geom.n_layer
    dev = state[1].device
    mk = lambda s: (torch.randn(bsz, 1, C, device=dev) * s).half()
    r, k, v, g, vf = mk(0.5), mk(0.5), mk(0.5), mk(1.0), mk(0.5)
    w = (torch.randn(bsz, 1, C, device=dev) - 1).half()
    a_sig, vg = torch.sigmoid(mk(1.0).float()).half(), torch.sigmoid(mk(1.0).float()).half()
    kk = torch.nn.functional.normalize(torch.randn(bsz, 1, H, 64, device=dev), dim=-1).view(bsz, 1, C)
    a, b = (-kk).half(), (kk * torch.sigmoid(torch.randn(bsz, 1, C, device=dev))).half()
    vec = lambda mean, std: (torch.randn(C, device=dev) * std + mean).half()
    k_k, k_a, r_k, ln_w, ln_b = vec(0.85, 0.1), vec(1.0, 0.1), vec(0.0, 0.1), vec(1.0, 0.05), vec(0.0, 0.02)
    y = torch.empty_like(r)
    et = torch.zeros(bsz, dtype=torch.int32, device=dev)
    scratch = state[1].clone()
    runs = {
        "fused": lambda: [ops.tmix_core_fused(scratch[l], r, w, k, v, a_sig, g, vg, vf, k_k, k_a, r_k, ln_w, ln_b, et, out=y)
                          for l in range(L)],
        "plain": lambda: [ops.wkv7_forward_seq(scratch[l], r, w, k, v, a, b, et, y=y) for l in range(L)],
    }
    out = {}
    for name, run in runs.items():
        run()
        torch.cuda.synchronize()
        g_ = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            run()
            with torch.cuda.graph(g_, stream=s):
                run()
        torch.cuda.current_stream().wait_stream(s)
        g_.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            g_.replay()
        e1.record()
        torch.cuda.synchronize()
        out[name] = e0.elapsed_time(e1) * 1e-3 / (reps * L)
        del g_
    del scratch
    return out


def run_ours(a, rank, world, device):
    from chirrup_b200 import _lib, ops

    geom, m = build_model(a.geom, device)
    B, V = a.bsz, geom.vocab_size
    state = m.generate_zero_state(B)
    gen = torch.Generator().manual_seed(1234 + rank)
    prompt = torch.randint(1, V, (B, 16), generator=gen)
    m.forward_seq_batch(prompt.tolist(), state)          # short synthetic prompt so the state is not all zeros
    tok = torch.randint(1, V, (B,), generator=gen).to(torch.int32).to(device)
    logits = torch.empty((B, V), dtype=torch.float16, device=device)
    nxt = torch.empty((B,), dtype=torch.int32, device=device)

    def step():   # one decode step, everything resident in HBM: forward -> logits -> greedy sample -> next token ids
        m._step(tok, B, 1, state, False, out=logits)
        ops.argmax(logits, out=nxt)
        tok.copy_(nxt)

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    launches_per_step = _lib.lib().b200_model_last_launch_count(m._handle) + 1   # + argmax (the token copy is a memcpy node)
    graph = None
    if not a.no_graph:
        try:
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                step()
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=s):
                    step()
            torch.cuda.current_stream().wait_stream(s)
            torch.cuda.synchronize()
        except Exception as e:  # noqa: BLE001 — report and time eager launches instead
            sys.stderr.write(f"[bench] CUDA graph capture failed ({e!r}); timing eager launches\n")
            graph = None
    do_step = graph.replay if graph is not None else step

    from chirrup_b200 import dist_util

    def barrier():
        torch.cuda.synchronize()
        dist_util.barrier()
        torch.cuda.synchronize()

    for _ in range(a.warmup):
        do_step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(device.index or 0) as clk:
        e0.record()
        for _ in range(a.steps):
            do_step()
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    barrier()

    # ---- end to end through the public API with host buffers (the call Chirrup's worker makes, chirrup/worker.py:704)
    host_out = torch.empty((B,), dtype=torch.int32).pin_memory()
    e2e_tokens = [torch.randint(1, V, (B, 1), generator=gen).tolist()]

    def e2e_step():
        # the call Chirrup's worker makes (chirrup/worker.py:690-704): token ids as a Python List[List[int]] — torch.tensor(list)
        # + pageable H2D inside the API — then logits -> greedy sample -> D2H of the ids -> the next step's list
        lg = m.forward_seq_batch_seperate(e2e_tokens[0], state)
        ops.argmax(lg, out=nxt)
        host_out.copy_(nxt, non_blocking=True)                                  # D2H of the sampled ids
        torch.cuda.current_stream().synchronize()
        e2e_tokens[0] = [[t] for t in host_out.tolist()]                        # next step's input comes from the host side

    for _ in range(max(3, a.warmup, getattr(m, "graph_after", 0) + 2)):   # past the point where the API captures its step graph
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    barrier()

    wkv_s = time_wkv_kernel(m, geom, state, B) if rank == 0 else {"fused": 0.0, "plain": 0.0}
    # in-step figure: the same step graph captured with the fused WKV launches knocked out (model.cu B200_DEBUG_SKIP=8; results
    # are garbage, only the time is read): step - knocked-out step = what the 32 WKV launches cost INSIDE the step, PDL overlap,
    # L2 state of the neighbours and clocks under load included (an upper bound: skipping also breaks the PDL chain)
    wkv_in_step_s = None
    if rank == 0 and graph is not None:
        try:
            os.environ["B200_DEBUG_SKIP"] = "8"
            s2 = torch.cuda.Stream()
            s2.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s2):
                step()
                g2 = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g2, stream=s2):
                    step()
            torch.cuda.current_stream().wait_stream(s2)
            os.environ.pop("B200_DEBUG_SKIP", None)

            def timed(gr, n):
                for _ in range(3):
                    gr.replay()
                torch.cuda.synchronize()
                a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a0.record()
                for _ in range(n):
                    gr.replay()
                a1.record()
                torch.cuda.synchronize()
                return a0.elapsed_time(a1) * 1e-3 / n

            n = max(10, a.steps)
            t_full, t_skip = timed(graph, n), timed(g2, n)
            wkv_in_step_s = max(t_full - t_skip, 0.0) / geom.n_layer
            del g2
        finally:
            os.environ.pop("B200_DEBUG_SKIP", None)
    ms, e2e_ms = dist_util.max_over_ranks([ms, e2e_s * 1e3], device=device)
    return dict(geom=geom, ms=ms, e2e_ms=e2e_ms, clocks=clk.summary(), launches=launches_per_step * a.steps,
                graph=graph is not None, wkv_s=wkv_s, B=B, wkv_in_step_s=wkv_in_step_s)


# ----------------------------------------------------------------------------------------------- engine / prefill / migration
def engine_run(geom_name: str, n_workers: int, slots: int, requests_per_worker: int, max_tokens: int = 256, load_timeout_s: float = 900.0):
    """BASELINE configs[2] as SURVEY §8(d) defines it: continuous batching through the AsyncEngineCore API — here
    MPEngineCore: one worker PROCESS per GPU, one shared task queue — with `slots` decode slots per worker, synthetic prompts of
    16-64 seeded tokens, max_tokens new tokens each, greedy, penalties 0, stop_tokens = [] (scripts/test_mmlu/test_mmlu.py:59-69).
    Whole-job wall clock from the first submission to the last finish event, tokens counted where a client receives them."""
    import asyncio

    from chirrup_b200.engine import ModelLoadConfig
    from chirrup_b200.engine_mp import MPEngineCore
    from chirrup_b200.synth import GEOMETRIES

    geom = GEOMETRIES[geom_name]

    async def main():
        eng = MPEngineCore()
        t_load = time.perf_counter()
        await eng.init(n_workers, ModelLoadConfig(vocab_size=geom.vocab_size, synthetic={"geom": geom_name, "seed": 0,
                                                                                           "recipe": "trained_like"}),
                       batch_size=slots, load_timeout_s=load_timeout_s, worker_options={"max_prefill_count": slots})
        t_load = time.perf_counter() - t_load
        g = torch.Generator().manual_seed(0)
        kw = dict(top_k=1, presence_penalty=0.0, frequency_penalty=0.0, stop_tokens=[], max_tokens=max_tokens)

        def make(n_req):
            out = []
            for _ in range(n_req):
                n = int(torch.randint(16, 65, (1,), generator=g))
                out.append(eng.completion(prefill_tokens=torch.randint(1, geom.vocab_size, (n,), generator=g).tolist(), **kw))
            return out

        async def drain(c):
            return len(await c.get_full_completion())

        # warm-up wave: step graphs captured, allocator settled, every worker's pipeline primed
        warm = make(n_workers * slots)
        await asyncio.gather(*[drain(c) for c in warm])
        comps = make(n_workers * requests_per_worker)
        t0 = time.perf_counter()
        counts = await asyncio.gather(*[drain(c) for c in comps])
        dt = time.perf_counter() - t0
        eng.shutdown()
        per_worker = {k: v["decode_tokens"] for k, v in eng.worker_steps.items()}
        steady = {k: (slots * v["full_steps"] / v["full_time_s"] if v.get("full_time_s") else None) for k, v in eng.worker_steps.items()}
        return sum(counts), dt, t_load, per_worker, steady, dict(eng.worker_steps)

    n_tok, dt, t_load, per_worker, steady, worker_stats = asyncio.run(main())
    ok = [v for v in steady.values() if v]
    gh = sum(v.get("graph_hits", 0) for v in worker_stats.values())
    gm = sum(v.get("graph_misses", 0) for v in worker_stats.values())
    return {"value": n_tok / dt, "steady_state_tok_per_s_per_worker": steady,
            "step_graph_hit_rate": gh / max(1, gh + gm), "step_graph_hits": gh, "step_graph_misses": gm,
            "steady_state_tok_per_s": sum(ok) if len(ok) == len(steady) and ok else None,
            "steady_state_definition": "loops in which every slot of the worker took a decode step (no admission, no prefill chunk in "
                                       "between): slots x such loops / their wall time, summed over workers", "unit": "tok/s", "tokens": n_tok, "wall_s": dt, "workers": n_workers, "slots_per_worker": slots,
            "requests": n_workers * requests_per_worker, "max_tokens": max_tokens, "load_s": t_load,
            "decode_tokens_per_worker_incl_warmup": per_worker,
            "api": "MPEngineCore.completion(prefill_tokens=...) async iteration: one worker process per GPU, one shared task "
                   "queue, one event batch per worker step; prompts H2D at admission, sampled ids D2H every step"}


def prefill_run(a, device):
    """BASELINE configs[3]: 7.2B, 32 prompts x 4096 seeded tokens through prefill() in 512-token chunks (chunked tensor-core WKV,
    persistent CTA-pair GEMMs), then the same with a shared 2048-token prefix served from the GPU-resident state cache."""
    from chirrup_b200 import state_move
    from chirrup_b200.state_cache import PrefixStateCache

    geom, m = build_model(a.geom, device)
    B, T, P, chunk = 32, 4096, 2048, 512
    g = torch.Generator().manual_seed(0)
    prefix = torch.randint(1, geom.vocab_size, (P,), generator=g)
    prompts = torch.cat([prefix.expand(B, -1), torch.randint(1, geom.vocab_size, (B, T - P), generator=g)], 1).to(torch.int32)
    host = prompts.pin_memory()

    def timed(fn, n):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            r = fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e-3 / n, r

    st = m.generate_zero_state(B)
    cold = lambda: m.prefill(host, st, chunk_len=chunk)          # host token ids in: the H2D copy is inside the timed region
    for _ in range(max(1, min(a.warmup, 2))):
        cold()
    steps = max(1, min(a.steps, 5))
    with ClockSampler(device.index or 0) as clk:
        t_cold, lg_cold = timed(cold, steps)
    os.environ["B200_WKV_SEQUENTIAL"] = "1"
    try:
        cold()
        t_seq, _ = timed(cold, 1)
    finally:
        os.environ.pop("B200_WKV_SEQUENTIAL", None)
    cache = PrefixStateCache(8)
    s1 = m.generate_zero_state(1)
    m.prefill(prompts[:1, :P], s1, chunk_len=chunk, want_logits=False)
    cache.cache(tuple(prefix.tolist()), state_move.extract_slot(s1, 0))
    st2 = m.generate_zero_state(B)

    def warm():
        for b in range(B):
            rest, packed, n = cache.check(host[b].tolist())
            assert n == P
            state_move.inject_slot(st2, b, packed)
        return m.prefill(host[:, P:], st2, chunk_len=chunk)

    warm()
    t_warm, lg_warm = timed(warm, 1)
    flops = geom.gemm_flops_per_token() * B * T - 2 * geom.n_embd * geom.vocab_size * B * (T - 1)   # the head runs on B rows only
    return dict(geom=geom, B=B, T=T, chunk=chunk, steps=steps, t_cold=t_cold, t_seq=t_seq, t_warm=t_warm, flops=flops,
                clocks=clk.summary(), logits_rel=float((lg_warm.float() - lg_cold.float()).abs().max() / lg_cold.float().abs().max()))


# ----------------------------------------------------------------------------------------------- main
def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    from chirrup_b200.synth import GEOMETRIES

    geom = GEOMETRIES[a.geom]
    config = {"workload": f"RWKV-7 g1 {a.geom} (L{geom.n_layer}, D{geom.n_embd}) fp16 decode, bsz {a.bsz}/GPU, T=1, "
                          f"greedy sampling, one replica per GPU (BASELINE.json configs[2])",
              "batch_per_gpu": a.bsz, "weights": "random-init trained_like recipe (chirrup_b200/synth.py), seed 0",
              "l2": "weights (14.4 GB at 7.2B) + state (4.3 GB) streamed every step: inputs larger than L2, no flush needed",
              "parallelism": f"replicas x{max(world, 1)} (requests sharded by rank, no collective on the step path)"}
    base = {"metric": METRIC, "unit": "tok/s", "n_gpus": max(world, 1), "steps": a.steps, "warmup": a.warmup,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
            "config": config}

    if a.impl == "reference":
        if rank != 0:
            return
        # The reference's own CPU implementation of the path on the box's host cores, on this arm's config (7.2B, bsz 256, T = 1),
        # full depth, measured; the number of steps actually timed is what "steps" says.  BASELINE.md §4's configs[0] figure
        # (0.1B, bsz 1, 16-token prompt, 3 warm-up + 128 greedy tokens) rides along as `cpu_baseline_configs0`.
        cb = cpu_reference_run(a.geom, a.bsz, 1, max(1, a.steps), max(1, min(a.warmup, 2)), budget_s=a.cpu_budget_s)
        out = dict(base)
        out.update({"impl": "reference", "value": cb["value"], "ms_per_step": cb["ms_per_step"], "steps": cb["steps_timed"],
                    "warmup": cb["warmup"], "n_gpus": max(world, 1), "dtype": "f16",
                    "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample", "extrapolated")},
                    "e2e": {"value": cb["value"], "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                    "gpu_launches": 0})
        if not a.skip_configs0:
            c0 = cpu_reference_run("0.1B", 1, 16, 128, 3, budget_s=60.0)
            out["cpu_baseline_configs0"] = {k: c0[k] for k in ("value", "unit", "cores", "kind", "sample", "steps_timed")}
            out["cpu_baseline_configs0"]["config"] = "BASELINE.json configs[0]"
        print(json.dumps(out))
        return

    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl ours needs a CUDA device: chirrup_b200 has no CPU path")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    hbm_peak, tf_peak, which = peaks()
    if a.config == "engine":
        if rank != 0:
            return
        n = max(a.gpus, 1)
        er = engine_run(a.geom, n, a.bsz, a.engine_requests)
        out = dict(base)
        out["config"] = dict(config, workload=f"RWKV-7 g1 {a.geom} continuous batching through the AsyncEngineCore worker loop, {a.bsz} "
                             f"slots per worker, one worker process per GPU on one task queue, synthetic 16-64-token prompts, "
                             f"256 new tokens, greedy (BASELINE.json configs[2] as SURVEY 8(d) defines it)")
        out.update({"n_gpus": n, "value": er["value"], "ms_per_step": 1e3 * n * a.bsz / er["value"], "steps": er["tokens"] // (n * a.bsz),
                    "e2e": dict(er, h2d_bytes_per_step=0, d2h_bytes_per_step=a.bsz * 4), "engine": er,
                    "gpu_launches": (geom.n_layer * 8 + 6) * (er["tokens"] // a.bsz)})
        print(json.dumps(out))
        return
    if a.config == "migration":
        if rank != 0:
            return
        import subprocess

        r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "bench_migration.py")], capture_output=True, text=True)
        line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
        mig = json.loads(line[-1]) if line else {"unavailable": r.stderr[-500:]}
        out = dict(base)
        out["config"] = {"workload": "RWKV-7 g1 13.3B (L61, D4096) fp16 decode bsz 512/GPU with state-slot migration between two "
                                     "workers over NVLink P2P (BASELINE.json configs[4])"}
        if "unavailable" not in mig:
            v = sum(mig["decode_under_migration_tok_per_s_per_gpu"])
            out.update({"n_gpus": 2, "value": v, "ms_per_step": max(mig["decode_under_migration_ms_per_step"]),
                        "e2e": {"value": v, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": None})
        out["migration"] = mig
        print(json.dumps(out))
        return
    if a.config == "prefill":
        if rank != 0:
            return
        r = prefill_run(a, device)
        g = r["geom"]
        tok = r["B"] * r["T"]
        out = dict(base)
        out["metric"] = "RWKV-7 7.2B chunked prefill tok/s @bsz32 x 4096 tokens (BASELINE.json configs[3])"
        out["config"] = dict(config, workload=f"RWKV-7 g1 {a.geom} prefill of {r['B']} x {r['T']}-token synthetic prompts in {r['chunk']}-token "
                             f"chunks (chunked tensor-core WKV), then with a shared 2048-token prefix served from the GPU state cache "
                             f"(BASELINE.json configs[3])", batch_per_gpu=r["B"])
        out.update({"n_gpus": 1, "steps": r["steps"], "value": tok / r["t_cold"], "ms_per_step": 1e3 * r["t_cold"],
                    "e2e": {"value": tok / r["t_cold"], "unit": "tok/s", "h2d_bytes_per_step": tok * 4, "d2h_bytes_per_step": 0,
                            "api": "RWKV_x070.prefill(pinned host token ids, chunk_len=512)"},
                    "clocks": r["clocks"],
                    "prefill": {"cold_s": r["t_cold"], "sequential_wkv_kernel_s": r["t_seq"], "chunked_over_sequential": r["t_seq"] / r["t_cold"],
                                "cache_hit_s": r["t_warm"], "cache_hit_speedup": r["t_cold"] / r["t_warm"],
                                "logits_rel_cached_vs_cold": r["logits_rel"]},
                    "roofline": {"bound": "tensor", "achieved": r["flops"] / r["t_cold"] / 1e12, "peak": tf_peak, "unit": "TFLOP/s",
                                 "frac": r["flops"] / r["t_cold"] / 1e12 / tf_peak, "traffic": None, "peak_source": which,
                                 "kernel": "whole prefill pass: GEMM flops of the step over its wall time (GEMMs + chunked WKV + glue)"},
                    "gpu_launches": (g.n_layer * 8 + 4) * (r["T"] // r["chunk"]) * r["steps"]})
        print(json.dumps(out))
        return
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.distributed.init_process_group("nccl", device_id=device)
        host_group = torch.distributed.new_group(backend="gloo")   # host-side waits (an NCCL barrier would spin on the GPUs)
    r = run_ours(a, rank, world, device)
    if rank == 0:
        n = max(world, 1)
        value = n * a.bsz * a.steps / (r["ms"] * 1e-3)
        e2e = n * a.bsz * a.steps / (r["e2e_ms"] * 1e-3)
        g = r["geom"]
        # algorithmic bytes per launch (one layer, bsz units): SURVEY.md §8(d) per-unit figure for the WKV op — state read + write,
        # r,w,k,v,a,b in, y out, elapsed_t; the fused kernel on the path really reads 8 rows (r,w,k,v,a,g,vgate,v_first) and
        # writes one (stated explicitly, as §8(d) asks of fused kernels).
        state_bytes = a.bsz * 2 * g.n_head * 64 * 64 * 2
        wkv_bytes = state_bytes + a.bsz * (7 * g.n_embd * 2 + 4)
        fused_bytes = state_bytes + a.bsz * (9 * g.n_embd * 2 + 4)
        achieved = fused_bytes / r["wkv_s"]["fused"] / 1e9
        achieved_plain = wkv_bytes / r["wkv_s"]["plain"] / 1e9
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "wkv_traffic.json"))).get(f"{a.geom}_bsz{a.bsz}")
        except Exception:
            pass
        out = dict(base)
        out.update({
            "value": value, "ms_per_step": r["ms"] / a.steps,
            "e2e": {"value": e2e, "unit": "tok/s", "h2d_bytes_per_step": a.bsz * 4, "d2h_bytes_per_step": a.bsz * 4,
                    "api": "RWKV_x070.forward_seq_batch_seperate(List[List[int]] token ids, as chirrup/worker.py:690-704 passes "
                           "them) + greedy sample + D2H of the ids, host-synchronised every step"},
            "gpu_launches": r["launches"], "cuda_graph": r["graph"], "clocks": r["clocks"],
            "roofline": {"bound": "hbm", "kernel": "wkv7_fused_one_kernel (T=1): kk-normalise + k,v mix + WKV-7 state update + GroupNorm + bonus + gate",
                         "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                         "frac_of_8TBps_nominal": achieved / 8000.0, "traffic": traffic, "peak_source": which,
                         "algorithmic_bytes_per_launch": fused_bytes, "us_per_launch": r["wkv_s"]["fused"] * 1e6,
                         "share_of_step": r["wkv_s"]["fused"] * g.n_layer / (r["ms"] * 1e-3 / a.steps),
                         "plain_wkv_kernel": {"achieved": achieved_plain, "frac": achieved_plain / hbm_peak,
                                              "algorithmic_bytes_per_launch": wkv_bytes,
                                              "us_per_launch": r["wkv_s"]["plain"] * 1e6}},
        })
        step_s = r["ms"] * 1e-3 / a.steps
        wkv_in = r.get("wkv_in_step_s")
        if wkv_in:
            out["roofline"]["in_step"] = {
                "us_per_launch": wkv_in * 1e6, "achieved": fused_bytes / wkv_in / 1e9, "frac": fused_bytes / wkv_in / 1e9 / hbm_peak,
                "frac_of_8TBps_nominal": fused_bytes / wkv_in / 1e9 / 8000.0,
                "how": "step graph minus the same graph with the 32 fused-WKV launches knocked out (upper bound on the kernel's "
                       "in-step cost: the knock-out also breaks the programmatic-dependent-launch chain around it)"}
        rest_s = step_s - (wkv_in or r["wkv_s"]["fused"]) * g.n_layer
        gf = g.gemm_flops_per_token() * a.bsz
        out["roofline_gemm"] = {"bound": "tensor", "achieved": gf / max(rest_s, 1e-9) / 1e12, "peak": tf_peak, "unit": "TFLOP/s",
                                "frac": gf / max(rest_s, 1e-9) / 1e12 / tf_peak, "traffic": None, "peak_source": which + " (sustained bf16)",
                                "flops_per_step": gf, "seconds": rest_s,
                                "kernel": "GEMM phase = the step minus the in-step WKV time: tcgen05 CTA-pair GEMMs (r/k/v + LoRA group, "
                                          "LoRA-up group, output projection, FFN key, FFN value, head) with the LayerNorm/shift/mix glue "
                                          "between them counted against it"}
        if not a.no_engine:
            # the same configuration through the continuous-batching worker loop (what a Chirrup deployment runs): one worker
            # process per GPU behind MPEngineCore, spawned by rank 0 on the GPUs the ranks just used (they are idle now)
            try:
                out["engine"] = engine_run(a.geom, n, a.bsz, a.engine_requests)
                out["engine"]["fraction_of_bare_step"] = out["engine"]["value"] / value
            except Exception as e:  # noqa: BLE001
                out["engine"] = {"value": None, "error": repr(e)[:400]}
        if not a.skip_cpu_baseline and world == 1:
            try:
                out["cpu_baseline"] = cpu_baseline_subprocess(a.geom, a.bsz)
            except Exception as e:  # noqa: BLE001
                out["cpu_baseline"] = {"value": None, "unit": "tok/s", "cores": host_threads(), "kind": "reference",
                                       "sample": f"failed: {e!r}"[:400]}
        print(json.dumps(out))
    if world > 1:
        del r
        torch.cuda.empty_cache()
        torch.distributed.barrier(group=host_group)      # ranks > 0 wait here (on the host) while rank 0 runs the engine phase
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
