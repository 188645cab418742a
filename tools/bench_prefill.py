"""BASELINE.json configs[3]: RWKV-7 7.2B chunked prefill of 4k-token synthetic prompts, bsz 32, with state-cache reuse.
Reports prefill tok/s (cold) and the speed-up when a shared 2048-token prefix is served from the GPU-resident state cache.
Usage (GPU box): python tools/bench_prefill.py [--geom 7.2B] [--bsz 32] [--T 4096] [--prefix 2048] [--chunk 512]"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chirrup_b200 import state_move
from chirrup_b200.rwkv7 import RWKV_x070, make_args
from chirrup_b200.state_cache import StateCache
from chirrup_b200.synth import GEOMETRIES, make_checkpoint


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--geom", default="7.2B"); ap.add_argument("--bsz", type=int, default=32)
    ap.add_argument("--T", type=int, default=4096); ap.add_argument("--prefix", type=int, default=2048)
    ap.add_argument("--chunk", type=int, default=512)
    a = ap.parse_args()
    geom = GEOMETRIES[a.geom]
    dev = torch.device("cuda")
    m = RWKV_x070(make_args(), state_dict=make_checkpoint(geom, seed=0, recipe="trained_like", device=dev, dtype=torch.float16))
    g = torch.Generator().manual_seed(0)
    prefix = torch.randint(1, geom.vocab_size, (a.prefix,), generator=g)
    prompts = torch.cat([prefix.expand(a.bsz, -1), torch.randint(1, geom.vocab_size, (a.bsz, a.T - a.prefix), generator=g)], 1)
    prompts = prompts.to(torch.int32)

    def timed(fn):
        torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); r = fn(); e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e-3, r

    # warm-up (small) then cold prefill of everything
    m.prefill(prompts[:, :a.chunk], m.generate_zero_state(a.bsz), chunk_len=a.chunk)
    st = m.generate_zero_state(a.bsz)
    t_cold, lg_cold = timed(lambda: m.prefill(prompts, st, chunk_len=a.chunk))
    # fill the cache with the shared prefix's state (computed once, bsz 1), then serve every request from it
    cache = StateCache(max_gpu_entries=8)
    s1 = m.generate_zero_state(1)
    m.prefill(prompts[:1, :a.prefix], s1, chunk_len=a.chunk, want_logits=False)
    cache.put(prefix.tolist(), state_move.extract_slot(s1, 0))
    st2 = m.generate_zero_state(a.bsz)

    def warm():
        for b in range(a.bsz):
            n, packed = cache.get(prompts[b].tolist(), max_len=a.T - 1)
            assert n == a.prefix
            state_move.inject_slot(st2, b, packed)
        return m.prefill(prompts[:, a.prefix:], st2, chunk_len=a.chunk)

    t_warm, lg_warm = timed(warm)
    rel = ((lg_warm.float() - lg_cold.float()).abs().max() / lg_cold.float().abs().max()).item()
    print(json.dumps({"config": f"{a.geom} prefill bsz {a.bsz} x {a.T} tokens, chunk {a.chunk}", "prefill_tok_per_s": a.bsz * a.T / t_cold,
                      "cold_s": t_cold, "cache_hit_s": t_warm, "cache_hit_speedup": t_cold / t_warm, "shared_prefix": a.prefix,
                      "gemm_tflops_cold": geom.gemm_flops_per_token() * a.bsz * a.T / t_cold / 1e12,
                      "logits_rel_cached_vs_cold": rel, "elapsed_equal": bool(torch.equal(st[2], st2[2])),
                      "slot_bytes": state_move.slot_bytes(st)}))


if __name__ == "__main__":
    main()
