"""Micro-benchmark of the fused TMix core for prefill: chunked tensor-core kernel vs the sequential kernel, one launch of
B x T tokens per layer-sized state slab (7.2B: 32 x 512 tokens, 64 heads).  Usage (GPU box): python tools/bench_wkv_chunk.py"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from chirrup_b200 import ops


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bsz", type=int, default=32)
    ap.add_argument("--T", type=int, default=512)
    ap.add_argument("--C", type=int, default=4096)
    ap.add_argument("--slabs", type=int, default=8)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--sequential", action="store_true", help="also time the sequential kernel")
    a = ap.parse_args()
    B, T, C, L = a.bsz, a.T, a.C, a.slabs
    H = C // 64
    dev = "cuda"
    torch.manual_seed(0)
    state = (torch.randn(L, B, H, 64, 64, device=dev) * 0.1).half()
    mk = lambda s=0.5: (torch.randn(B, T, C, device=dev) * s).half()
    r, k, v, g, vg, vf = mk(), mk(), mk(), mk(), torch.sigmoid(mk(1.0)), mk()
    w = (torch.randn(B, T, C, device=dev) - 1).half()
    aa = torch.sigmoid(torch.randn(B, T, C, device=dev)).half()
    vec = lambda s=1.0: (torch.randn(C, device=dev) * s).half()
    k_k, k_a, r_k, ln_w, ln_b = vec(), vec(0.2), vec(0.1), vec(), vec(0.1)
    et = torch.zeros(B, dtype=torch.int32, device=dev)
    out = torch.empty_like(r)
    bytes_per_launch = B * (2 * H * 64 * 64 * 2) + B * T * 9 * C * 2 + B * 4

    def run(chunked):
        for l in range(L):
            ops.tmix_core_fused(state[l], r, w, k, v, aa, g, vg, vf, k_k, k_a, r_k, ln_w, ln_b, et, out=out, chunked=chunked)

    res = {}
    for name, chunked in [("chunked", True)] + ([("sequential", False)] if a.sequential else []):
        run(chunked)
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(a.iters):
            run(chunked)
        e.record()
        torch.cuda.synchronize()
        us = s.elapsed_time(e) * 1e3 / (a.iters * L)
        res[name] = dict(us_per_launch=us, GBps=bytes_per_launch / us / 1e3)
    print(json.dumps({"bsz": B, "T": T, "C": C, "bytes_per_launch": bytes_per_launch, "results": res}))


if __name__ == "__main__":
    main()
