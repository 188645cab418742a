"""Micro-benchmark of the tcgen05 linear op on the decode-step shapes: back-to-back launches in one CUDA graph, each launch
streaming a DIFFERENT weight matrix (n_w of them, > L2 in total), so the figure is the warm in-stream cost per launch that the
model step actually pays (ncu's per-launch times are cold and serialised).
--cublas adds, beside each of our launches, the SAME-BOX library bar SURVEY §2a names: torch.nn.functional.linear (cuBLAS /
cuBLASLt fp16, what Albatross/rwkv7.py:625-630,649,678-679,551 dispatches to) on the same operands in the same graph-replay
harness — the plain GEMM only; the reference runs its bias / activation / residual as further ATen launches on top of it.
Usage (GPU box): python tools/bench_gemm.py [--M 256] [--C 4096] [--nw 24] [--cublas] [--json out.json]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from chirrup_b200 import ops


def time_graph(fn, iters=5):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            fn()
        g.replay()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(s)
        for _ in range(iters):
            g.replay()
        b.record(s)
        torch.cuda.synchronize()
    return a.elapsed_time(b) / iters  # ms per fn()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--M", type=int, default=256)
    ap.add_argument("--C", type=int, default=4096)
    ap.add_argument("--nw", type=int, default=24)
    ap.add_argument("--json", default=None)
    ap.add_argument("--shapes", default="all")
    ap.add_argument("--cublas", action="store_true")
    a = ap.parse_args()
    M, C, F = a.M, a.C, 4 * a.C
    dev = "cuda"
    shapes = [("cxc", C, C, ops.EPI_NONE), ("cxc_resadd", C, C, ops.EPI_RESADD), ("ffn_key", F, C, ops.EPI_RELU_SQ),
              ("ffn_value", C, F, ops.EPI_RESADD), ("lora_down_128", 128, C, ops.EPI_TANH), ("lora_down_480", 480, C, ops.EPI_SIGMOID),
              ("lora_down_96", 96, C, ops.EPI_NONE), ("lora_up_128", C, 128, ops.EPI_SIGMOID), ("lora_up_96", C, 96, ops.EPI_SIGMOID),
              ("lora_up_480", C, 480, ops.EPI_NONE), ("head", 65536, C, ops.EPI_NONE)]
    out = []
    for name, N, K, epi in shapes:
        if a.shapes != "all" and name not in a.shapes.split(","):
            continue
        nw = a.nw if N * K >= C * C else 32
        if N == 65536:
            nw = 2
        Ws = [(torch.randn(N, K, device=dev) / K ** 0.5).half() for _ in range(nw)]
        x = torch.randn(M, K, device=dev).half()
        res = torch.randn(M, N, device=dev).half() if epi == ops.EPI_RESADD else None
        bias = torch.randn(N, device=dev).half() if epi in (ops.EPI_TANH, ops.EPI_SIGMOID) else None
        y = torch.empty(M, N, device=dev, dtype=torch.float16)

        def fn():
            for W in Ws:
                ops.linear(x, W, bias, epi, res, out=y)

        ms = time_graph(fn)
        us = ms * 1e3 / nw
        rec = {"shape": name, "M": M, "N": N, "K": K, "us_per_launch": round(us, 2),
               "tflops": round(2.0 * M * N * K / us / 1e6, 1), "weight_GBps": round(N * K * 2 / us / 1e3, 1)}
        if a.cublas:
            def fn_lib():
                for W in Ws:
                    torch.nn.functional.linear(x, W)

            us_lib = time_graph(fn_lib) * 1e3 / nw
            rec.update({"cublas_us_per_launch": round(us_lib, 2), "cublas_tflops": round(2.0 * M * N * K / us_lib / 1e6, 1),
                        "ours_over_cublas_time": round(us / us_lib, 3)})
        print(json.dumps(rec), flush=True)
        out.append(rec)
        del Ws
    if a.json:
        json.dump(out, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
