"""Host-side cost of one decode step's launches (7.2B, bsz 256): time spent inside RWKV_x070._step on the CPU (asynchronous
launches through the C ABI, tensor maps encoded per call) vs the GPU time of the step.  If the CPU cost approaches the GPU
time, an eager (non-graph) caller starves the GPU.  Usage (GPU box): python tools/bench_launch_cost.py [--geom 7.2B] [--bsz 256]"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--geom", default="7.2B")
    ap.add_argument("--bsz", type=int, default=256)
    ap.add_argument("--steps", type=int, default=30)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    geom, m = bench.build_model(a.geom, dev)
    B, V = a.bsz, geom.vocab_size
    state = m.generate_zero_state(B)
    tok = torch.randint(1, V, (B,)).to(torch.int32).to(dev)
    logits = torch.empty((B, V), dtype=torch.float16, device=dev)
    for _ in range(3):
        m._step(tok, B, 1, state, False, out=logits)
    torch.cuda.synchronize()
    cpu, tot = [], []
    for _ in range(a.steps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        m._step(tok, B, 1, state, False, out=logits)
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        cpu.append((t1 - t0) * 1e3)
        tot.append((t2 - t0) * 1e3)
    cpu.sort(), tot.sort()
    print(json.dumps({"geom": a.geom, "bsz": B, "cpu_ms_in_step_call_median": cpu[len(cpu) // 2],
                      "wall_ms_step_eager_median": tot[len(tot) // 2]}))


if __name__ == "__main__":
    main()
