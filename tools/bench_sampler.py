"""Times the fused sampling kernel (b200_sample_f16) on the headline shape: 256 rows x 65536 logits, with penalties.
Usage (GPU box): python tools/bench_sampler.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chirrup_b200 import ops

B, V = 256, 65536
dev = "cuda"
logits0 = (torch.randn(B, V, device=dev) * 3).half()
occ = torch.zeros(B, V, device=dev); alpha = torch.zeros(B, V, device=dev)
f = lambda v: torch.full((B,), v, device=dev)
res = {}
for name, (T, P, K, pen) in {"greedy_no_pen": (1.0, 1.0, 1, False), "top_p0.9_k50_pen": (1.0, 0.9, 50, True),
                            "top_p0.95_pen": (1.0, 0.95, 0, True), "nofilter_pen": (1.0, 1.0, 0, True)}.items():
    kw = dict(occurrence=occ, presence=alpha, penalty_decay=f(0.996), frequency_penalty=f(0.3), presence_penalty=f(0.4)) if pen else {}
    tk = torch.full((B,), K, dtype=torch.int32, device=dev)
    lg = logits0.clone()
    for i in range(3):
        ops.sample(lg, f(T), f(P), tk, seed=1, step=i, **kw)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 20
    a.record()
    for i in range(n):
        ops.sample(lg, f(T), f(P), tk, seed=1, step=10 + i, **kw)
    b.record(); torch.cuda.synchronize()
    res[name] = round(a.elapsed_time(b) * 1e3 / n, 1)
print(json.dumps({"rows": B, "V": V, "us_per_call": res}))
