"""How far apart are two legitimate evaluations of the SAME reference math on this GPU?  Compares, on one synthetic checkpoint:
  ours            vs oracle (torch fp16 ops / cuBLAS + reference WKV kernel)
  oracle fp32-GEMM (sgemm, one rounding) vs oracle cuBLAS-fp16            <- GEMM summation-order noise
  oracle batched  vs oracle one sequence at a time                         <- the reference's own batch non-invariance
Prints max|a-b|/max|b| for logits and WKV state.  (GPU box; test infrastructure.)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chirrup_b200.rwkv7 import RWKV_x070, make_args
from chirrup_b200.synth import GEOMETRIES, make_checkpoint
from oracle import rwkv7_oracle as O

geom_name, B, P, D = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
recipe = sys.argv[5] if len(sys.argv) > 5 else "plain"
n_layer = int(sys.argv[6]) if len(sys.argv) > 6 else None
geom = GEOMETRIES[geom_name]
raw = make_checkpoint(geom, seed=21, recipe=recipe, n_layer=n_layer)
m = RWKV_x070(make_args(), state_dict={k: v.clone() for k, v in raw.items()})
z = O.transform_checkpoint({k: v.cuda() for k, v in raw.items()}, device="cuda")
og = O.Geometry(z)
prompt = torch.randint(1, geom.vocab_size, (B, P), generator=torch.Generator().manual_seed(5)).tolist()
rel = lambda a, b: ((a.float() - b.float()).abs().max() / b.float().abs().max()).item()

def run(kind):
    if kind == "ours":
        st = m.generate_zero_state(B); f = lambda t, s: m.forward_seq_batch(t, s)
    elif kind == "cublas":
        st = O.zero_state(og, B, "cuda"); f = lambda t, s: O.forward_seq_batch(z, og, t, s, wkv="ref_cuda")
    elif kind == "fp32gemm":
        st = O.zero_state(og, B, "cuda"); f = lambda t, s: O.forward_seq_batch(z, og, t, s, wkv="ref_cuda", fast=True)
    elif kind == "single":
        st = O.zero_state(og, B, "cuda")
        def f(t, s):
            outs = []
            for i in range(B):
                sub = [s[0][:, :, i:i+1], s[1][:, i:i+1], s[2][i:i+1]]
                outs.append(O.forward_seq_batch(z, og, [t[i]], sub, wkv="ref_cuda"))
            return torch.cat(outs, 0)
    outs = [f(prompt, st)]
    return outs, st, f

ref_out, ref_st, ref_f = run("cublas")
runs = {k: run(k) for k in ("ours", "fp32gemm", "single")}
print(f"{geom_name} B{B} P{P} D{D} {recipe}: prefill logits rel:", {k: f"{rel(v[0][0], ref_out[0]):.2e}" for k, v in runs.items()})
for step in range(D):
    nxt = [[int(t)] for t in ref_out[-1].float().argmax(-1)]
    ref_out.append(ref_f(nxt, ref_st))
    for k, (o, s, f) in runs.items():
        o.append(f(nxt, s))
print("  last logits rel:", {k: f"{rel(v[0][-1], ref_out[-1]):.2e}" for k, v in runs.items()})
print("  state1 rel:", {k: f"{rel(v[1][1], ref_st[1]):.2e}" for k, v in runs.items()})
print("  argmax agree (all steps):", {k: sum(int((a.float().argmax(-1) == b.float().argmax(-1)).sum()) for a, b in zip(v[0], ref_out)) for k, v in runs.items()}, "of", B * (D + 1))
lg = ref_out[-1].float(); top2 = lg.topk(2, -1).values
print("  ref logits std %.3f, top1-top2 margin min %.3f mean %.3f" % (lg.std().item(), (top2[:, 0] - top2[:, 1]).min().item(), (top2[:, 0] - top2[:, 1]).mean().item()))
