"""Debug aid: run a 1-layer model and compare the workspace intermediates with the oracle's taps (GPU box)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chirrup_b200 import _lib
from chirrup_b200.rwkv7 import RWKV_x070, make_args
from chirrup_b200.synth import GEOMETRIES, make_checkpoint
from oracle import rwkv7_oracle as O

geom_name, B, T, L = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
geom = GEOMETRIES[geom_name]
raw = make_checkpoint(geom, seed=2, n_layer=L)
m = RWKV_x070(make_args(), state_dict={k: v.clone() for k, v in raw.items()})
z = O.transform_checkpoint({k: v.cuda() for k, v in raw.items()}, device="cuda")
og = O.Geometry(z)
toks = torch.randint(1, geom.vocab_size, (B, T), generator=torch.Generator().manual_seed(1)).tolist()
sa, sb = m.generate_zero_state(B), O.zero_state(og, B, "cuda")
taps = []
la = m.forward_seq_batch(toks, sa)
lb = O.forward_seq_batch(z, og, toks, sb, wkv="ref_cuda", taps=taps)
torch.cuda.synchronize()
M, C = B * T, geom.n_embd
al = lambda x: (x + 255) // 256 * 256
off = al(_lib.lib().b200_linear_workspace_bytes(M, geom.vocab_size))
names = ["x", "xr", "xw", "xk", "xv", "xa", "xg", "r", "k", "v", "w", "a", "g", "vg", "k2", "v2", "an", "bn", "y", "yg", "vfirst", "kx", "xo"]
bufs = {}
for n in names:
    bufs[n] = m._ws[off:off + M * C * 2].view(torch.float16).view(M, C)
    off += al(M * C * 2)
def rel(a, b):
    return ((a.float() - b.float()).abs().max() / b.float().abs().max()).item()
t = taps[-1]
pairs = [("r", t["r"]), ("w", t["w"]), ("k2", t["k"]), ("v2" if L > 1 else "vfirst", t["v"]), ("an", t["a"]), ("bn", t["b"]), ("g", t["g"])]
for n, ref in pairs:
    print(f"{n:8s} rel {rel(bufs[n], ref.reshape(M, C)):.3e}")
print("logits rel", rel(la, lb), "state1 rel", rel(sa[1], sb[1]), "state0 rel", rel(sa[0], sb[0]))
for n in names:
    b = bufs[n].float()
    print(f"  buf {n:7s} nan {int(torch.isnan(b).sum()):6d} absmax {b[~torch.isnan(b)].abs().max().item() if (~torch.isnan(b)).any() else -1:.3e}")
nanrows = torch.isnan(bufs["vg"].float()).any(1).nonzero().flatten().tolist()
nancols = torch.isnan(bufs["vg"].float()).any(0).nonzero().flatten().tolist()
print("vg nan rows", nanrows[:40], "cols", nancols[:40], len(nancols))
