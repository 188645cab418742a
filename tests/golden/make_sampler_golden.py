"""Generates tests/golden/sampler_*.npz from the UNMODIFIED reference sampler (chirrup/utils/samplers.py, loaded by file path
from /root/reference) on CPU, and pins oracle/sampler_oracle.py against it: under the same torch seed the oracle must draw
the SAME tokens, and the penalty step must leave bit-identical logits / occurrence.  Build container only
(``python tests/golden/make_sampler_golden.py``); the vectors are committed because /root/reference is absent on the GPU box.
"""
import importlib.util
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import sampler_oracle as SO

REF = os.environ.get("REF_ROOT", "/root/reference")
OUT = os.path.dirname(os.path.abspath(__file__))

spec = importlib.util.spec_from_file_location("ref_samplers", os.path.join(REF, "chirrup", "utils", "samplers.py"))
R = importlib.util.module_from_spec(spec)
spec.loader.exec_module(R)

CASES = [  # name, B, V, logit scale, temperature, top_p, top_k
    ("nofilter", 4, 512, 3.0, 1.0, 1.0, 0),
    ("topk50_p09", 4, 512, 3.0, 1.0, 0.9, 50),
    ("lowtemp_p05", 4, 1024, 2.0, 0.5, 0.5, 0),
    ("hightemp_k10", 4, 1024, 4.0, 2.0, 1.0, 10),
    ("greedy_k1", 4, 512, 3.0, 1.0, 1.0, 1),
    ("tiny_p", 4, 512, 3.0, 1.0, 0.01, 0),
    ("mixed_rows", 6, 2048, 3.0, None, None, None),
]


def main():
    for name, B, V, scale, T, P, K in CASES:
        g = torch.Generator().manual_seed(hash(name) % (1 << 31) if False else sum(map(ord, name)))
        logits = (torch.randn(B, V, generator=g) * scale).half()
        if T is None:  # per-row parameters
            temp = torch.tensor([[1.0], [0.7], [1.5], [1.0], [0.3], [1.0]])
            top_p = torch.tensor([[1.0], [0.8], [0.95], [0.3], [1.0], [0.6]])
            top_k = torch.tensor([[0], [20], [0], [5], [100], [1]], dtype=torch.int32)
        else:
            temp = torch.full((B, 1), T)
            top_p = torch.full((B, 1), P)
            top_k = torch.full((B, 1), K, dtype=torch.int32)
        # penalty state of a short history (worker.py:527-547 applied a few times), then one decode step (:719-728)
        occ = torch.zeros(B, V)
        alpha = torch.zeros(B, V)
        decay = torch.full((B, 1), 0.996)
        freq = torch.full((B, 1), 0.3)
        pres = torch.full((B, 1), 0.4)
        hist = torch.randint(0, V, (B, 12), generator=g)
        for b in range(B):
            for t in hist[b].tolist():
                occ[b, t] += 1.0
                alpha[b, t] = pres[b, 0]
        forbidden = [[int(x) for x in torch.randint(0, V, (2,), generator=g)] for _ in range(B)]
        # --- the reference's own ops (worker.py:719-728 transcribed in oracle.penalise is checked against this literal form)
        ref_logits, ref_occ = logits.clone(), occ.clone()
        for b in range(B):
            for t in forbidden[b]:
                ref_logits[b][t] -= 1e10
        ref_occ[:, :] *= decay
        ref_logits -= (alpha + ref_occ * freq)
        o_logits, o_occ = logits.clone(), occ.clone()
        SO.penalise(o_logits, o_occ, alpha, decay, freq, forbidden)
        assert torch.equal(o_logits, ref_logits) and torch.equal(o_occ, ref_occ)
        # --- the sampler: same seed, same tokens
        draws = []
        for seed in range(8):
            torch.manual_seed(seed)
            ref_tok = R.sample_logits_real_batch(ref_logits.clone(), temp, top_p, top_k)
            torch.manual_seed(seed)
            my_tok = SO.sample_logits_real_batch(o_logits.clone(), temp, top_p, top_k)
            assert torch.equal(ref_tok, my_tok), (name, seed)
            draws.append(ref_tok.numpy())
        dist = SO.kept_distribution(o_logits, temp, top_p, top_k)
        # the reference draws only from the oracle's kept set
        for d in draws:
            assert all(dist[b, d[b]] > 0 for b in range(B))
        np.savez_compressed(os.path.join(OUT, f"sampler_{name}.npz"), logits=logits.numpy(), occ=occ.numpy(), alpha=alpha.numpy(),
                            decay=decay.numpy(), freq=freq.numpy(), pres=pres.numpy(), forbidden=np.array(forbidden, dtype=np.int32),
                            temp=temp.numpy(), top_p=top_p.numpy(), top_k=top_k.numpy(), pen_logits=ref_logits.numpy(),
                            pen_occ=ref_occ.numpy(), draws=np.stack(draws), dist=dist.numpy())
        print(name, "ok; kept per row:", (dist > 0).sum(1).tolist())


def main_pip():
    """The sampler the worker imports (chirrup/worker.py:12 -> samplers.py:171-255) with the parameter tensors in the dtypes the
    worker builds at admission (worker.py:611-640: temperature / top_p / frequency penalty / decay FP16, presence fp32)."""
    for name, B, V, scale, T, P, K in CASES:
        g = torch.Generator().manual_seed(sum(map(ord, name)) + 7)
        logits = (torch.randn(B, V, generator=g) * scale).half()
        if T is None:
            temp = torch.tensor([[1.0], [0.7], [1.5], [1.0], [0.3], [1.0]])
            top_p = torch.tensor([[1.0], [0.8], [0.95], [0.3], [1.0], [0.6]])
            top_k = torch.tensor([[0], [20], [0], [5], [100], [1]], dtype=torch.int32)
        else:
            temp, top_p, top_k = torch.full((B, 1), T), torch.full((B, 1), P), torch.full((B, 1), K, dtype=torch.int32)
        temp, top_p = temp.half(), top_p.half()
        occ, alpha = torch.zeros(B, V), torch.zeros(B, V)
        decay, freq, pres = torch.full((B, 1), 0.996).half(), torch.full((B, 1), 0.3).half(), torch.full((B, 1), 0.4)
        hist = torch.randint(0, V, (B, 12), generator=g)
        for b in range(B):
            for t in hist[b].tolist():
                occ[b, t] += 1.0
                alpha[b, t] = pres[b, 0]
        forbidden = [[int(x) for x in torch.randint(0, V, (2,), generator=g)] for _ in range(B)]
        ref_logits, ref_occ = logits.clone(), occ.clone()
        for b in range(B):
            for t in forbidden[b]:
                ref_logits[b][t] -= 1e10
        ref_occ[:, :] *= decay                                   # worker.py:725 with its fp16 decay tensor
        ref_logits -= (alpha + ref_occ * freq)                   # :726-729 with its fp16 frequency tensor
        o_logits, o_occ = logits.clone(), occ.clone()
        SO.penalise(o_logits, o_occ, alpha, decay, freq, forbidden)
        assert torch.equal(o_logits, ref_logits) and torch.equal(o_occ, ref_occ)
        draws = []
        for seed in range(8):
            torch.manual_seed(seed)
            ref_tok = R.sample_logits_rwkv_pip_compatible(ref_logits.clone(), temp, top_p, top_k)
            torch.manual_seed(seed)
            my_tok = SO.sample_logits_rwkv_pip_compatible(o_logits.clone(), temp, top_p, top_k)
            assert torch.equal(ref_tok, my_tok), (name, seed)
            draws.append(ref_tok.numpy())
        dist = SO.kept_distribution_pip(o_logits, temp, top_p, top_k)
        for d in draws:
            assert all(dist[b, d[b]] > 0 for b in range(B))
        np.savez_compressed(os.path.join(OUT, f"samplerpip_{name}.npz"), logits=logits.numpy(), occ=occ.numpy(), alpha=alpha.numpy(),
                            decay=decay.numpy(), freq=freq.numpy(), pres=pres.numpy(), forbidden=np.array(forbidden, dtype=np.int32),
                            temp=temp.numpy(), top_p=top_p.numpy(), top_k=top_k.numpy(), pen_logits=ref_logits.numpy(),
                            pen_occ=ref_occ.numpy(), draws=np.stack(draws), dist=dist.numpy())
        print("pip", name, "ok; kept per row:", (dist > 0).sum(1).tolist())


if __name__ == "__main__":
    import sys
    if "--pip-only" not in sys.argv:
        main()
    main_pip()
