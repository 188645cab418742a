"""GPU parity of the fused sampling kernel (C ABI b200_sample_f16) against oracle/sampler_oracle.py, the restatement of the
reference's penalty block + sample_logits_real_batch pinned by tests/golden/sampler_*.npz.

Integer / byte work is bit-exact: the penalised fp16 logits, the decayed occurrence, the penalty-state update, the greedy
tokens and the SIZE of the kept set.  The draw uses a different RNG than torch.multinomial, so it is checked the way the
reference checks its own two samplers (tests/test_sampler_equivalence.py:110-143): empirical frequencies over many draws vs
the kept distribution, tolerance 0.05 absolute here tightened to 5 sigma of the binomial error + 2e-3."""
import glob
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "sampler_*.npz")))
# mode 1: the sampler the worker imports (samplers.py:171-255), parameters recorded in the worker's fp16 dtypes
GOLD_PIP = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "samplerpip_*.npz")))


def _check_frequencies(freq, dist, values, n_draws):
    """Empirical frequencies vs the oracle's kept distribution, compared per class of EQUAL fp16 logit value: which members of
    a tie straddling the top-k / top-p boundary are kept depends on the (unspecified) order torch.sort gives equal keys, the
    class's total mass does not."""
    for b in range(freq.shape[0]):
        _, inv = torch.unique(values[b].float(), return_inverse=True)
        n = int(inv.max()) + 1
        f = torch.zeros(n).index_add_(0, inv, freq[b])
        d = torch.zeros(n).index_add_(0, inv, dist[b].float())
        assert ((f > 0) & (d == 0)).sum() == 0, f"row {b}: drew a value class outside the reference's kept set"
        sigma = torch.sqrt(d * (1 - d).clamp(min=0) / n_draws)
        assert ((f - d).abs() <= 5 * sigma + 2e-3).all(), (b, (f - d).abs().max().item())


def _exact_pip_kept(logits, temp, top_p, top_k):
    """Kept set of samplers.py:171-255 evaluated in float64 (the rule itself, free of fp32 cumsum round-off): probability >= the
    probability at the first sorted position where the cumulative sum reaches top_p, and sort rank < top_k."""
    B, V = logits.shape
    p = torch.softmax(logits.double(), -1)
    sp, si = torch.sort(p, descending=True, dim=-1)
    cum = torch.cumsum(sp, -1)
    keep = torch.zeros(B, V, dtype=torch.bool)
    for b in range(B):
        tp = float(top_p.reshape(-1)[b]) if float(temp.reshape(-1)[b]) != 0 else 0.0
        pos = int(torch.searchsorted(cum[b], torch.tensor(tp, dtype=torch.float64) * (1 - 1e-12)).clamp(max=V - 1)) if tp < 1 else V - 1
        kb = p[b] >= sp[b, pos]
        k = int(top_k.reshape(-1)[b])
        if k > 0:
            inside = torch.zeros(V, dtype=torch.bool)
            inside[si[b, :k]] = True
            kb &= inside
        keep[b] = kb
    return keep


def _load(path):
    d = np.load(path)
    return {k: torch.from_numpy(d[k]) for k in d.files}


@pytest.mark.parametrize("path", GOLD + GOLD_PIP, ids=[os.path.basename(p)[:-4] for p in GOLD + GOLD_PIP])
def test_penalties_kept_set_and_distribution(path):
    from chirrup_b200 import ops

    d = _load(path)
    mode = 1 if os.path.basename(path).startswith("samplerpip_") else 0
    B, V = d["logits"].shape
    dev = "cuda"
    flat = lambda k, dt=torch.float32: d[k].reshape(-1).to(dt).to(dev).contiguous()
    forb = d["forbidden"].to(torch.int32).to(dev).contiguous()
    dist = d["dist"]
    n_draws = 4000
    counts = torch.zeros(B, V)
    for it in range(n_draws):
        logits = d["logits"].to(dev).clone()
        occ, alpha = d["occ"].to(dev).clone(), d["alpha"].to(dev).clone()
        tok, kept = ops.sample(logits, flat("temp"), flat("top_p"), flat("top_k", torch.int32), occurrence=occ, presence=alpha,
                               penalty_decay=flat("decay"), frequency_penalty=flat("freq"), presence_penalty=flat("pres"),
                               forbidden=forb, seed=1234, step=it, return_kept=True, mode=mode)
        if it == 0:
            torch.cuda.synchronize()
            assert torch.equal(logits.cpu(), d["pen_logits"]), "penalised logits differ from the reference's decode_out"
            # kept-set size: the oracle's, except that elements whose fixed-point mass underflows (p < 2^-32 of the top
            # token) cannot be told apart from an exact cumsum tie
            want = (dist > 0).sum(1).to(torch.int32)
            if mode == 1:
                # the reference's fp32 cumsum can reach top_p a few tail tokens early (with top_p = 1 it stops where the running
                # sum first rounds to 1.0): compare with the exact (float64) evaluation of the same rule, and check that what
                # separates the reference's kept set from the exact one carries no mass
                want64 = _exact_pip_kept(d["pen_logits"], d["temp"], d["top_p"], d["top_k"])
                p64 = torch.softmax(d["pen_logits"].double(), -1)
                for b in range(B):
                    dropped = want64[b] & ~(dist[b] > 0)
                    assert float(p64[b][dropped].sum()) < 1e-5 and not bool(((dist[b] > 0) & ~want64[b]).any())
                want = want64.sum(1).to(torch.int32)
            assert (kept.cpu() - want).abs().max() <= max(2, int(0.01 * want.max())), (kept.cpu().tolist(), want.tolist())
            # penalty state after the step: decayed occurrence + 1 at the drawn token, presence set there (worker.py:527-547)
            exp_occ = d["pen_occ"].clone()
            exp_alpha = d["alpha"].clone()
            for b, t in enumerate(tok.cpu().tolist()):
                exp_occ[b, t] += 1.0
                exp_alpha[b, t] = d["pres"][b, 0]
            assert torch.equal(occ.cpu(), exp_occ) and torch.equal(alpha.cpu(), exp_alpha)
        counts[torch.arange(B), tok.cpu().long()] += 1
    if mode == 1:   # tail tokens the exact rule keeps and the reference's fp32 cumsum drops (mass < 1e-5) may legitimately be drawn
        dist = dist + 1e-12 * _exact_pip_kept(d["pen_logits"], d["temp"], d["top_p"], d["top_k"]).float()
    _check_frequencies(counts / n_draws, dist, d["pen_logits"], n_draws)


def test_greedy_rows_and_determinism():
    from chirrup_b200 import ops

    g = torch.Generator().manual_seed(5)
    B, V = 16, 65536
    logits = (torch.randn(B, V, generator=g) * 3).half().cuda()
    one = torch.ones(B, device="cuda")
    for kw in (dict(top_k=torch.ones(B, dtype=torch.int32, device="cuda")), dict(top_p=torch.zeros(B, device="cuda")),
               dict(temperature=torch.zeros(B, device="cuda"))):
        args = dict(temperature=one, top_p=one, top_k=torch.zeros(B, dtype=torch.int32, device="cuda"))
        args.update(kw)
        tok = ops.sample(logits.clone(), args["temperature"], args["top_p"], args["top_k"], seed=7, step=3)
        assert torch.equal(tok.cpu().long(), logits.float().argmax(-1).cpu())          # greedy exact (lowest index)
    a = ops.sample(logits.clone(), one, one * 0.9, torch.full((B,), 40, dtype=torch.int32, device="cuda"), seed=11, step=5)
    b = ops.sample(logits.clone(), one, one * 0.9, torch.full((B,), 40, dtype=torch.int32, device="cuda"), seed=11, step=5)
    c = ops.sample(logits.clone(), one, one * 0.9, torch.full((B,), 40, dtype=torch.int32, device="cuda"), seed=11, step=6)
    assert torch.equal(a, b) and not torch.equal(a, c)                                  # pure function of (seed, step, row)


def test_full_vocab_distribution_and_slots():
    """V = 65536 (the model's vocabulary), slot indirection, no penalties: frequencies of the top tokens vs the oracle."""
    from chirrup_b200 import ops
    from oracle import sampler_oracle as SO

    g = torch.Generator().manual_seed(9)
    B, V, S = 4, 65536, 7
    logits = (torch.randn(B, V, generator=g) * 2.5).half()
    slot = torch.tensor([5, 0, 3, 6], dtype=torch.int32)
    temp = torch.ones(S); temp[5], temp[0], temp[3], temp[6] = 1.0, 0.8, 1.3, 1.0
    top_p = torch.ones(S); top_p[5], top_p[0], top_p[3], top_p[6] = 0.9, 1.0, 0.5, 0.95
    top_k = torch.zeros(S, dtype=torch.int32); top_k[5], top_k[0], top_k[3], top_k[6] = 0, 64, 0, 200
    sel = slot.long()
    dist = SO.kept_distribution(logits, temp[sel, None], top_p[sel, None], top_k[sel, None])
    n_draws = 3000
    counts = torch.zeros(B, V)
    lg = logits.cuda()
    for it in range(n_draws):
        tok = ops.sample(lg, temp.cuda(), top_p.cuda(), top_k.cuda(), slot=slot.cuda(), seed=99, step=it)
        counts[torch.arange(B), tok.cpu().long()] += 1
    assert torch.equal(lg.cpu(), logits)                                                # no penalties: logits untouched
    _check_frequencies(counts / n_draws, dist, logits, n_draws)
