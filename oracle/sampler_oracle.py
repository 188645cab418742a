"""CPU restatement of the reference's decode SAMPLING step.  TEST INFRASTRUCTURE ONLY: imported by tests/, by
``tests/golden/make_sampler_golden.py`` and nowhere on the product path (chirrup_b200/ never imports oracle/).

Follows, op for op with the same torch calls (so that on CPU, with the same ``torch.manual_seed``, the drawn tokens are
IDENTICAL to the reference's):

* ``penalise``                 <- chirrup/worker.py:719-728  (forbidden tokens, decay, presence + frequency penalties, in place)
* ``sample_logits_real_batch`` <- chirrup/utils/samplers.py:98-168
* ``kept_distribution``        the renormalised distribution that function draws from (``sorted_probs`` after step 6, :156-160,
                               scattered back to token order) — what a sampler with a different RNG must reproduce
* ``update_penalty``           <- chirrup/worker.py:527-547 (``_batch_update_penalty``)

Pinned: ``tests/golden/make_sampler_golden.py`` imports the UNMODIFIED ``chirrup/utils/samplers.py`` from /root/reference,
asserts token equality with this file under equal seeds, and commits the vectors (tests/golden/sampler_*.npz);
``tests/test_sampler_cpu.py`` re-checks this file against them.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch
import torch.nn.functional as F


def penalise(logits: torch.Tensor, occurrence: torch.Tensor, alpha_presence: torch.Tensor, penalty_decay: torch.Tensor,
             frequency_penalty: torch.Tensor, forbidden: Optional[Sequence[Sequence[int]]] = None) -> None:
    """worker.py:719-728 on the decode rows.  ``logits`` fp16 (B,V) and ``occurrence`` fp32 (B,V) are updated IN PLACE;
    ``penalty_decay`` / ``frequency_penalty`` are (B,1) fp32 like the worker's tensors."""
    if forbidden is not None:
        for b, toks in enumerate(forbidden):                       # :719-722
            for t in toks:
                logits[b][t] -= 1e10
    occurrence *= penalty_decay                                    # :725
    logits -= alpha_presence + occurrence * frequency_penalty      # :726-729 (fp32 right-hand side, fp16 in-place result)


def _filtered_sorted(logits, temperature, top_p, top_k, min_tokens_to_keep: int = 1):
    bsz, vocab_size = logits.shape
    logits = logits / temperature                                  # :113
    probs = F.softmax(logits.float(), dim=-1)                      # :116
    sorted_probs, sorted_indices = torch.sort(probs, descending=True, dim=-1)  # :119
    clamped_top_k = torch.clamp(top_k, min=min_tokens_to_keep, max=vocab_size).long()             # :125
    effective_top_k = torch.where(top_k > 0, clamped_top_k, torch.full_like(clamped_top_k, vocab_size))  # :126
    ranks = torch.arange(vocab_size, device=logits.device).expand(bsz, vocab_size)                # :129
    topk_remove_mask = ranks >= effective_top_k                    # :132
    effective_top_p = torch.where(top_p < 1.0, top_p, torch.full_like(top_p, 2.0))                # :136
    cumulative_probs = torch.cumsum(sorted_probs, dim=-1)          # :139
    topp_remove_mask = cumulative_probs > effective_top_p          # :142
    min_keep_mask = ranks < min_tokens_to_keep                     # :145
    topp_remove_mask = topp_remove_mask & (~min_keep_mask)         # :146
    combined_remove_mask = topk_remove_mask | topp_remove_mask     # :150
    sorted_probs = sorted_probs.masked_fill(combined_remove_mask, 0.0)                            # :153
    probs_sum = sorted_probs.sum(dim=-1, keepdim=True)             # :156
    sorted_probs = sorted_probs / torch.where(probs_sum > 0, probs_sum, torch.ones_like(probs_sum))  # :157
    return sorted_probs, sorted_indices


def sample_logits_real_batch(logits: torch.Tensor, temperature: torch.Tensor, top_p: torch.Tensor, top_k: torch.Tensor,
                             min_tokens_to_keep: int = 1) -> torch.Tensor:
    """samplers.py:98-168.  temperature / top_p / top_k are (B,1) tensors like the worker's."""
    sorted_probs, sorted_indices = _filtered_sorted(logits, temperature, top_p, top_k, min_tokens_to_keep)
    sampled_sorted_idx = torch.multinomial(sorted_probs, num_samples=1)           # :160
    return torch.gather(sorted_indices, dim=1, index=sampled_sorted_idx).squeeze(-1)  # :163


def kept_distribution(logits: torch.Tensor, temperature: torch.Tensor, top_p: torch.Tensor, top_k: torch.Tensor) -> torch.Tensor:
    """(B,V) fp32: probability of each token under ``sample_logits_real_batch`` (zero outside the kept set)."""
    sorted_probs, sorted_indices = _filtered_sorted(logits, temperature, top_p, top_k)
    return torch.zeros_like(sorted_probs).scatter_(1, sorted_indices, sorted_probs)


def update_penalty(occurrence: torch.Tensor, alpha_presence: torch.Tensor, presence_penalty: torch.Tensor, slots: List[int],
                   tokens: List[int], no_penalty_token_ids: Sequence[int] = ()) -> None:
    """worker.py:506-510 (weight 0 for no-penalty ids) + :527-547."""
    if not slots:
        return
    s = torch.tensor(slots, dtype=torch.long)
    t = torch.tensor(tokens, dtype=torch.long)
    w = torch.tensor([0.0 if tok in no_penalty_token_ids else 1.0 for tok in tokens], dtype=torch.float32)
    occurrence[s, t] += w                                          # :541
    alpha_presence[s, t] = presence_penalty[s, 0]                  # :546-547


# ---------------------------------------------------------------------------------------------------------------------------
# The sampler the worker ACTUALLY imports: ``from chirrup.utils.samplers import sample_logits_rwkv_pip_compatible as
# sample_logits_real_batch`` (chirrup/worker.py:12) — chirrup/utils/samplers.py:171-255.  Different semantics from the
# single-sort sampler above: softmax WITHOUT temperature, top-p by cutoff VALUE (every token whose probability is >= the
# probability at the first sorted position where the cumulative sum reaches top_p stays: ties at the cutoff are all kept),
# top-k on the sort rank, temperature applied AFTER the filters as probs ** (1/T), multinomial.
def _pip_filtered(logits, temperature, top_p, top_k):
    bsz, vocab_size = logits.shape
    zero_temp_mask = temperature == 0                                              # :193
    temperature = torch.where(zero_temp_mask, torch.ones_like(temperature), temperature)   # :194
    top_p = torch.where(zero_temp_mask, torch.zeros_like(top_p), top_p)            # :195
    probs = F.softmax(logits.float(), dim=-1)                                      # :198
    sorted_probs, sorted_indices = torch.sort(probs, descending=True, dim=-1)      # :201
    cumulative_probs = torch.cumsum(sorted_probs, dim=-1)                          # :204
    cutoff_positions = torch.searchsorted(cumulative_probs, top_p)                 # :209
    cutoff_positions = torch.clamp(cutoff_positions, max=vocab_size - 1)           # :210
    cutoff_values = torch.gather(sorted_probs, dim=1, index=cutoff_positions)      # :213
    probs = torch.where(probs < cutoff_values, torch.zeros_like(probs), probs)     # :216
    top_k_long = top_k.long()                                                      # :219
    active_top_k = top_k_long > 0                                                  # :220
    if active_top_k.any():                                                         # :222
        effective_k = torch.where(active_top_k, top_k_long, torch.full_like(top_k_long, vocab_size))   # :225-227
        ranks = torch.arange(vocab_size, device=logits.device).expand(bsz, vocab_size)                 # :229
        outside_topk_sorted = ranks >= effective_k                                 # :231
        outside_topk = torch.zeros_like(probs, dtype=torch.bool).scatter_(dim=1, index=sorted_indices, src=outside_topk_sorted)  # :234-236
        probs.masked_fill_(outside_topk, 0.0)                                      # :238
    need_temp = temperature != 1.0                                                 # :241
    if need_temp.any():                                                            # :242
        inv_temp = 1.0 / temperature                                               # :244
        adjusted_probs = probs ** inv_temp                                         # :247
        probs = torch.where(need_temp.expand_as(probs), adjusted_probs, probs)     # :248
    return probs


def sample_logits_rwkv_pip_compatible(logits: torch.Tensor, temperature: torch.Tensor, top_p: torch.Tensor,
                                      top_k: torch.Tensor) -> torch.Tensor:
    """samplers.py:171-255; temperature / top_p (B,1) in the worker's dtypes (fp16, worker.py:611-620), top_k (B,1) int32."""
    return torch.multinomial(_pip_filtered(logits, temperature, top_p, top_k), num_samples=1).squeeze(-1)    # :253


def kept_distribution_pip(logits: torch.Tensor, temperature: torch.Tensor, top_p: torch.Tensor, top_k: torch.Tensor) -> torch.Tensor:
    """(B,V) fp32: the normalised distribution ``sample_logits_rwkv_pip_compatible`` draws from."""
    p = _pip_filtered(logits, temperature, top_p, top_k)
    return p / p.sum(dim=-1, keepdim=True)


def worker_sampling_params(temperature: float, top_p: float, top_k: int, frequency_penalty: float, penalty_decay: float,
                           presence_penalty: float):
    """The per-slot tensors the worker builds at admission (chirrup/worker.py:611-640): temperature (<= 0 becomes 1.0), top_p,
    frequency penalty and decay as FP16, presence penalty as fp32, top_k int32 — each (1,1)."""
    h = lambda v: torch.tensor([[v]], dtype=torch.float16)
    return dict(temperature=h(temperature if temperature > 0 else 1.0), top_p=h(top_p),
                top_k=torch.tensor([[top_k]], dtype=torch.int32), frequency_penalty=h(frequency_penalty),
                penalty_decay=h(penalty_decay), presence_penalty=torch.tensor([[presence_penalty]], dtype=torch.float32))
