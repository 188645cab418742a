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
