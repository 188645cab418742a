"""Seeded WKV-7 operand generator shaped like what TMix produces: a = -kk, b = kk*sigmoid(.)
with kk unit-norm per head — arbitrary a,b make the recurrence blow up."""
import torch


def make_wkv_inputs(B, T, H, seed=0, device="cpu", n_slots=None, state_std=0.1):
    g = torch.Generator().manual_seed(seed)
    C = H * 64
    n_slots = n_slots or B

    def rn(*s, std=1.0, mean=0.0):
        return torch.randn(*s, generator=g) * std + mean

    r = rn(B, T, C, std=0.5)
    k = rn(B, T, C, std=0.5)
    v = rn(B, T, C, std=0.5)
    w = rn(B, T, C, mean=-1.0)
    kk = torch.nn.functional.normalize(rn(B, T, H, 64), dim=-1).view(B, T, C)
    a = -kk
    b = kk * torch.sigmoid(rn(B, T, C))
    state = rn(n_slots, H, 64, 64, std=state_std)
    out = [t.to(torch.float16).to(device) for t in (state, r, w, k, v, a, b)]
    return out
