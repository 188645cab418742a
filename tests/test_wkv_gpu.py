"""GPU parity of the WKV-7 kernel, through the C ABI: bit-exact against the C oracle (oracle/wkv7_oracle.c) and against the
reference's own CUDA kernel built for sm_100a (oracle/_ref, /root/reference/Albatross/cuda/rwkv7_state_fwd_fp16.cu:26-97)."""
import os

import pytest
import torch

from wkv_inputs import make_wkv_inputs

pytestmark = pytest.mark.gpu

REF_SO = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libref_wkv7.so")


def _run_ours(state, r, w, k, v, a, b, et, slot_idx=None):
    from chirrup_b200 import ops

    st = state.clone()
    y = ops.wkv7_forward_seq(st, r, w, k, v, a, b, et, slot_idx)
    torch.cuda.synchronize()
    return y, st


def _frac_diff(x, y):
    return (x.view(torch.int16) != y.view(torch.int16)).float().mean().item()


@pytest.mark.parametrize("B,T,H,t0", [(1, 1, 1, 0), (3, 1, 2, 7), (2, 5, 3, (1 << 20) + 3), (4, 16, 12, 0), (64, 1, 32, 7),
                                      (2, 100, 4, 123456789)])
def test_wkv_matches_c_oracle(B, T, H, t0):
    """Bit-exact state; y and state equal to the C restatement except where MUFU.EX2/RCP differ from libm by an ulp
    (the oracle header documents this: w~ may differ by one f16 ulp on a small fraction of elements)."""
    from oracle import rwkv7_oracle as O

    ins = make_wkv_inputs(B, T, H, seed=B * 1000 + T)
    et = torch.full((B,), t0, dtype=torch.int32) + torch.arange(B, dtype=torch.int32)
    st_cpu = ins[0].clone()
    y_cpu = O.wkv_cpu(st_cpu, *ins[1:], et)
    dev = [t.cuda() for t in ins]
    y, st = _run_ours(*dev, et.cuda())
    # the decay goes through MUFU approximations on the GPU: allow a tiny fraction of 1-ulp differences, no more
    fy, fs = _frac_diff(y.cpu(), y_cpu), _frac_diff(st.cpu(), st_cpu)
    assert fy < 0.02 and fs < 0.02, (fy, fs)
    assert torch.allclose(y.cpu().float(), y_cpu.float(), rtol=4e-3, atol=4e-3)
    assert torch.allclose(st.cpu().float(), st_cpu.float(), rtol=4e-3, atol=2e-3)


@pytest.mark.skipif(not os.path.exists(REF_SO), reason="oracle/_ref/libref_wkv7.so not built")
@pytest.mark.parametrize("B,T,H,t0", [(1, 1, 1, 0), (3, 1, 2, 7), (2, 5, 3, (1 << 20) + 3), (4, 16, 12, 0), (64, 1, 32, 7),
                                      (64, 1, 32, (1 << 20) + 3), (256, 1, 64, 5), (8, 37, 64, 99), (2, 100, 4, 2 ** 31 - 50)])
def test_wkv_bit_exact_vs_reference_cuda_kernel(B, T, H, t0):
    """BASELINE.json configs[1] (1.5B geometry, bsz 64, T = 1, elapsed_t in {0, 7, 2^20+3}) and friends: every bit of y and of
    the new state equals the reference kernel's."""
    from oracle import rwkv7_oracle as O

    ins = [t.cuda() for t in make_wkv_inputs(B, T, H, seed=B * 7 + T)]
    et = (torch.full((B,), t0, dtype=torch.int64) + torch.arange(B)).to(torch.int32).cuda()
    st_ref = ins[0].clone()
    y_ref = O.wkv_ref_cuda(st_ref, *ins[1:], et)
    y, st = _run_ours(*ins, et)
    assert torch.equal(y.view(torch.int16), y_ref.view(torch.int16))
    assert torch.equal(st.view(torch.int16), st_ref.view(torch.int16))


@pytest.mark.skipif(not os.path.exists(REF_SO), reason="oracle/_ref/libref_wkv7.so not built")
def test_wkv_one_matches_reference_one_kernel():
    from chirrup_b200 import ops
    from oracle import rwkv7_oracle as O

    B, H = 5, 12
    ins = [t.cuda() for t in make_wkv_inputs(B, 1, H, seed=11)]
    et = torch.arange(B, dtype=torch.int32).cuda() * 1000
    st_ref = ins[0].clone()
    y_ref = O.wkv_one_ref_cuda(st_ref, *[t.view(B, -1) for t in ins[1:]], et)
    st = ins[0].clone()
    y = ops.wkv7_forward_one(st, *[t.view(B, -1) for t in ins[1:]], et)
    torch.cuda.synchronize()
    assert torch.equal(y.view(torch.int16), y_ref.view(torch.int16))
    assert torch.equal(st.view(torch.int16), st_ref.view(torch.int16))


def test_wkv_slot_indirection_and_views():
    """State addressed through slot_idx == gathering the slots, running, scattering back; untouched slots keep every bit.
    Also the worker's B-axis view of a larger allocation (chirrup/worker.py:697-701)."""
    from chirrup_b200 import ops

    B, T, H, n_slots = 3, 2, 4, 7
    ins = [t.cuda() for t in make_wkv_inputs(B, T, H, seed=5, n_slots=n_slots)]
    big = ins[0]
    et = torch.tensor([3, 4, 5], dtype=torch.int32).cuda()
    idx = torch.tensor([5, 0, 2], dtype=torch.int32).cuda()
    gathered = big[idx.long()].clone()
    y_a = ops.wkv7_forward_seq(gathered, *ins[1:], et)
    st_b = big.clone()
    y_b = ops.wkv7_forward_seq(st_b, *ins[1:], et, slot_idx=idx)
    torch.cuda.synchronize()
    assert torch.equal(y_a, y_b)
    assert torch.equal(st_b[idx.long()], gathered)
    mask = torch.ones(n_slots, dtype=torch.bool)
    mask[idx.long().cpu()] = False
    assert torch.equal(st_b[mask.cuda()], big[mask.cuda()])
    # view: slots 2..4 of the big tensor
    st_c = big.clone()
    y_c = ops.wkv7_forward_seq(st_c[2:5], *ins[1:], et)
    st_d = big[2:5].clone()
    y_d = ops.wkv7_forward_seq(st_d, *ins[1:], et)
    torch.cuda.synchronize()
    assert torch.equal(y_c, y_d) and torch.equal(st_c[2:5], st_d) and torch.equal(st_c[:2], big[:2]) and torch.equal(st_c[5:], big[5:])


def test_wkv_empty_and_linearity_in_v():
    """B = 0 is a no-op; with a zero state and zero `a`/`b` correction the readout is linear in v: y(2v) == 2 y(v) up to the
    fp16 subnormal grid (scaling by two commutes with every rounding except in the subnormal range)."""
    from chirrup_b200 import ops

    H = 2
    st, r, w, k, v, a, b = [t.cuda() for t in make_wkv_inputs(2, 3, H, seed=9)]
    et = torch.zeros(2, dtype=torch.int32).cuda()
    e = lambda t: t[:0]
    ops.wkv7_forward_seq(st[:0], e(r), e(w), e(k), e(v), e(a), e(b), et[:0])
    z = torch.zeros_like(a)
    s1, s2 = torch.zeros_like(st), torch.zeros_like(st)
    y1 = ops.wkv7_forward_seq(s1, r, w, k, v, z, z, et)
    y2 = ops.wkv7_forward_seq(s2, r, w, k, (v.float() * 2).half(), z, z, et)
    torch.cuda.synchronize()
    assert torch.allclose(y1.float() * 2, y2.float(), rtol=0, atol=2e-5)
    assert torch.allclose(s1.float() * 2, s2.float(), rtol=0, atol=2e-6)
