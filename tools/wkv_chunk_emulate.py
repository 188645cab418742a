"""CPU emulation of the CHUNKED WKV-7 prefill algorithm (csrc/wkv7_chunk.cu) with the kernel's operand roundings, compared with
the sequential fp16 recurrence of the reference (oracle/wkv7_oracle.c).  Development aid: validates the algebra, the mid-chunk
rescaling and the fp16-operand / fp32-accumulate error level before any CUDA is written.

Recurrence per head (N = 64; Albatross/cuda/rwkv7_state_fwd_fp16.cu:56-82), S[i][j] (i value row, j key column):
    d_t = 1 + w~_t                                   (w~ = fp16 decay-1 with the elapsed_t dither, .cu:20-23, :59)
    u_t = S_{t-1} a_t                                ("sa")
    S_t = S_{t-1} diag(d_t) + u_t b_t^T + v_t k_t^T
    y_t = S_t r_t
Chunk of T tokens from S_0, with c_t = sum_{s<=t} log d_s (per channel), reference point m (per channel):
    a~_t = a_t exp(c_{t-1} - m)    r~_t = r_t exp(c_t - m)    b~_s = b_s exp(m - c_s)    k~_s = k_s exp(m - c_s)    S0' = S_0 diag(exp(m))
    U = (S0' A~ + V Aak) (I - Aab)^-1       Aab[s,t] = b~_s . a~_t (s < t)   Aak[s,t] = k~_s . a~_t (s < t)
    Y = S0' R~ + U Abr + V Akr               Abr[s,t] = b~_s . r~_t (s <= t)  Akr[s,t] = k~_s . r~_t (s <= t)
    S_T = S_0 diag(exp(c_T)) + (U B~^T + V K~^T) diag(exp(c_T - m))
Usage: python tools/wkv_chunk_emulate.py [--T 16] [--tokens 512] [--heads 4]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch

from oracle import rwkv7_oracle as O
from wkv_inputs import make_wkv_inputs

h16 = lambda x: x.to(torch.float16).float()      # one rounding to binary16, kept in fp32 containers


def decay(w, t):
    """d = 1 + w~ exactly as the kernel computes w~ (fp32 math, one fp16 rounding); w (..., C) fp16, t int32 scalar per row."""
    import ctypes
    out = torch.empty_like(w)
    lib = O._c_oracle()
    flat_w, flat_o = w.contiguous().view(-1, w.shape[-1]), out.view(-1, w.shape[-1])
    for i in range(flat_w.shape[0]):
        lib.wkv7_oracle_decay(flat_w.shape[1], flat_w[i].data_ptr(), int(t[i]), flat_o[i].data_ptr())
    return 1.0 + out.float()


def chunked(state, r, w, k, v, a, b, elapsed, Tc=16, operand=h16, state_operand=h16, tensor_solve=False, small=lambda x: x):
    """state (B,H,64,64) fp16 in/out (rounded to fp16 only at the END: the kernel keeps it fp32 on chip across chunks);
    r.. (B,T,C) fp16.  Returns y (B,T,C) fp32."""
    B, T, C = r.shape
    H = C // 64
    tt = (elapsed.view(B, 1) + torch.arange(T).view(1, T)).to(torch.int32)
    d = decay(w.view(B * T, C), tt.reshape(-1)).view(B, T, H, 64)
    f = lambda x: x.float().view(B, T, H, 64)
    r_, k_, v_, a_, b_ = f(r), f(k), f(v), f(a), f(b)
    S = state.float().clone()                                  # (B,H,i,j)
    y = torch.empty(B, T, H, 64)
    for t0 in range(0, T, Tc):
        t1 = min(T, t0 + Tc)
        n = t1 - t0
        ld = torch.log(d[:, t0:t1])                            # (B,n,H,64)
        c = torch.cumsum(ld, dim=1)                            # inclusive
        cprev = c - ld
        m = c[:, (n - 1) // 2]                                 # reference point: the middle of the chunk (B,H,64)
        mm = m.unsqueeze(1)
        At = operand(a_[:, t0:t1] * torch.exp(cprev - mm))     # (B,n,H,64)
        Rt = operand(r_[:, t0:t1] * torch.exp(c - mm))
        Bt = operand(b_[:, t0:t1] * torch.exp(mm - c))
        Kt = operand(k_[:, t0:t1] * torch.exp(mm - c))
        V = v_[:, t0:t1]                                       # exact fp16 values
        S0p = state_operand(S * torch.exp(m).unsqueeze(2))     # columns j scaled
        ein = torch.einsum
        Aab = ein("bshj,bthj->bhst", Bt, At)
        Aak = ein("bshj,bthj->bhst", Kt, At)
        Abr = ein("bshj,bthj->bhst", Bt, Rt)
        Akr = ein("bshj,bthj->bhst", Kt, Rt)
        strict = torch.triu(torch.ones(n, n), diagonal=1)
        incl = torch.triu(torch.ones(n, n), diagonal=0)
        Aab, Aak, Abr, Akr = Aab * strict, Aak * strict, Abr * incl, Akr * incl
        SA = ein("bhij,bthj->bhit", S0p, At)                   # (B,H,i,t)
        SR = ein("bhij,bthj->bhit", S0p, Rt)
        Vt = V.permute(0, 2, 3, 1)                             # (B,H,i,s)
        rhs = SA + ein("bhis,bhst->bhit", Vt, small(Aak))
        if not tensor_solve:
            U = torch.zeros_like(rhs)
            for t in range(n):                                 # forward substitution, fp32
                U[..., t] = rhs[..., t] + (U[..., :t] * Aab[:, :, :t, t].unsqueeze(2)).sum(-1)
        else:                                                  # Tinv = (I - Aab)^-1 in fp32 (one warp), then U = rhs Tinv on the tensor cores
            Tinv = torch.linalg.inv(torch.eye(n) - Aab)
            U = ein("bhis,bhst->bhit", operand(rhs), small(Tinv))
        Uo = operand(U)
        Y = SR + ein("bhis,bhst->bhit", Uo if tensor_solve else U, small(Abr)) + ein("bhis,bhst->bhit", Vt, small(Akr))
        y[:, t0:t1] = Y.permute(0, 3, 1, 2)
        upd = ein("bhis,bshj->bhij", Uo, Bt) + ein("bhis,bshj->bhij", Vt, Kt)
        cT = c[:, -1]
        S = S * torch.exp(cT).unsqueeze(2) + upd * torch.exp(cT - m).unsqueeze(2)
    state.copy_(S.to(torch.float16))
    return y.reshape(B, T, C)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--T", type=int, default=16)
    ap.add_argument("--tokens", type=int, default=512)
    ap.add_argument("--heads", type=int, default=4)
    ap.add_argument("--B", type=int, default=2)
    a = ap.parse_args()
    ins = make_wkv_inputs(a.B, a.tokens, a.heads, seed=3)
    et = torch.tensor([0, (1 << 20) + 3][:a.B], dtype=torch.int32)
    st_ref = ins[0].clone()
    y_ref = O.wkv_cpu(st_ref, *ins[1:], et).float()
    rel = lambda x, z: ((x - z).norm() / z.norm()).item()
    relmax = lambda x, z: ((x - z).abs().max() / z.abs().max()).item()
    ident = lambda x: x
    for name, op, sop, ts, sm in (("fp32 everywhere (algebra check)", ident, ident, False, ident),
                                  ("fp16 operands, fp32 accumulate, scalar solve", h16, h16, False, ident),
                                  ("fp16 operands incl. TxT matrices, tensor-core solve", h16, h16, True, h16)):
        st = ins[0].clone()
        y = chunked(st, *ins[1:], et, Tc=a.T, operand=op, state_operand=sop, tensor_solve=ts, small=sm)
        print(f"T={a.T} {name}: y rel-l2 {rel(y, y_ref):.2e} max-rel {relmax(y, y_ref):.2e}; "
              f"state rel-l2 {rel(st.float(), st_ref.float()):.2e} max-rel {relmax(st.float(), st_ref.float()):.2e}")
    # how far is the sequential fp16 kernel from an exact (fp64) evaluation of the same recurrence?  (the reference's own noise)
    st64 = ins[0].double().clone()
    B, T, C = ins[1].shape
    H = C // 64
    tt = (et.view(B, 1) + torch.arange(T).view(1, T)).to(torch.int32)
    d = decay(ins[2].view(B * T, C), tt.reshape(-1)).view(B, T, H, 64).double()
    f = lambda x: x.double().view(B, T, H, 64)
    r_, k_, v_, a_, b_ = (f(x) for x in (ins[1], ins[3], ins[4], ins[5], ins[6]))
    y64 = torch.empty(B, T, H, 64, dtype=torch.float64)
    for t in range(T):
        u = torch.einsum("bhij,bhj->bhi", st64, a_[:, t])
        st64 = st64 * d[:, t].unsqueeze(2) + u.unsqueeze(3) * b_[:, t].unsqueeze(2) + v_[:, t].unsqueeze(3) * k_[:, t].unsqueeze(2)
        y64[:, t] = torch.einsum("bhij,bhj->bhi", st64, r_[:, t])
    print(f"reference sequential fp16 kernel vs exact fp64 recurrence: y rel-l2 {rel(y_ref.double(), y64.view(B, T, C)):.2e}; "
          f"state rel-l2 {rel(st_ref.double(), st64):.2e}")
    st = ins[0].clone()
    y = chunked(st, *ins[1:], et, Tc=a.T, tensor_solve=True, small=h16)
    print(f"chunked (fp16 operands, tensor-core solve) vs exact fp64 recurrence: y rel-l2 {rel(y.double(), y64.view(B, T, C)):.2e}; "
          f"state rel-l2 {rel(st.double(), st64):.2e}")


if __name__ == "__main__":
    main()
