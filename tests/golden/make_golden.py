"""Generates tests/golden/*.npz by running the UNMODIFIED reference model (imported from /root/reference) on CPU.

Run in the build container only (``python tests/golden/make_golden.py``); /root/reference does not exist on the GPU
box, which is why the vectors are committed.  The seam is the one described in SURVEY.md §8(c):

* ``RWKV_x070(args, auto_load=False)``  — skips the hard-coded ``"cuda"`` of ``Albatross/rwkv7.py:203``;
* the reference's own load-time transform (``load_weights_to_device`` ``:211-221``) is applied on CPU and the
  post-load fixes of ``:206-209`` are replayed, then the whole dict is assigned back (``m.z = zz``);
* state tensors are built by hand (``generate_zero_state`` hard-codes ``"cuda"``, ``:228-234``);
* the reference has no CPU body for its WKV custom op, so ``oracle/wkv7_oracle.c`` is registered on the
  reference's own ``mylib::RWKV7_BATCH_OP`` (``:154-157``).  Everything else that runs is reference code.

Each case also runs ``oracle/rwkv7_oracle.py`` and asserts it is BIT-IDENTICAL to the reference before saving;
that is the pin of the oracle's Python part.  The C part (WKV body) is pinned on the GPU box against the
reference's real CUDA kernel (tests/test_wkv_gpu.py).
"""
import hashlib
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0a")
os.environ.setdefault("TORCH_EXTENSIONS_DIR", "/tmp/ref_torch_ext")

import numpy as np
import torch

from chirrup_b200.synth import GEOMETRIES, make_checkpoint
from oracle import rwkv7_oracle as O

REF_ROOT = os.environ.get("REF_ROOT", "/root/reference")
OUT = os.path.dirname(os.path.abspath(__file__))


def load_reference(raw, geom, tmpdir="/tmp/golden_ckpt"):
    sys.path.insert(0, REF_ROOT)
    import Albatross.rwkv7 as R  # JIT-builds the reference extension (CPU box: the < sm_80 branch)

    os.makedirs(tmpdir, exist_ok=True)
    path = os.path.join(tmpdir, f"{geom.name}")
    torch.save(raw, path + ".pth")
    args = types.SimpleNamespace(vocab_size=geom.vocab_size, head_size=64, MODEL_NAME=path)
    m = R.RWKV_x070(args, auto_load=False)
    # reference transform, on CPU, on the module's own dict copy
    zz = dict(m.z)
    m_z_backup = m.z
    holder = types.SimpleNamespace(z=zz)
    R.RWKV_x070.load_weights_to_device(holder, list(zz.keys()), "cpu")
    zz = holder.z
    zz["emb.weight"] = torch.nn.functional.layer_norm(
        zz["emb.weight"], (geom.n_embd,), weight=zz["blocks.0.ln0.weight"], bias=zz["blocks.0.ln0.bias"])
    zz["blocks.0.att.v0"] = zz["blocks.0.att.a0"]
    zz["blocks.0.att.v1"] = zz["blocks.0.att.a1"]
    zz["blocks.0.att.v2"] = zz["blocks.0.att.a2"]
    m.z = zz
    if not getattr(load_reference, "_registered", False):
        R.RWKV7_BATCH_OP.register_kernel("cpu")(
            lambda state, r, w, k, v, a, b, elapsed_t: O.wkv_cpu(state, r, w, k, v, a, b, elapsed_t))
        load_reference._registered = True
    return R, m


def sha(t: torch.Tensor) -> str:
    return hashlib.sha256(t.contiguous().numpy().tobytes()).hexdigest()


def run_case(name, geom_name, seed, B, prompt_len, n_decode, recipe="plain", store_state=True, big_state_offset=None):
    geom = GEOMETRIES[geom_name]
    raw = make_checkpoint(geom, seed=seed, recipe=recipe)
    R, m = load_reference({k: v.clone() for k, v in raw.items()}, geom)
    z = O.transform_checkpoint(raw)
    og = O.Geometry(z)
    gen = torch.Generator().manual_seed(1000 + seed)
    prompt = torch.randint(1, geom.vocab_size, (B, prompt_len), generator=gen).tolist()

    def fresh():
        s = O.zero_state(og, B)
        if big_state_offset is not None:
            s[2] += big_state_offset     # exercise the int32 wrap of the dither (.cu:20-23)
        return s

    s_ref, s_orc = fresh(), fresh()
    logits_ref = [m.forward_seq_batch_seperate(prompt, s_ref)]
    logits_orc = [O.forward_seq_batch(z, og, prompt, s_orc)]
    toks = []
    for _ in range(n_decode):
        nxt = logits_ref[-1].float().argmax(-1)
        toks.append(nxt.clone())
        step = [[int(t)] for t in nxt]
        logits_ref.append(m.forward_seq_batch_seperate(step, s_ref))
        logits_orc.append(O.forward_seq_batch(z, og, step, s_orc))
    for a, b in zip(logits_ref, logits_orc):
        assert torch.equal(a, b), f"{name}: oracle != reference (logits)"
    for a, b in zip(s_ref, s_orc):
        assert torch.equal(a, b), f"{name}: oracle != reference (state)"
    out = dict(
        geom=geom_name, seed=seed, recipe=recipe, B=B, prompt=np.array(prompt, dtype=np.int64),
        tokens=torch.stack(toks, 1).numpy() if toks else np.zeros((B, 0), np.int64),
        logits_prefill=logits_ref[0].numpy(), logits_last=logits_ref[-1].numpy(),
        elapsed=s_ref[2].numpy(), state_offset=np.int64(big_state_offset or 0),
        sha_state0=sha(s_ref[0]), sha_state1=sha(s_ref[1]),
    )
    if store_state:
        out["state0"] = s_ref[0].numpy()
        out["state1"] = s_ref[1].numpy()
    else:
        out["state1_sample"] = s_ref[1].flatten()[::997].numpy()
        out["state0_sample"] = s_ref[0].flatten()[::97].numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(f"[golden] {name}: ok, logits std {logits_ref[-1].float().std():.3f}, "
          f"|state1| max {s_ref[1].float().abs().max():.3f}")


def run_one_case(name, geom_name, seed, n_steps):
    """bsz-1 ``forward`` path (forward_one + TMix_one/CMix_one; the sparse FFN product is the reference's Triton/CUDA op,
    which cannot run here, so the fixture only pins the oracle's own output for regression; flagged unpinned)."""
    geom = GEOMETRIES[geom_name]
    raw = make_checkpoint(geom, seed=seed)
    z = O.transform_checkpoint(raw)
    og = O.Geometry(z)
    st = O.zero_state(og, 0)
    gen = torch.Generator().manual_seed(2000 + seed)
    toks = torch.randint(1, geom.vocab_size, (n_steps,), generator=gen).tolist()
    for t in toks:
        logits = O.forward_one(z, og, t, st)
    # cross-check with the batched path (same math, B = 1): must agree to fp16 noise
    st_b = O.zero_state(og, 1)
    logits_b = O.forward_seq_batch(z, og, [toks], st_b)
    err = (logits.float() - logits_b[0].float()).abs().max().item()
    assert err < 0.05, err
    np.savez_compressed(os.path.join(OUT, name + ".npz"), geom=geom_name, seed=seed, tokens=np.array(toks),
                        logits=logits.numpy(), state0=st[0].numpy(), state1=st[1].numpy(), pinned=False)
    print(f"[golden] {name}: ok (oracle-only, one-vs-batch max abs {err:.4f})")


if __name__ == "__main__":
    torch.manual_seed(0)
    only = set(sys.argv[1:])
    _run_case, _run_one_case = run_case, run_one_case
    run_case = lambda name, *a, **k: _run_case(name, *a, **k) if (not only or name in only) else None
    run_one_case = lambda name, *a, **k: _run_one_case(name, *a, **k) if (not only or name in only) else None
    run_case("tiny_b3_p5_d4", "tiny", seed=0, B=3, prompt_len=5, n_decode=4)
    run_case("tiny_b2_p1_d6_wrap", "tiny", seed=1, B=2, prompt_len=1, n_decode=6, big_state_offset=(1 << 20) + 3)
    run_case("small_b4_p16_d8", "small", seed=2, B=4, prompt_len=16, n_decode=8)
    run_case("small_margin_b2_p4_d32", "small", seed=3, B=2, prompt_len=4, n_decode=32, recipe="margin")
    run_case("g01B_b1_p16_d8", "0.1B", seed=0, B=1, prompt_len=16, n_decode=8, store_state=False)
    run_one_case("tiny_one_s6", "tiny", seed=4, n_steps=6)
    # the parity bar's recipe (chirrup_b200/synth.py "trained_like"): greedy tokens must be reproduced exactly
    run_case("small_tl_b4_p8_d32", "small", seed=5, B=4, prompt_len=8, n_decode=32, recipe="trained_like")
    run_case("g01B_tl_b2_p8_d8", "0.1B", seed=6, B=2, prompt_len=8, n_decode=8, recipe="trained_like", store_state=False)
