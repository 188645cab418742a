"""CPU tier: oracle/sampler_oracle.py against the vectors recorded from the UNMODIFIED reference sampler
(tests/golden/make_sampler_golden.py): same seed -> same tokens; penalty step bit-identical."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import sampler_oracle as SO

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "sampler_*.npz")))


def _t(d, k, dtype=None):
    t = torch.from_numpy(d[k])
    return t.to(dtype) if dtype is not None else t


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p)[8:-4] for p in GOLD])
def test_oracle_matches_reference_recording(path):
    d = np.load(path)
    logits, occ = _t(d, "logits").clone(), _t(d, "occ").clone()
    SO.penalise(logits, occ, _t(d, "alpha"), _t(d, "decay"), _t(d, "freq"), d["forbidden"].tolist())
    assert torch.equal(logits, _t(d, "pen_logits")) and torch.equal(occ, _t(d, "pen_occ"))
    for seed in range(d["draws"].shape[0]):
        torch.manual_seed(seed)
        tok = SO.sample_logits_real_batch(logits.clone(), _t(d, "temp"), _t(d, "top_p"), _t(d, "top_k"))
        assert np.array_equal(tok.numpy(), d["draws"][seed])
    dist = SO.kept_distribution(logits, _t(d, "temp"), _t(d, "top_p"), _t(d, "top_k"))
    assert torch.allclose(dist, _t(d, "dist"), rtol=0, atol=1e-7)
    assert torch.allclose(dist.sum(1), torch.ones(dist.shape[0]), atol=1e-5)


GOLD_PIP = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "samplerpip_*.npz")))


@pytest.mark.parametrize("path", GOLD_PIP, ids=[os.path.basename(p)[11:-4] for p in GOLD_PIP])
def test_pip_oracle_matches_reference_recording(path):
    """The sampler the worker imports (chirrup/worker.py:12 -> samplers.py:171-255), parameter tensors in the worker's dtypes."""
    d = np.load(path)
    assert d["temp"].dtype == np.float16 and d["top_p"].dtype == np.float16 and d["decay"].dtype == np.float16
    logits, occ = _t(d, "logits").clone(), _t(d, "occ").clone()
    SO.penalise(logits, occ, _t(d, "alpha"), _t(d, "decay"), _t(d, "freq"), d["forbidden"].tolist())
    assert torch.equal(logits, _t(d, "pen_logits")) and torch.equal(occ, _t(d, "pen_occ"))
    for seed in range(d["draws"].shape[0]):
        torch.manual_seed(seed)
        tok = SO.sample_logits_rwkv_pip_compatible(logits.clone(), _t(d, "temp"), _t(d, "top_p"), _t(d, "top_k"))
        assert np.array_equal(tok.numpy(), d["draws"][seed])
    dist = SO.kept_distribution_pip(logits, _t(d, "temp"), _t(d, "top_p"), _t(d, "top_k"))
    assert torch.allclose(dist, _t(d, "dist"), rtol=0, atol=1e-7)


def test_worker_sampling_params_dtypes():
    p = SO.worker_sampling_params(0.0, 0.3, 0, 0.5, 0.996, 0.5)
    assert p["temperature"].dtype == torch.float16 and float(p["temperature"]) == 1.0       # worker.py:611-613: <= 0 -> 1.0
    assert float(p["penalty_decay"]) == 0.99609375 and p["presence_penalty"].dtype == torch.float32


def test_update_penalty():
    occ, alpha = torch.zeros(3, 16), torch.zeros(3, 16)
    pres = torch.tensor([[0.1], [0.2], [0.3]])
    SO.update_penalty(occ, alpha, pres, [0, 2, 2], [5, 7, 9], no_penalty_token_ids=(9,))
    assert occ[0, 5] == 1 and occ[2, 7] == 1 and occ[2, 9] == 0
    assert alpha[0, 5] == pytest.approx(0.1) and alpha[2, 9] == pytest.approx(0.3)
