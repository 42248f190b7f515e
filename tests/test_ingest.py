"""Ingest (int16 PCM -> fp32 batch, data_utils_SSL.pad) and calculate_confusion_matrix: oracle vs the golden vectors produced
by the unmodified reference (oracle/gen_ingest_golden.py), and the CUDA path (C ABI) vs the oracle, bit-exact."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import scoring as osc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def kat():
    d = json.load(open(os.path.join(ROOT, "tests", "golden", "ingest_kat.json")))
    rng = np.random.default_rng(7)                    # replay the generator's draw order
    xs = [rng.integers(-32768, 32768, c["n"], dtype=np.int16) for c in d["pad_cases"]]
    t = rng.normal(1, 1, 1000).astype(np.float32).astype(np.float64)
    nt = rng.normal(-1, 1, 3000).astype(np.float32).astype(np.float64)
    return d, xs, t, nt


def _check_pad_case(c, x, y):
    assert y.shape == (c["max_len"],) and y.dtype == np.float32
    if "y" in c:
        assert x.tolist() == c["x"]
        assert [float(v) for v in y] == c["y"]
    else:
        assert x[:8].tolist() == c["x_first"]
        assert float(np.sum(y.astype(np.float64))) == c["y_sum"]
        assert [float(v) for v in y[-4:]] == c["y_tail"]


def test_oracle_pad_and_confusion_vs_reference_golden(kat):
    d, xs, t, nt = kat
    for c, x in zip(d["pad_cases"], xs):
        _check_pad_case(c, x, osc.pad(osc.pcm16_to_float(x), c["max_len"]))
    for item in d["cm"]:
        assert list(osc.calculate_confusion_matrix(t, nt, item["thr"])) == item["counts"]


@pytest.mark.gpu
def test_gpu_ingest_matches_reference_pad(kat):
    from occm_b200 import ingest
    d, xs, _, _ = kat
    for c, x in zip(d["pad_cases"], xs):
        y = ingest.pad(x, c["max_len"]).cpu().numpy()
        _check_pad_case(c, x, y)
    # one ragged batch, both modes
    wav, lens = ingest.batch_from_pcm16(xs[:5], 64, "repeat")
    for b, x in enumerate(xs[:5]):
        assert np.array_equal(wav[b].cpu().numpy(), osc.pad(osc.pcm16_to_float(x), 64))
    assert lens == [64] * 5
    wav0, lens0 = ingest.batch_from_pcm16(xs[:5], 16, "zero")
    for b, x in enumerate(xs[:5]):
        want = np.zeros(16, np.float32)
        n = min(len(x), 16)
        want[:n] = osc.pcm16_to_float(x)[:n]
        assert np.array_equal(wav0[b].cpu().numpy(), want) and lens0[b] == n
    with pytest.raises(ValueError):
        ingest.batch_from_pcm16([np.zeros(0, np.int16)], 16)
    with pytest.raises(ValueError):
        ingest.batch_from_pcm16([np.zeros(4, np.float32)], 16)


@pytest.mark.gpu
def test_gpu_confusion_matrix_matches_reference(kat):
    from occm_b200.evaluate_metrics import calculate_confusion_matrix
    d, _, t, nt = kat
    for item in d["cm"]:
        assert list(calculate_confusion_matrix(t, nt, item["thr"])) == item["counts"]
    dev = torch.device("cuda", 0)
    tt, nn = torch.from_numpy(t).to(dev), torch.from_numpy(nt).to(dev)
    assert list(calculate_confusion_matrix(tt, nn, 0.0)) == list(osc.calculate_confusion_matrix(t, nt, 0.0))


@pytest.mark.gpu
def test_gpu_ingest_feeds_the_model():
    """PCM16 utterance -> pad -> AModel.forward equals the same fp32 waveform handed over directly."""
    from occm_b200 import ingest
    from occm_b200.models.sslassist import AModel
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(3)
    x = (rng.normal(0, 3000, 9000)).astype(np.int16)
    wav, _ = ingest.batch_from_pcm16([x], 16000, "repeat")
    ref = torch.from_numpy(osc.pad(osc.pcm16_to_float(x), 16000)).to(dev)[None]
    m = AModel(None, dev, dtype="fp32").to(dev).eval()
    with torch.no_grad():
        e0, o0 = m(wav)
        e1, o1 = m(ref)
    assert torch.equal(e0, e1) and torch.equal(o0, o1)
    m._engine.close()
