"""CPU: the oracle (oracle/*.py) against the committed golden vectors that oracle/gen_golden.py produced from the
UNMODIFIED reference (fairseq stood in by torchaudio), and against torchaudio itself where that is cheap."""
import numpy as np
import pytest
import torch

from occm_b200 import synth
from oracle import backend as obe, frontend as ofe, scoring as osc


def test_eer_kats_oracle(eer_kats):
    for k, t, n in eer_kats:
        assert osc.compute_eer(t, n) == (k["eer"], k["thr"]), k["name"]
        if "eer_swapped" in k:
            assert osc.compute_eer(n, t) == (k["eer_swapped"], k["thr_swapped"])


def test_num_frames():
    for n, t in ((64600, 201), (16000, 49), (160000, 499), (480000, 1499)):
        assert synth.num_frames(n) == t
        assert int(ofe.frame_lengths(torch.tensor([n]))[0]) == t


def test_oracle_full_vs_reference_golden(golden, synth_sd):
    """Config 1 (first two utterances; the reference is batch-invariant to 2e-7, PINNING.json)."""
    wav = synth.make_waveforms(seed=1, batch=8, n_samples=64600)[:2]
    emb, out = osc.amodel_forward(synth_sd, wav)
    assert (emb - golden["c1_emb"][:2]).abs().max() < 2e-6
    assert (out - golden["c1_out"][:2]).abs().max() < 2e-6
    d = osc.one_class_scores(golden["c1_refemb"], emb)
    assert torch.allclose(d, golden["c1_dist"][:2], rtol=1e-4, atol=1e-6)


def test_oracle_varlen_vs_reference_golden(golden, synth_sd):
    w = synth.make_waveforms(seed=30, batch=1, n_samples=16000)
    emb, out = osc.amodel_forward(synth_sd, w)
    assert (emb - golden["vl_emb"][:1]).abs().max() < 2e-6
    assert (out - golden["vl_out"][:1]).abs().max() < 2e-6


def test_reference_embedding_and_lines(golden):
    embs = [e[None] for e in golden["c1_ref_embs"]]
    ref, thr = osc.reference_embedding(embs)
    assert torch.equal(ref, golden["c1_refemb"]) and torch.equal(thr, golden["c1_thr"])
    assert osc.score_lines_1c(golden["c1_dist"].tolist(), float(golden["c1_thr"])) == golden["c1_lines"]
    x = torch.zeros(1, 160)
    assert abs(float(osc.one_class_scores(x, x)) - 1.2649e-5) < 1e-8       # eps is added to the difference


def test_score_line_kat():
    line = osc.score_lines_1c([float(np.float32(0.07951234))], 0.05)[0]
    assert line == "0.0795123428106308, 1 \n"
    assert float(line.split(",")[0]) == float(np.float32(0.07951234))


def test_frontend_oracle_vs_torchaudio_small(synth_sd):
    """The frontend restatement against torchaudio's own wav2vec2_xlsr_300m on the same weights (bit-exact on CPU)."""
    torchaudio = pytest.importorskip("torchaudio")
    m = torchaudio.models.wav2vec2_xlsr_300m().eval()
    m.load_state_dict(ofe.to_torchaudio_state_dict(synth_sd), strict=True)
    wav = synth.make_waveforms(seed=9, batch=1, n_samples=4000)
    with torch.no_grad():
        want, _ = m(wav)
        got = ofe.xlsr_forward(synth_sd, wav)
    assert (got - want).abs().max() <= 1e-6


def test_backend_dead_bn1_is_ignored(synth_sd):
    """Residual_block.bn1 is computed and discarded in the reference (sslassist.py:408-415)."""
    sd2 = dict(synth_sd)
    for k in list(sd2):
        if k.startswith("encoder.") and ".bn1." in k and sd2[k].dtype == torch.float32:
            sd2[k] = sd2[k] * 3.0 + 1.0
    feat = torch.randn(1, 49, 1024, generator=torch.Generator().manual_seed(3))
    a = obe.aasist_backend(synth_sd, feat)
    b = obe.aasist_backend(sd2, feat)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])


def test_ocsoftmax_oracle_properties():
    g = torch.Generator().manual_seed(0)
    x = torch.randn(5, 160, generator=g)
    w = torch.randn(160, generator=g)
    s = osc.ocsoftmax_scores(x, w)
    assert s.abs().max() <= 1.0 + 1e-6
    assert torch.allclose(osc.ocsoftmax_scores(3 * x, 0.5 * w), s, atol=1e-6)   # scale invariant
    assert abs(float(osc.ocsoftmax_scores(w[None], w)) - 1.0) < 1e-6
