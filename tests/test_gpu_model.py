"""GPU (-m gpu): the scoring path through the reference-facing surface (AModel / SSLModel / scoring) and the
C-ABI stage entry points, against the oracle and the golden vectors produced from the unmodified reference.

Tolerances (rel = max|got-want| / max|want| per tensor):
  * fp32 mode (split-bf16 3-MMA GEMMs): north_star's 1e-3 on emb / out / distance; measured ~3e-6 .. 3e-5.
  * bf16 mode: ONE stated tensor-level bound, 1e-1 on emb / out (the bound smoke() uses too; top-k graph pooling turns a
    rounding-level change of a node score into a different node, so the per-tensor error is jumpy: measured 9.3e-3 on config 1,
    2.1e-2 on the 2 x 16 000-sample smoke input, 6.9e-2 on the B = 64 batch) and 2e-2 on frontend features (measured 9.3e-3);
    the statement that matters is at SCORE level and lives in tests/test_gpu_round2.py (distance distribution, Spearman, dEER).
"""
import ctypes as C

import numpy as np
import pytest
import torch

from occm_b200 import _lib, synth
from oracle import backend as obe, frontend as ofe, scoring as osc

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)
FP32_TOL = 1e-3
BF16_TOL_EMB = 1e-1
BF16_TOL_FEAT = 2e-2


def rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.fixture(scope="module")
def model_fp32():
    from occm_b200.models.sslassist import AModel
    m = AModel(None, DEV, dtype="fp32").to(DEV).eval()
    yield m
    if m._engine is not None:
        m._engine.close()


@pytest.fixture(scope="module")
def model_bf16():
    from occm_b200.models.sslassist import AModel
    m = AModel(None, DEV, dtype="bf16").to(DEV).eval()
    yield m
    if m._engine is not None:
        m._engine.close()


@pytest.fixture(scope="module")
def small(synth_sd):
    """B=2 x 16 000 samples with oracle stage tensors."""
    wav = synth.make_waveforms(5, 2, 16000)
    with torch.no_grad():
        feats = ofe.conv_feature_encoder(synth_sd, wav)
        proj = ofe.feature_projection(synth_sd, feats)
        x0 = proj + ofe.pos_conv(synth_sd, proj)
        x2 = x0
        for i in range(2):
            x2 = ofe.encoder_layer(synth_sd, x2, i, None)
        p = ofe.P
        y2 = torch.nn.functional.layer_norm(x2, (1024,), synth_sd[p + "encoder.layer_norm.weight"], synth_sd[p + "encoder.layer_norm.bias"])
        emb, out = obe.aasist_backend(synth_sd, y2)
    return dict(wav=wav, proj=proj, x0=x0, x2=x2, y2=y2, emb=emb, out=out)


# ------------------------------------------------------------------------------------ stage entry points (C ABI)
@pytest.mark.parametrize("mode", ["fp32", "bf16"])
def test_stage_entry_points_vs_oracle(mode, small, model_fp32, model_bf16):
    model = model_fp32 if mode == "fp32" else model_bf16
    tol = FP32_TOL if mode == "fp32" else BF16_TOL_FEAT
    eng = model._get_engine(2, 16000, DEV)
    lib, h = eng.lib, eng.handle
    B, T = 2, small["proj"].shape[1]
    buf = lambda: torch.full((B, T, 1024), float("nan"), device=DEV)          # noqa: E731
    out = buf()
    wav = small["wav"].to(DEV)
    assert lib.occ_conv_encoder(h, wav.data_ptr(), B, 16000, out.data_ptr(), None) == 0, _lib.last_error(h)
    assert rel(out, small["proj"]) < tol
    x = small["proj"].to(DEV).contiguous()
    out = buf()
    assert lib.occ_pos_conv(h, x.data_ptr(), B, T, out.data_ptr(), None) == 0, _lib.last_error(h)
    assert rel(out, small["x0"]) < tol
    x = small["x0"].to(DEV).contiguous()
    out, hid = buf(), torch.full((2, B, T, 1024), float("nan"), device=DEV)
    assert lib.occ_transformer(h, x.data_ptr(), B, T, 2, out.data_ptr(), hid.data_ptr(), None) == 0, _lib.last_error(h)
    assert rel(out, small["y2"]) < tol and rel(hid[1], small["x2"]) < tol
    f = small["y2"].to(DEV).contiguous()
    emb, o2 = torch.full((B, 160), float("nan"), device=DEV), torch.full((B, 2), float("nan"), device=DEV)
    assert lib.occ_aasist_backend(h, f.data_ptr(), B, T, emb.data_ptr(), o2.data_ptr(), None) == 0, _lib.last_error(h)
    assert rel(emb, small["emb"]) < 1e-4 and rel(o2, small["out"]) < 1e-4      # backend is fp32 in both modes


def test_backend_long_utterance_vs_oracle(synth_sd, model_fp32):
    """10 s utterance (T = 499 frames, 166 temporal nodes): the graph no longer fits the one-CTA GAT kernel (row-block
    fallback) and the residual-encoder taps span more than the 128-row halo (per-tap loads); both against the oracle."""
    B, T = 1, 499
    g = torch.Generator().manual_seed(77)
    feat = torch.randn(B, T, 1024, generator=g)
    with torch.no_grad():
        want_emb, want_out = obe.aasist_backend(synth_sd, feat)
    eng = model_fp32._get_engine(B, 160000, DEV)
    lib, h = eng.lib, eng.handle
    f = feat.to(DEV).contiguous()
    emb, out = torch.full((B, 160), float("nan"), device=DEV), torch.full((B, 2), float("nan"), device=DEV)
    assert lib.occ_aasist_backend(h, f.data_ptr(), B, T, emb.data_ptr(), out.data_ptr(), None) == 0, _lib.last_error(h)
    assert rel(emb, want_emb) < 1e-4 and rel(out, want_out) < 1e-4


# ------------------------------------------------------------------------------------ config 1: B=8 x 64 600
def test_config1_fp32_vs_reference_golden(model_fp32, golden):
    """BASELINE config 1 against tensors produced by the UNMODIFIED reference (tests/golden/model_golden.pt)."""
    from occm_b200.scoring import one_class_scores, format_score_lines_1c
    wav = synth.make_waveforms(seed=1, batch=8, n_samples=64600).to(DEV)
    with torch.no_grad():
        emb, out = model_fp32(wav)
    assert emb.shape == (8, 160) and out.shape == (8, 2) and emb.dtype == torch.float32
    assert rel(emb, golden["c1_emb"]) < FP32_TOL and rel(out, golden["c1_out"]) < FP32_TOL
    d = one_class_scores(golden["c1_refemb"].to(DEV), emb)
    assert rel(d, golden["c1_dist"]) < FP32_TOL
    # same decisions as the reference's score file (distance > threshold), oc_classifier.py:262-265
    got = [l.split(",")[1] for l in format_score_lines_1c(d.cpu().tolist(), float(golden["c1_thr"]))]
    assert got == [l.split(",")[1] for l in golden["c1_lines"]]
    # [B,N,1] input (sslassist.py:42-43, 508)
    e3, _ = model_fp32(wav[:2].unsqueeze(-1))
    assert torch.equal(e3, emb[:2])


def test_config1_frontend_stage_goldens(model_fp32, golden):
    wav = synth.make_waveforms(seed=1, batch=8, n_samples=64600).to(DEV)
    feat, hid = model_fp32.ssl_model.extract_feat(wav, all_layers=True)
    assert feat.shape == (8, 201, 1024) and hid.shape == (24, 8, 201, 1024)
    sub = lambda t: t[:, ::10, ::16]                                             # noqa: E731
    assert rel(sub(feat), golden["c1_feat_sub"]) < FP32_TOL
    assert rel(sub(hid[0]), golden["c1_l0_sub"]) < FP32_TOL
    assert rel(sub(hid[11]), golden["c1_l11_sub"]) < FP32_TOL


def test_config2_bf16_tolerance(model_bf16, golden):
    """bf16 compute at B=8 (the B=64 shape is in test_gpu_round2.py): stated tolerance on emb/out, distances within 5 %."""
    from occm_b200.scoring import one_class_scores
    wav = synth.make_waveforms(seed=1, batch=8, n_samples=64600).to(DEV)
    with torch.no_grad():
        emb, out = model_bf16(wav)
    assert rel(emb, golden["c1_emb"]) < BF16_TOL_EMB and rel(out, golden["c1_out"]) < BF16_TOL_EMB
    d = one_class_scores(golden["c1_refemb"].to(DEV), emb)
    assert float(((d.cpu() - golden["c1_dist"]).abs() / golden["c1_dist"]).max()) < 0.05


def test_batch_invariance_bit_exact(model_bf16, model_fp32):
    """An utterance's result must not depend on batch composition (-> world-size-invariant scores/EER)."""
    wav = synth.make_waveforms(seed=21, batch=5, n_samples=32000).to(DEV)
    for m in (model_bf16, model_fp32):
        e_all, o_all = m(wav)
        e1, o1 = m(wav[3:4])
        e2, o2 = m(wav[2:5])
        assert torch.equal(e1[0], e_all[3]) and torch.equal(o1[0], o_all[3])
        assert torch.equal(e2, e_all[2:5]) and torch.equal(o2, o_all[2:5])


# ------------------------------------------------------------------------------------ config 4: variable length
def test_config4_variable_length_vs_reference_batch1(model_fp32, golden):
    """Each utterance's truth = the reference alone at batch 1, un-padded (golden vl_emb); here they run as ONE
    padded batch with per-utterance lengths (masks in the conv encoder / attention / backend)."""
    lens = golden["vl_lens"].tolist()
    N = max(lens)
    wav = torch.zeros(len(lens), N)
    for j, L in enumerate(lens):
        wav[j, :L] = synth.make_waveforms(seed=30 + j, batch=1, n_samples=L)[0]
        wav[j, L:] = 0.3          # padding content must not matter
    with torch.no_grad():
        emb, out = model_fp32(wav.to(DEV), lengths=lens)
    for j in range(len(lens)):
        assert rel(emb[j], golden["vl_emb"][j]) < FP32_TOL, (j, lens[j])
        assert rel(out[j], golden["vl_out"][j]) < FP32_TOL


def test_ragged_and_limits(model_fp32):
    with pytest.raises(ValueError):
        model_fp32(torch.zeros(1, 300, device=DEV))                  # too short for the backend
    with pytest.raises(ValueError):
        model_fp32(torch.zeros(2, 16000, device=DEV), lengths=[16000])
    with pytest.raises(ValueError):
        model_fp32(torch.zeros(2, 16000, device=DEV), lengths=[16000, 17000])
    e, o = model_fp32(torch.zeros(1, 16000, device=DEV))             # all-zero waveform: LN of a constant row is finite
    assert torch.isfinite(e).all() and torch.isfinite(o).all()


# ------------------------------------------------------------------------------------ config 5: hidden states
def test_config5_hidden_states_vs_torchaudio_golden(model_fp32, golden):
    w5 = synth.make_waveforms(seed=50, batch=2, n_samples=48000).to(DEV)
    feat, hid = model_fp32.ssl_model.extract_feat(w5, all_layers=True)
    got = hid[:, :, ::10, ::32]
    assert got.shape == golden["c5_layers_sub"].shape
    for l in range(24):
        assert rel(got[l], golden["c5_layers_sub"][l]) < FP32_TOL, l


# ------------------------------------------------------------------------------------ scoring drivers / host path
def test_scoring_drivers_and_score_file(tmp_path, model_bf16):
    from occm_b200 import scoring
    from occm_b200.evaluate_metrics import compute_eer
    train = synth.make_waveforms(seed=2, batch=6, n_samples=16000)
    evalw = synth.make_waveforms(seed=3, batch=10, n_samples=16000)
    evalw[5:] *= 3.0
    ref, thr = scoring.create_reference_embedding(model_bf16, train, DEV, batch_size=4)
    assert ref.shape == (1, 160)
    p = tmp_path / "scores.txt"
    d = scoring.score_eval_set_1c(model_bf16, evalw, DEV, ref, thr, str(p), batch_size=4)
    lines = p.read_text().splitlines(keepends=True)
    assert lines == scoring.format_score_lines_1c(d.cpu().tolist(), float(thr))
    back = scoring.load_score_file(str(p))
    assert back == [float(v) for v in d.cpu().tolist()]                      # repr(float) round-trips the fp32 value
    labels = ["bonafide"] * 5 + ["spoof"] * 5
    eer, th = scoring.calculate_EER(back, labels)
    s = np.array(back)
    assert (eer, th) == osc.compute_eer(s[5:], s[:5]) == compute_eer(s[5:], s[:5])
    s2 = scoring.score_eval_set_2c(model_bf16, evalw[:3], DEV, str(p))
    assert p.read_text() == "".join(f"{float(v)}\n" for v in s2.cpu().tolist())


def test_forward_host_matches_device_path(model_bf16):
    wav = synth.make_waveforms(seed=8, batch=3, n_samples=16000)
    ref = torch.randn(160, generator=torch.Generator().manual_seed(1)) * 0.1
    emb_h, out_h, score_h = model_bf16.forward_host(wav, ref)
    emb_d, out_d = model_bf16(wav.to(DEV))
    assert torch.equal(emb_h, emb_d.cpu()) and torch.equal(out_h, out_d.cpu())
    from occm_b200.scoring import one_class_scores
    assert torch.equal(score_h, one_class_scores(ref.to(DEV), emb_d).cpu())


def test_load_state_dict_rebuilds_engine(model_bf16):
    wav = synth.make_waveforms(seed=4, batch=1, n_samples=16000).to(DEV)
    e0, _ = model_bf16(wav)
    orig = model_bf16.state_dict()["LL.bias"].clone()
    model_bf16.load_state_dict({"LL.bias": orig + 0.05}, strict=False)
    e1, _ = model_bf16(wav)
    assert not torch.equal(e0, e1)
    model_bf16.load_state_dict({"LL.bias": orig}, strict=False)
    e2, _ = model_bf16(wav)
    assert torch.equal(e0, e2)


def test_score_host_batches_pipelined_equals_forward_host(model_bf16):
    """The pipelined end-to-end driver (bench.py's e2e) returns exactly what the synchronous host call returns, in order,
    including a ragged last batch."""
    from occm_b200.scoring import score_host_batches
    w = synth.make_waveforms(seed=123, batch=11, n_samples=16000)
    ref = 0.1 * torch.randn(1, 160, generator=torch.Generator().manual_seed(4))
    batches = [w[0:4].contiguous().pin_memory(), w[4:8].contiguous().pin_memory(), w[8:11].contiguous().pin_memory()]
    got = list(score_host_batches(model_bf16, batches, ref, DEV))
    assert [g[0].shape[0] for g in got] == [4, 4, 3]
    for hb, (e, o, d) in zip(batches, got):
        e2, o2, d2 = model_bf16.forward_host(hb, ref)
        assert torch.equal(e, e2) and torch.equal(o, o2) and torch.equal(d, d2)


@pytest.mark.gpu
def test_create_reference_embedding2_files_and_cache(tmp_path, model_bf16):
    from occm_b200 import scoring
    """oc_classifier.py:159-202 side effects: reference_embedding.pt [1,160], threshold.pt (0-d), one distances.txt line per
    bona fide utterance in the reference's ``f"{float(d)}\\n"`` form; the second call is served from the cache."""
    g = torch.Generator(device="cuda").manual_seed(11)
    wav = torch.empty(6, 16000, device="cuda").normal_(0, 0.1, generator=g)
    ref, thr = scoring.create_reference_embedding2(model_bf16, wav, "cuda", cache_dir=str(tmp_path))
    emb, _ = model_bf16(wav)
    assert torch.allclose(ref.cpu(), emb.mean(0, keepdim=True).cpu(), rtol=1e-6, atol=1e-7)
    lines = (tmp_path / "distances.txt").read_text().splitlines()
    d = [float(x) for x in lines]
    assert len(d) == 6 and max(d) == float(thr) and all(repr(v) == s for v, s in zip(d, lines))
    want = torch.nn.functional.pairwise_distance(ref.cpu().double(), emb.cpu().double()).float()
    assert torch.allclose(torch.tensor(d), want, rtol=1e-5)
    sref, sthr = torch.load(tmp_path / "reference_embedding.pt"), torch.load(tmp_path / "threshold.pt")
    assert sref.shape == (1, 160) and sthr.ndim == 0
    r2, t2 = scoring.create_reference_embedding2(None, None, "cuda", cache_dir=str(tmp_path))
    assert torch.equal(r2.cpu(), ref.cpu()) and float(t2) == float(thr)
    assert len((tmp_path / "distances.txt").read_text().splitlines()) == 6
