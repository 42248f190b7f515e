"""Generate `tests/golden/*` from the UNMODIFIED reference, in the build container.

Run:  python -m oracle.gen_golden            (needs /root/reference; ~2 min on 8 cores)

What it does:
  * imports `/root/reference/models/sslassist.py` as-is, with `fairseq` replaced by a
    `sys.modules` shim whose checkpoint loader returns torchaudio's `wav2vec2_xlsr_300m()`
    (fairseq@a540213 is not installable here);
  * imports `/root/reference/evaluate_metrics.py` as-is;
  * loads the seeded synthetic checkpoint of `occm_b200.synth.make_state_dict(seed)` into the
    reference model, runs it on seeded waveforms, and stores emb / out / distances plus
    sub-sampled stage tensors;
  * checks the restatement in `oracle/` against the reference on the same inputs and records
    the max-abs differences in `tests/golden/PINNING.json`;
  * stores EER known-answer vectors computed by the reference's own `compute_eer`.

The GPU box has no /root/reference: tests there read only the committed files.
"""
from __future__ import annotations

import importlib.util
import json
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")


def load_reference_modules():
    import torchaudio

    class _W(torch.nn.Module):  # stands in for fairseq Wav2Vec2Model
        def __init__(self):
            super().__init__()
            self.m = torchaudio.models.wav2vec2_xlsr_300m()

        def forward(self, x, mask=False, features_only=True):
            y, _ = self.m(x)
            return {"x": y}

    fs = types.ModuleType("fairseq")
    cu = types.ModuleType("fairseq.checkpoint_utils")
    cu.load_model_ensemble_and_task = lambda paths: ([_W()], None, None)
    fs.checkpoint_utils = cu
    sys.modules["fairseq"] = fs
    sys.modules["fairseq.checkpoint_utils"] = cu
    spec = importlib.util.spec_from_file_location("ref_sslassist", os.path.join(REF, "models", "sslassist.py"))
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    spec2 = importlib.util.spec_from_file_location("ref_evaluate_metrics", os.path.join(REF, "evaluate_metrics.py"))
    em = importlib.util.module_from_spec(spec2)
    spec2.loader.exec_module(em)
    return ref, em


def load_into_reference(model, sd):
    """Feed the fairseq-named synthetic checkpoint to the shimmed reference model."""
    from oracle.frontend import to_torchaudio_state_dict

    ta = to_torchaudio_state_dict(sd)
    missing, unexpected = model.ssl_model.model.m.load_state_dict(ta, strict=True)
    be = {k: v for k, v in sd.items() if not k.startswith("ssl_model.")}
    res = model.load_state_dict(be, strict=False)
    assert not res.unexpected_keys, res.unexpected_keys
    assert all(k.startswith("ssl_model.") for k in res.missing_keys), res.missing_keys


def sub(t, *steps):
    idx = tuple(slice(None, None, s) for s in steps)
    return t[idx].contiguous().clone()


def main():
    from occm_b200 import synth
    from oracle import backend as obe, frontend as ofe, scoring as osc

    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(os.cpu_count())
    ref, em = load_reference_modules()
    pin = {"torch": torch.__version__, "reference_commit": "9141a39"}

    # ------------------------------------------------------------------ EER KATs
    kats = []

    def kat(name, t, n):
        t = np.asarray(t, dtype=np.float64)
        n = np.asarray(n, dtype=np.float64)
        e, th = em.compute_eer(t, n)
        e2, th2 = osc.compute_eer(t, n)
        assert (e, th) == (e2, th2), (name, e, th, e2, th2)
        return e, th

    e, th = kat("KAT1", [0.9, 0.8, 0.35, 0.7], [0.1, 0.4, 0.35, 0.2, 0.05])
    kats.append(dict(name="KAT1", target=[0.9, 0.8, 0.35, 0.7], nontarget=[0.1, 0.4, 0.35, 0.2, 0.05], eer=e, thr=th))
    e, th = kat("KAT2", [0.5] * 3, [0.5] * 2)
    kats.append(dict(name="KAT2", target=[0.5] * 3, nontarget=[0.5] * 2, eer=e, thr=th))
    rng = np.random.default_rng(1234)
    t = rng.normal(1, 1, 7355).astype(np.float32).astype(np.float64)
    n = rng.normal(-1, 1, 63882).astype(np.float32).astype(np.float64)
    e, th = kat("KAT3", t, n)
    e_sw, th_sw = kat("KAT3sw", n, t)
    kats.append(dict(name="KAT3", gen="default_rng(1234): normal(1,1,7355) then normal(-1,1,63882), f32->f64",
                     eer=e, thr=th, eer_swapped=e_sw, thr_swapped=th_sw))
    # tie-heavy seeded vectors (quantised scores): exercises mergesort stability
    for seed, nt, nn, q in ((7, 300, 500, 16), (8, 1, 1000, 4), (9, 2048, 2048, 64), (10, 5000, 3, 8)):
        r = np.random.default_rng(seed)
        tt = np.round(r.normal(0.3, 1, nt) * q) / q
        nn_ = np.round(r.normal(-0.3, 1, nn) * q) / q
        e, th = kat(f"ties{seed}", tt, nn_)
        kats.append(dict(name=f"ties{seed}", gen=f"default_rng({seed}): round(normal(.3,1,{nt})*{q})/{q}; "
                                                 f"round(normal(-.3,1,{nn})*{q})/{q}",
                         seed=seed, nt=nt, nn=nn, q=q, eer=e, thr=th))
    with open(os.path.join(OUT, "eer_kat.json"), "w") as f:
        json.dump(kats, f, indent=1)
    print("EER KATs written:", [k["name"] for k in kats])

    # ------------------------------------------------------------------ model goldens
    sd = synth.make_state_dict(seed=0)
    model = ref.AModel(None, "cpu").eval()
    load_into_reference(model, sd)

    gold = {}
    hooks = {}

    def grab(name):
        def h(mod, inp, out):
            hooks[name] = out.detach() if torch.is_tensor(out) else out
        return h

    ta = model.ssl_model.model.m
    hs = [ta.feature_extractor.register_forward_hook(lambda m, i, o: hooks.__setitem__("conv", o[0].detach())),
          ta.encoder.feature_projection.register_forward_hook(grab("proj")),
          ta.encoder.transformer.pos_conv_embed.register_forward_hook(grab("pos")),
          ta.encoder.transformer.layers[0].register_forward_hook(lambda m, i, o: hooks.__setitem__("l0", o[0].detach())),
          ta.encoder.transformer.layers[11].register_forward_hook(lambda m, i, o: hooks.__setitem__("l11", o[0].detach())),
          model.LL.register_forward_hook(grab("LL")),
          model.encoder.register_forward_hook(grab("enc")),
          model.GAT_layer_S.register_forward_hook(grab("gat_S")),
          model.GAT_layer_T.register_forward_hook(grab("gat_T")),
          model.pool_S.register_forward_hook(grab("pool_S")),
          model.pool_T.register_forward_hook(grab("pool_T"))]

    # config 1: B=8 x 64600 (fp32, CPU)
    wav = synth.make_waveforms(seed=1, batch=8, n_samples=64600)
    with torch.no_grad():
        emb, out = model(wav)
        feat = model.ssl_model.extract_feat(wav)
    gold["c1_emb"] = emb.clone()
    gold["c1_out"] = out.clone()
    gold["c1_conv_sub"] = sub(hooks["conv"], 1, 10, 8)        # [8,21,64]
    gold["c1_proj_sub"] = sub(hooks["proj"], 1, 10, 16)
    gold["c1_pos_sub"] = sub(hooks["pos"], 1, 10, 16)
    gold["c1_l0_sub"] = sub(hooks["l0"], 1, 10, 16)
    gold["c1_l11_sub"] = sub(hooks["l11"], 1, 10, 16)
    gold["c1_feat_sub"] = sub(feat, 1, 10, 16)
    gold["c1_LL_sub"] = sub(hooks["LL"], 1, 5, 4)
    gold["c1_enc_b0"] = sub(hooks["enc"][:2], 1, 4, 3, 3)
    gold["c1_gat_S"] = hooks["gat_S"][:2].clone()
    gold["c1_gat_T"] = hooks["gat_T"][:2].clone()
    gold["c1_pool_S"] = hooks["pool_S"][:2].clone()
    gold["c1_pool_T"] = hooks["pool_T"][:2].clone()

    # restatement vs reference, same inputs
    with torch.no_grad():
        o_feat = ofe.xlsr_forward(sd, wav)
        st = {}
        o_emb, o_out = obe.aasist_backend(sd, o_feat, st)
        # backend alone on the reference's own features (isolates backend restatement)
        b_emb, b_out = obe.aasist_backend(sd, feat)
    pin["frontend_feat_maxabs_vs_torchaudio"] = float((o_feat - feat).abs().max())
    pin["backend_emb_maxabs_vs_reference_same_feat"] = float((b_emb - emb).abs().max())
    pin["backend_out_maxabs_vs_reference_same_feat"] = float((b_out - out).abs().max())
    pin["full_emb_maxabs_vs_reference"] = float((o_emb - emb).abs().max())
    pin["full_out_maxabs_vs_reference"] = float((o_out - out).abs().max())
    print(json.dumps(pin, indent=1))

    # reference embedding from a separate seeded batch (seed 2, 32 utterances) -- 8 here
    wav_ref = synth.make_waveforms(seed=2, batch=8, n_samples=64600)
    with torch.no_grad():
        embs = [model(wav_ref[i:i + 1])[0] for i in range(8)]
    refemb = torch.mean(torch.stack(embs), dim=0)                       # oc_classifier.py:189
    import torch.nn.functional as F
    dist = torch.cat([F.pairwise_distance(refemb, emb[i:i + 1], p=2) for i in range(8)])
    thr = torch.max(torch.stack([F.pairwise_distance(refemb, e, p=2) for e in embs]))
    gold["c1_refemb"] = refemb.clone()
    gold["c1_ref_embs"] = torch.cat(embs).clone()
    gold["c1_dist"] = dist.clone()
    gold["c1_thr"] = thr.clone()
    o_ref, o_thr = osc.reference_embedding(embs)
    assert torch.equal(o_ref, refemb) and torch.equal(o_thr, thr)
    assert torch.equal(osc.one_class_scores(refemb, emb), dist)
    gold["c1_lines"] = osc.score_lines_1c(dist.tolist(), float(thr))
    # batch-1 == batch-8 in the reference?  (truth for batched GPU results)
    with torch.no_grad():
        e1, o1 = model(wav[3:4])
    pin["ref_batch1_vs_batch8_emb_maxabs"] = float((e1 - emb[3:4]).abs().max())

    # config 4: variable length, each utterance alone at batch 1 (reference truth)
    lens = [16000, 23456, 40000, 64600, 100001]
    gold["vl_lens"] = torch.tensor(lens)
    vl_emb, vl_out = [], []
    for j, L in enumerate(lens):
        w = synth.make_waveforms(seed=30 + j, batch=1, n_samples=L)
        with torch.no_grad():
            e, o = model(w)
            oe, oo = osc.amodel_forward(sd, w)
        pin[f"vl{L}_emb_maxabs_vs_reference"] = float((oe - e).abs().max())
        vl_emb.append(e)
        vl_out.append(o)
    gold["vl_emb"] = torch.cat(vl_emb)
    gold["vl_out"] = torch.cat(vl_out)

    # config 5 (small): all hidden states, frontend only, 2 x 48000
    w5 = synth.make_waveforms(seed=50, batch=2, n_samples=48000)
    with torch.no_grad():
        hs5, _ = ta.extract_features(w5)
        y5, o5 = ofe.xlsr_forward(sd, w5, all_layers=True)
    pin["c5_layers_maxabs_vs_torchaudio"] = float(max((a - b).abs().max() for a, b in zip(hs5, o5)))
    gold["c5_layers_sub"] = torch.stack([sub(h, 1, 10, 32) for h in hs5])   # [24,2,15,32]

    for h in hs:
        h.remove()
    torch.save({k: (v if not torch.is_tensor(v) else v.contiguous()) for k, v in gold.items()},
               os.path.join(OUT, "model_golden.pt"))
    with open(os.path.join(OUT, "PINNING.json"), "w") as f:
        json.dump(pin, f, indent=1)
    print(json.dumps(pin, indent=1))
    print("golden bytes:", os.path.getsize(os.path.join(OUT, "model_golden.pt")))


if __name__ == "__main__":
    main()


def dump_state_dict_keys():
    """tests/golden/state_dict_keys.json: the reference AModel's backend state_dict keys/shapes (from the
    unmodified reference class) and the fairseq-named frontend keys the checkpoint carries under
    ``ssl_model.model.`` (fairseq itself is not importable; names follow import_fairseq.py's map, checked by
    loading them into torchaudio strictly in ``load_into_reference``)."""
    from occm_b200 import synth
    ref, _ = load_reference_modules()
    model = ref.AModel(None, "cpu")
    be = {k: list(v.shape) for k, v in model.state_dict().items() if not k.startswith("ssl_model.")}
    fe = {synth.SSL_PREFIX + k: list(s) for k, s in synth.frontend_shapes().items()}
    with open(os.path.join(OUT, "state_dict_keys.json"), "w") as f:
        json.dump({"backend": be, "frontend": fe}, f, indent=0)
    print("state_dict keys:", len(be), "backend,", len(fe), "frontend")
