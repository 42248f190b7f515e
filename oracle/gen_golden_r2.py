"""Round-2 golden vectors from the UNMODIFIED reference, generated in the build container.

Run:  python -m oracle.gen_golden_r2          (needs /root/reference; ~3 min on 8 cores)

Adds (without touching the round-1 files, which reproduce bit-for-bit):
  tests/golden/model_golden_r2.pt
    * c2_*    BASELINE config 2's benchmarked shape: B = 64 x 64 600 samples through the reference AModel (fp32, CPU)
    * l160_* / l480_*   one 10 s (T = 499) and one 30 s (T = 1 499) utterance: reference emb / out and torchaudio's
                        extract_features (24 hidden states, sub-sampled) + final features  (configs 4 / 5 at their stated sizes)
    * c4_*    16 utterances with N_i ~ U{16 000..160 000} (seed 3), each ALONE at batch 1 through the reference (config 4 truth)
    * c3_*    32 utterances of the config-3 synthetic set (16 bona fide + 16 spoof, generated on the CPU) and their reference
              embeddings, so the GPU scores of the same waveforms can be checked against the reference
  tests/golden/s2_join_kat.json
    * the stdout of the reference's own `calculate_eer.py` (pandas merge + compute_eer) on a small protocol / score-file pair
      with shuffled order, missing and duplicated utterances
  tests/golden/det_kat.json
    * the reference's compute_det_curve (frr, far, thresholds) on KAT vectors
The GPU box has no /root/reference: tests there read only the committed files.
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
import tempfile

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")


def sub(t, *steps):
    idx = tuple(slice(None, None, s) for s in steps)
    return t[idx].contiguous().clone()


def s2_join_kat(em):
    rng = np.random.default_rng(2024)
    n = 400
    utts = [f"LA_E_{1000000 + i}" for i in range(n)]
    labels = ["bonafide" if rng.random() < 0.25 else "spoof" for _ in range(n)]
    scores = np.where(np.array(labels) == "bonafide", rng.normal(1.0, 1.0, n), rng.normal(-0.5, 1.0, n)).astype(np.float32)
    proto_lines = [f"LA_{rng.integers(1, 50):04d} {u} - {'-' if l == 'bonafide' else 'A' + str(rng.integers(7, 20)).zfill(2)} {l}"
                   for u, l in zip(utts, labels)]
    order = rng.permutation(n)
    keep = [i for i in order if i % 17 != 3]                    # some utterances have no score
    score_lines = [f"{utts[i]} {float(scores[i])}" for i in keep]
    score_lines.insert(5, f"{utts[keep[40]]} {float(scores[keep[40]]) + 0.25}")     # one utterance scored twice
    score_lines.append("LA_E_9999999 0.5")                     # a score without a protocol row
    with tempfile.TemporaryDirectory() as td:
        pp, sp = os.path.join(td, "proto.txt"), os.path.join(td, "scores.txt")
        open(pp, "w").write("\n".join(proto_lines) + "\n")
        open(sp, "w").write("\n".join(score_lines) + "\n")
        r = subprocess.run([sys.executable, os.path.join(REF, "calculate_eer.py"), "--eval_protocol_file", pp, "--score_file", sp],
                           cwd=REF, capture_output=True, text=True, check=True)
    last = [l for l in r.stdout.splitlines() if l.startswith("EER = ")][-1]
    return dict(protocol=proto_lines, scores=score_lines, stdout_last_line=last)


def main():
    from occm_b200 import synth
    from oracle.gen_golden import load_reference_modules, load_into_reference
    from oracle import scoring as osc

    torch.set_num_threads(os.cpu_count())
    ref, em = load_reference_modules()

    # ------------------------------------------------------------------ file-format / curve KATs (no model)
    with open(os.path.join(OUT, "s2_join_kat.json"), "w") as f:
        json.dump(s2_join_kat(em), f, indent=0)
    dets = []
    for name, t, n in (("KAT1", [0.9, 0.8, 0.35, 0.7], [0.1, 0.4, 0.35, 0.2, 0.05]), ("KAT2", [0.5] * 3, [0.5] * 2)):
        frr, far, thr = em.compute_det_curve(np.asarray(t, dtype=np.float64), np.asarray(n, dtype=np.float64))
        dets.append(dict(name=name, target=t, nontarget=n, frr=frr.tolist(), far=far.tolist(), thr=thr.tolist()))
    r = np.random.default_rng(7)
    tt = np.round(r.normal(0.3, 1, 300) * 16) / 16
    nn = np.round(r.normal(-0.3, 1, 500) * 16) / 16
    frr, far, thr = em.compute_det_curve(tt, nn)
    o_frr, o_far, o_thr = osc.compute_det_curve(tt, nn)
    assert np.array_equal(frr, o_frr) and np.array_equal(far, o_far) and np.array_equal(thr, o_thr)
    dets.append(dict(name="ties7", gen="default_rng(7): round(normal(.3,1,300)*16)/16; round(normal(-.3,1,500)*16)/16",
                     frr=frr.tolist(), far=far.tolist(), thr=thr.tolist()))
    with open(os.path.join(OUT, "det_kat.json"), "w") as f:
        json.dump(dets, f)
    print("s2_join_kat.json, det_kat.json written")

    # ------------------------------------------------------------------ model goldens
    sd = synth.make_state_dict(seed=0)
    model = ref.AModel(None, "cpu").eval()
    load_into_reference(model, sd)
    ta = model.ssl_model.model.m
    gold = {}
    g1 = torch.load(os.path.join(OUT, "model_golden.pt"), weights_only=False)
    import torch.nn.functional as F

    # config 2: the benchmarked shape, B = 64 x 64 600
    wav = synth.make_waveforms(seed=64, batch=64, n_samples=64600)
    with torch.no_grad():
        emb, out = model(wav)
    gold["c2_emb"], gold["c2_out"] = emb.clone(), out.clone()
    gold["c2_dist"] = torch.cat([F.pairwise_distance(g1["c1_refemb"], emb[i:i + 1], p=2) for i in range(64)])
    print("c2 done", flush=True)

    # long utterances: 10 s and 30 s
    for tag, n in (("l160", 160000), ("l480", 480000)):
        w = synth.make_waveforms(seed=n // 1000, batch=1, n_samples=n)
        with torch.no_grad():
            e, o = model(w)
            hs, _ = ta.extract_features(w)
            feat = model.ssl_model.extract_feat(w)
        gold[f"{tag}_emb"], gold[f"{tag}_out"] = e.clone(), o.clone()
        gold[f"{tag}_layers_sub"] = torch.stack([sub(h, 1, 10, 32) for h in hs])       # [24,1,ceil(T/10),32]
        gold[f"{tag}_feat_sub"] = sub(feat, 1, 10, 16)
        print(tag, "done", tuple(feat.shape), flush=True)

    # config 4: N_i ~ U{16 000..160 000}, seed 3, each utterance alone at batch 1 (the reference's own regime)
    lens = synth.config4_lengths(16)
    gold["c4_lens"] = torch.tensor(lens)
    ce, co = [], []
    for j, L in enumerate(lens):
        w = synth.make_waveforms(seed=4000 + j, batch=1, n_samples=L)
        with torch.no_grad():
            e, o = model(w)
        ce.append(e)
        co.append(o)
    gold["c4_emb"], gold["c4_out"] = torch.cat(ce), torch.cat(co)
    print("c4 done", flush=True)

    # config 3 subset: 16 bona fide + 16 spoof utterances of the synthetic set (CPU generator), reference embeddings
    nb = 16
    w3 = synth.config3_batch(0, 32, nb, "cpu")
    tr = synth.config3_train_batch("cpu", 32)
    with torch.no_grad():
        e3 = torch.cat([model(w3[i:i + 8])[0] for i in range(0, 32, 8)])
        etr = torch.cat([model(tr[i:i + 8])[0] for i in range(0, 32, 8)])
    gold["c3_emb"] = e3.clone()
    gold["c3_ref"] = etr.mean(dim=0, keepdim=True)
    d3 = torch.cat([F.pairwise_distance(gold["c3_ref"], e3[i:i + 1], p=2) for i in range(32)])
    gold["c3_dist"] = d3.clone()
    eer, thr = em.compute_eer(d3[nb:].double().numpy(), d3[:nb].double().numpy())
    gold["c3_eer_32"] = (float(eer), float(thr))
    print("c3 subset: EER over 16+16 =", eer, flush=True)

    torch.save({k: (v.contiguous() if torch.is_tensor(v) else v) for k, v in gold.items()}, os.path.join(OUT, "model_golden_r2.pt"))
    print("golden r2 bytes:", os.path.getsize(os.path.join(OUT, "model_golden_r2.pt")))


if __name__ == "__main__":
    main()
