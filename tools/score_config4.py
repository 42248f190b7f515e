#!/usr/bin/env python
"""BASELINE.json configs[3]: variable-length utterances 1-10 s, length-bucketed with padding masks.

    python tools/score_config4.py [--utterances 2048] [--dtype bf16] [--max-batch 64]

N_i ~ U{16 000..160 000} samples (seed 3, occm_b200.synth.config4_lengths), white noise sigma 0.1, generated on the device
before the timed region.  scoring.score_varlen sorts by length, cuts buckets (scoring.bucket_by_length), pads each bucket to
its longest utterance and passes the lengths (masks in the conv encoder / attention / backend) -- each utterance's result is
the reference's batch-1 un-padded one (oc_classifier.py:350,356; parity: tests/test_gpu_round2.py).  Prints one JSON line:
utt/s AND audio-seconds/s (device-timed), padding waste, bucket count.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--utterances", type=int, default=2048)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--max-batch", type=int, default=64)
    ap.add_argument("--repeats", type=int, default=3)
    a = ap.parse_args()
    import torch
    from occm_b200 import scoring, synth
    from occm_b200.models.sslassist import AModel

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    lens = synth.config4_lengths(a.utterances)
    gen = torch.Generator(device=dev)
    waves = []
    for i, L in enumerate(lens):
        gen.manual_seed(4000 + i)
        waves.append(torch.empty(L, device=dev).normal_(0.0, 0.1, generator=gen))
    model = AModel(None, dev, dtype=a.dtype).to(dev).eval()
    budget = 64 * 64600
    model.reserve(a.max_batch, max(lens), dev)
    ref = torch.zeros(1, 160, device=dev)
    buckets = scoring.bucket_by_length(lens, a.max_batch, budget)
    padded = sum(len(b) * max(lens[i] for i in b) for b in buckets)
    scoring.score_varlen(model, waves, dev, ref, a.max_batch, budget)          # warm-up pass
    torch.cuda.synchronize(dev)
    best = None
    for _ in range(a.repeats):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        emb, out, d = scoring.score_varlen(model, waves, dev, ref, a.max_batch, budget)
        e1.record()
        torch.cuda.synchronize(dev)
        ms = e0.elapsed_time(e1)
        best = ms if best is None else min(best, ms)
    audio_s = sum(lens) / 16000.0
    print(json.dumps({"config": "BASELINE.json configs[3] (variable length 1-10 s, length-bucketed, padding masks)", "dtype": a.dtype,
                      "n_utterances": a.utterances, "audio_seconds": audio_s, "seconds": best / 1e3,
                      "utterances_per_s": a.utterances / (best / 1e3), "audio_seconds_per_s": audio_s / (best / 1e3),
                      "buckets": len(buckets), "padding_waste": 1.0 - sum(lens) / padded, "max_batch": a.max_batch,
                      "max_padded_samples": budget, "finite": bool(torch.isfinite(emb).all()),
                      "note": "timed region = bucket padding copies + forward + one-class score, best of %d passes, CUDA events" % a.repeats}),
          flush=True)


if __name__ == "__main__":
    main()
