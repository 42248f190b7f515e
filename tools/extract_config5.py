#!/usr/bin/env python
"""BASELINE.json configs[4]: XLS-R frontend only, 30 s audio, batch 16, all 24 hidden states exported.

    python tools/extract_config5.py [--batch 16] [--samples 480000] [--dtype bf16] [--steps 5]

SSLModel.extract_feat(x, all_layers=True) (models/xlsr.py:39-52 surface; torchaudio components.py:445-463 semantics): T = 1 499
frames, attention on the KV-chunked tcgen05 kernel, every layer's output written by the FC2 epilogue into the caller's
[24, B, T, 1024] fp32 tensor (2.36 GB per batch).  Prints one JSON line: audio-s/s, utt/s, ms/step, export write GB/s
(export bytes / whole step, and the per-family device times).
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
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--samples", type=int, default=480000)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--steps", type=int, default=5)
    a = ap.parse_args()
    import torch
    from occm_b200.models.xlsr import SSLModel

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    m = SSLModel(dev, dtype=a.dtype)
    m.reserve(a.batch, a.samples, dev)
    gen = torch.Generator(device=dev).manual_seed(5)
    x = torch.empty(a.batch, a.samples, device=dev).normal_(0.0, 0.1, generator=gen)
    res = {}
    for all_layers in (True, False):
        for _ in range(2):
            m.extract_feat(x, all_layers=all_layers)
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            y = m.extract_feat(x, all_layers=all_layers)
        e1.record()
        torch.cuda.synchronize(dev)
        res[all_layers] = e0.elapsed_time(e1) / a.steps
    feat, hid = m.extract_feat(x, all_layers=True)
    eng = m._engine
    eng.profile(True)
    m.extract_feat(x, all_layers=True)
    torch.cuda.synchronize(dev)
    eng.profile(False)
    prof = {k: round(v["ms"], 3) for k, v in eng.profile_read().items()}
    T = feat.shape[1]
    export_bytes = 24 * a.batch * T * 1024 * 4
    ms = res[True]
    print(json.dumps({"config": "BASELINE.json configs[4] (XLS-R frontend only, long-form audio, all 24 hidden states)", "dtype": a.dtype,
                      "batch": a.batch, "n_samples": a.samples, "frames": T, "ms_per_step": ms, "ms_per_step_without_export": res[False],
                      "utterances_per_s": a.batch / (ms / 1e3), "audio_seconds_per_s": a.batch * a.samples / 16000.0 / (ms / 1e3),
                      "export_bytes_per_step": export_bytes, "export_cost_ms": ms - res[False],
                      "export_write_gbs_over_its_cost": export_bytes / max(ms - res[False], 1e-3) / 1e6,
                      "family_ms": prof, "finite": bool(torch.isfinite(hid).all() and torch.isfinite(feat).all()),
                      "tflops_model": a.batch * 1300.2e9 * (a.samples / 480000.0) / (ms / 1e3) / 1e12}), flush=True)


if __name__ == "__main__":
    main()
