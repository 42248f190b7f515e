#!/usr/bin/env python
"""SURVEY 8(f) N4 measurement: device-timed utterances/s of the other backends on the shared frontend (ssl_lcnn, ssl_resnet34, OCCM)
next to AModel, B = 64 x 64 600 samples, bf16 frontend, one B200; plus the backend-only time (features resident).

    python tools/bench_backends.py [--batch 64] [--steps 10]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed(fn, steps, dev):
    import torch
    for _ in range(3):
        fn()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize(dev)
    return e0.elapsed_time(e1) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=10)
    a = ap.parse_args()
    import torch
    from occm_b200 import synth
    from occm_b200.models.lcnn import ssl_lcnn
    from occm_b200.models.occm import OCCM
    from occm_b200.models.senet import ssl_resnet34
    from occm_b200.models.sslassist import AModel
    dev = torch.device("cuda", 0)
    B = a.batch
    wav = synth.make_waveforms(1000, B, 64600).to(dev)
    out = {}
    with torch.no_grad():
        m = AModel(None, dev, dtype="bf16").to(dev).eval()
        out["AModel (AASIST)"] = timed(lambda: m(wav), a.steps, dev)
        feat = m.ssl_model.extract_feat(wav)
        out["frontend only"] = timed(lambda: m.ssl_model.extract_feat(wav), a.steps, dev)
        m._engine.close()
        m = ssl_lcnn(dev, dtype="bf16")
        out["ssl_lcnn"] = timed(lambda: m(wav), a.steps, dev)
        out["LCNN backend alone"] = timed(lambda: m.lcnn_forward(feat), a.steps, dev)
        m._engine.close()
        m = ssl_resnet34(dev, dtype="bf16")
        out["ssl_resnet34"] = timed(lambda: m(wav), a.steps, dev)
        out["SE-ResNet34 backend alone"] = timed(lambda: m.resnet34_forward(feat), a.steps, dev)
        m._engine.close()
        m = OCCM(dev, dtype="bf16")
        out["OCCM (both branches)"] = timed(lambda: m(wav), a.steps, dev)
        m._engine.close()
    print(json.dumps({"config": f"B = {B} x 64 600 samples, bf16 frontend, one B200, CUDA events, {a.steps} steps after 3 warm-ups",
                      "ms_per_step": {k: round(v, 3) for k, v in out.items()},
                      "utterances_per_s": {k: round(B / (v / 1e3), 1) for k, v in out.items()}}), flush=True)


if __name__ == "__main__":
    main()
