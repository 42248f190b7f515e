"""One scoring step for ncu: warm-up forwards, then ONE forward bracketed by cudaProfilerStart/Stop.

    python tools/prof_step.py [--batch 64] [--dtype bf16] [--warm 2]
    ncu --profile-from-start off ... python tools/prof_step.py ...

Development tool only (not on the product path).  Prints the per-family CUDA-event split of one step.
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
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--samples", type=int, default=64600)
    ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--warm", type=int, default=2)
    a = ap.parse_args()
    import torch
    from occm_b200 import synth
    from occm_b200.models.sslassist import AModel
    from occm_b200.scoring import one_class_scores
    dev = torch.device("cuda", 0)
    model = AModel(None, dev, dtype=a.dtype).to(dev).eval()
    model.reserve(a.batch, a.samples, dev)
    wav = synth.make_waveforms(1000, a.batch, a.samples).to(dev)
    ref = torch.zeros(1, 160, device=dev)
    with torch.no_grad():
        for _ in range(a.warm):
            emb, out = model(wav)
        torch.cuda.synchronize()
        eng = model._engine
        eng.profile(True)
        model(wav)
        torch.cuda.synchronize()
        eng.profile(False)
        prof = eng.profile_read()
        torch.cuda.profiler.start()
        emb, out = model(wav)
        d = one_class_scores(ref, emb)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    print(json.dumps({k: {"ms": round(v["ms"], 3), "n": v["launches"],
                          "tflops": round(v["flops"] / max(v["ms"], 1e-9) / 1e9, 1)} for k, v in prof.items()}))
    import hashlib
    print("emb finite:", bool(torch.isfinite(emb).all()), "score0:", float(d[0]),
          "emb sha:", hashlib.sha256(emb.cpu().numpy().tobytes()).hexdigest()[:16])


if __name__ == "__main__":
    main()
