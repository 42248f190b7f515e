#!/usr/bin/env python
"""The informative GPU bar (SURVEY section 2 / BASELINE.md section 3): the UNMODIFIED reference AModel (oracle/_ref, fairseq stood in
by torchaudio's XLS-R-300M) run by PyTorch eager on the same B200, fp32 and bf16 autocast, batch 8 and 64 x 64 600 samples.

    python tools/ref_eager_gpu.py            # one JSON line per (dtype, batch)

This is test / benchmark infrastructure (it imports oracle/): it times the reference, never the product.
"""
from __future__ import annotations

import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.nn.functional as F
    from occm_b200 import synth
    from oracle import ref_runner

    dev = torch.device("cuda", 0)
    model = ref_runner.load_reference_amodel(dev)
    ref = torch.zeros(1, 160, device=dev)
    for dtype in ("fp32", "tf32", "bf16_autocast"):
        torch.backends.cuda.matmul.allow_tf32 = dtype == "tf32"
        torch.backends.cudnn.allow_tf32 = dtype == "tf32"
        for B in (8, 64):
            wav = synth.make_waveforms(1, B, 64600).to(dev)

            def step():
                with torch.no_grad():
                    if dtype == "bf16_autocast":
                        with torch.autocast("cuda", dtype=torch.bfloat16):
                            emb, out = model(wav)
                    else:
                        emb, out = model(wav)
                    return F.pairwise_distance(ref, emb.float(), p=2)
            for _ in range(3):
                step()
            torch.cuda.synchronize(dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            K = 10
            e0.record()
            for _ in range(K):
                d = step()
            e1.record()
            torch.cuda.synchronize(dev)
            ms = e0.elapsed_time(e1) / K
            print(json.dumps({"impl": "reference eager on B200 (unmodified models/sslassist.py, torchaudio XLS-R stand-in for fairseq)",
                              "dtype": dtype, "batch": B, "ms_per_step": ms, "utterances_per_s": B / (ms / 1e3),
                              "finite": bool(torch.isfinite(d).all())}), flush=True)


if __name__ == "__main__":
    main()
