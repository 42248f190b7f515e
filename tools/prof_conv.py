"""The conv feature encoder alone for ncu (occ_conv_encoder through the C ABI): warm-up, then ONE stage call bracketed by
cudaProfilerStart/Stop.   ncu --profile-from-start off ... python tools/prof_conv.py [--batch 64]      (development tool)"""
from __future__ import annotations

import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--samples", type=int, default=64600)
    a = ap.parse_args()
    import torch
    from occm_b200 import _lib, synth
    from occm_b200.models.sslassist import AModel
    dev = torch.device("cuda", 0)
    model = AModel(None, dev, dtype="bf16").to(dev).eval()
    model.reserve(a.batch, a.samples, dev)
    eng = model._engine
    wav = synth.make_waveforms(1000, a.batch, a.samples).to(dev)
    T = eng.lib.occ_num_frames(a.samples)
    out = torch.empty(a.batch, T, 1024, device=dev)
    for _ in range(2):
        _lib.check(eng.lib.occ_conv_encoder(eng.handle, wav.data_ptr(), a.batch, a.samples, out.data_ptr(), None), eng.handle, "conv")
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    _lib.check(eng.lib.occ_conv_encoder(eng.handle, wav.data_ptr(), a.batch, a.samples, out.data_ptr(), None), eng.handle, "conv")
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print("finite:", bool(torch.isfinite(out).all()))


if __name__ == "__main__":
    main()
