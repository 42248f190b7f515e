"""LCNN-backend golden vectors from the UNMODIFIED reference `models/lcnn.py`, in the build container.

Run:  python -m oracle.gen_golden_r3      -> tests/golden/lcnn_golden.pt

The reference `LCNN(asoftmax=False)` is built from the unmodified file (fairseq stood in by the torchaudio shim, because
`lcnn.py` imports `xlsr.SSLModel` at module level), loaded with the seeded checkpoint `synth.make_lcnn_state_dict(0)`, and
run (a) on random feature tensors [B,T,1024] (backend alone) and (b) behind the torchaudio XLS-R stand-in on seeded waveforms
(`ssl_lcnn` surface).  The restatement `oracle/lcnn.py` is checked against it here (max-abs printed and stored).
"""
from __future__ import annotations

import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tests", "golden")


def main():
    from occm_b200 import synth
    from oracle.gen_golden import load_reference_modules
    from oracle import frontend as ofe, lcnn as olc

    torch.set_num_threads(os.cpu_count())
    load_reference_modules()                                  # installs the fairseq stand-in into sys.modules
    sys.path.insert(0, "/root/reference/models")
    ref_lcnn = importlib.import_module("lcnn")
    net = ref_lcnn.lcnn_net(asoftmax=False).eval()
    sd_l = synth.make_lcnn_state_dict(0)
    res = net.load_state_dict({k[len(synth.LCNN_PREFIX):]: v for k, v in sd_l.items()}, strict=True)
    gold = {"keys": {k: list(v.shape) for k, v in net.state_dict().items()}}
    g = torch.Generator().manual_seed(7)
    for tag, (B, T) in (("f49", (3, 49)), ("f201", (2, 201)), ("f499", (1, 499))):
        feat = torch.randn(B, T, 1024, generator=g)
        with torch.no_grad():
            want = net(feat.unsqueeze(1))
        got = olc.lcnn_backend(sd_l, feat)
        # (the feature tensors are regenerated in the tests: torch.Generator().manual_seed(7), shapes in this order)
        gold[f"{tag}_out"] = want.clone()
        gold[f"{tag}_oracle_maxabs"] = float((got - want).abs().max())
        print(tag, tuple(want.shape), "oracle vs reference max-abs", gold[f"{tag}_oracle_maxabs"], flush=True)
    # ssl_lcnn surface: waveform -> XLS-R (seeded synthetic checkpoint) -> LCNN
    sd = synth.make_state_dict(0, frontend=True, backend=False)
    for tag, (B, N) in (("w16k", (2, 16000)), ("w64k", (2, 64600))):
        wav = synth.make_waveforms(seed=60 + B + N // 1000, batch=B, n_samples=N)
        with torch.no_grad():
            feat = ofe.xlsr_forward(sd, wav)
            want = net(feat.unsqueeze(1))
        gold[f"{tag}_seed"] = 60 + B + N // 1000
        gold[f"{tag}_out"] = want.clone()
        print(tag, want.tolist(), flush=True)
    torch.save(gold, os.path.join(OUT, "lcnn_golden.pt"))
    print("lcnn_golden.pt bytes:", os.path.getsize(os.path.join(OUT, "lcnn_golden.pt")))

    # ------------------------------------------------------------------ SE-ResNet34 backend (models/senet.py), same recipe
    from oracle import senet as ose
    ref_senet = importlib.import_module("senet")
    rnet = ref_senet.se_resnet34().eval()
    sd_r = synth.make_resnet34_state_dict(0)
    rnet.load_state_dict({k[len(synth.RESNET_PREFIX):]: v for k, v in sd_r.items()}, strict=True)
    gr = {"keys": {k: list(v.shape) for k, v in rnet.state_dict().items()}}
    g = torch.Generator().manual_seed(8)
    for tag, (B, T) in (("f49", (3, 49)), ("f201", (2, 201)), ("f312", (1, 312))):
        feat = torch.randn(B, T, 1024, generator=g)
        with torch.no_grad():
            com, des = rnet(feat.unsqueeze(1))
        ocom, odes = ose.resnet34_backend(sd_r, feat)
        gr[f"{tag}_com"], gr[f"{tag}_des"] = com.clone(), des.clone()
        gr[f"{tag}_oracle_maxabs"] = float(max((ocom - com).abs().max(), (odes - des).abs().max()))
        print("senet", tag, "oracle vs reference max-abs", gr[f"{tag}_oracle_maxabs"], flush=True)
    for tag, (B, N) in (("w16k", (2, 16000)), ("w64k", (2, 64600))):
        wav = synth.make_waveforms(seed=60 + B + N // 1000, batch=B, n_samples=N)
        with torch.no_grad():
            feat = ofe.xlsr_forward(sd, wav)
            com, des = rnet(feat.unsqueeze(1))
        gr[f"{tag}_seed"] = 60 + B + N // 1000
        gr[f"{tag}_com"], gr[f"{tag}_des"] = com.clone(), des.clone()
    torch.save(gr, os.path.join(OUT, "senet_golden.pt"))
    print("senet_golden.pt bytes:", os.path.getsize(os.path.join(OUT, "senet_golden.pt")))


if __name__ == "__main__":
    main()
