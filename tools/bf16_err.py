"""dev: bf16-mode error of the frontend features (smooth) and of (emb, out) (top-k pooling makes these jumpy) against the
fp32 mode of the same library and the committed golden tensors of the reference.   python tools/bf16_err.py [repo_root]"""
import os, sys, torch
ROOT = os.path.abspath(sys.argv[1]) if len(sys.argv) > 1 else os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from occm_b200 import synth
from occm_b200.models.sslassist import AModel
HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
g = torch.load(os.path.join(HERE, "tests/golden/model_golden.pt"))
dev = torch.device("cuda", 0)
rel = lambda a, b: float((a.float().cpu() - b.float().cpu()).abs().max() / b.float().cpu().abs().max())
res = {}
for name, wav in (("config1 8x64600", synth.make_waveforms(seed=1, batch=8, n_samples=64600)), ("smoke 2x16000", synth.make_waveforms(seed=11, batch=2, n_samples=16000))):
    wav = wav.to(dev)
    out = {}
    for mode in ("fp32", "bf16"):
        m = AModel(None, dev, dtype=mode).to(dev).eval()
        with torch.no_grad():
            emb, o = m(wav)
            feat = m.ssl_model.extract_feat(wav)
        out[mode] = (emb, o, feat)
        if m._engine is not None: m._engine.close()
    print(name, "bf16 vs fp32 mode: feat %.3e  emb %.3e  out %.3e" % (rel(out["bf16"][2], out["fp32"][2]), rel(out["bf16"][0], out["fp32"][0]), rel(out["bf16"][1], out["fp32"][1])))
    if name.startswith("config1"):
        print("   vs reference golden: fp32 emb %.3e   bf16 emb %.3e" % (rel(out["fp32"][0], g["c1_emb"]), rel(out["bf16"][0], g["c1_emb"])))
