"""dev: bf16-mode error of (emb, out) on BASELINE config 1 inputs against the committed golden tensors of the reference."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from occm_b200 import synth
from occm_b200.models.sslassist import AModel
g = torch.load(os.path.join(ROOT, "tests/golden/model_golden.pt"))
dev = torch.device("cuda", 0)
wav = synth.make_waveforms(seed=1, batch=8, n_samples=64600).to(dev)
for mode in ("bf16", "fp32"):
    m = AModel(None, dev, dtype=mode).to(dev).eval()
    with torch.no_grad():
        emb, out = m(wav)
    rel = lambda a, b: float((a.cpu() - b).abs().max() / b.abs().max())
    print(mode, "emb rel-inf err %.3e  out rel-inf err %.3e" % (rel(emb, g["c1_emb"]), rel(out, g["c1_out"])))
