"""dev: bf16-mode embedding error against fp32 mode on the config-3 set, with and without the LayerNorm fold (same process)."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from occm_b200 import synth, scoring
from occm_b200.models.sslassist import AModel
dev = torch.device("cuda", 0)
n, nb = 1024, synth.config3_n_bonafide(1024)
def run(dtype, fold):
    os.environ["OCC_LN_FOLD"] = str(fold)
    m = AModel(None, dev, dtype=dtype).to(dev).eval()
    m.reserve(64, 64600, dev)
    with torch.no_grad():
        tr = m(synth.config3_train_batch(dev))[0]
        embs = torch.cat([m(synth.config3_batch(a, min(a + 64, n), nb, dev))[0] for a in range(0, n, 64)])
    m._engine.close()
    return tr, embs
tr32, e32 = run("fp32", 0)
ref32 = tr32.mean(0, keepdim=True)
d32 = scoring.one_class_scores(ref32, e32)
for fold in (int(os.environ.get("DIAG_FOLD", "3")),):
    tr16, e16 = run("bf16", fold)
    l2 = ((e16 - e32).norm(dim=1) / e32.norm(dim=1))
    dd_true = scoring.one_class_scores(ref32, e16)
    dd_own = scoring.one_class_scores(tr16.mean(0, keepdim=True), e16)
    mv_true = ((dd_true - d32).abs() / d32)
    mv_own = ((dd_own - d32).abs() / d32)
    sgn = ((dd_own - d32) / d32)
    print(f"fold={fold}: emb L2 rel err median {l2.median():.3e} p90 {l2.quantile(0.9):.3e}; |emb err| abs median {(e16-e32).norm(dim=1).median():.3e}; "
          f"d32 median {d32.median():.4f}; moved(true ref) median {mv_true.median():.3e} >2%: {(mv_true>0.02).float().mean():.3f}; "
          f"moved(own ref) median {mv_own.median():.3e} >2%: {(mv_own>0.02).float().mean():.3f} signed mean {sgn.mean():+.3e}; ref shift {(tr16.mean(0)-tr32.mean(0)).norm():.3e}", flush=True)
