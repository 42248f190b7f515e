"""dev probe: pos-conv stage vs oracle under the current OCC_HALO_BO / OCC_GEMM_NO_HALO setting."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from occm_b200 import synth, _lib
from occm_b200.models.sslassist import AModel
from oracle import frontend as ofe
DEV = torch.device("cuda", 0)
sd = synth.make_state_dict(0)
model = AModel(None, DEV, dtype="fp32").to(DEV).eval()
eng = model._get_engine(2, 16000, DEV)
lib, h = eng.lib, eng.handle
torch.manual_seed(5)
B, T = 2, 49
proj = torch.randn(B, T, 1024)
with torch.no_grad():
    want = proj + ofe.pos_conv(sd, proj)
out = torch.full((B, T, 1024), float("nan"), device=DEV)
x = proj.to(DEV).contiguous()
assert lib.occ_pos_conv(h, x.data_ptr(), B, T, out.data_ptr(), None) == 0
torch.cuda.synchronize()
d = (out.cpu() - want).abs()
print("env BO=%s NO_HALO=%s  rel err %.3e" % (os.environ.get("OCC_HALO_BO"), os.environ.get("OCC_GEMM_NO_HALO"), float(d.max() / want.abs().max())))
print("per-row max err (b=0):", [round(float(v), 4) for v in d[0].amax(dim=1)[:24]])
