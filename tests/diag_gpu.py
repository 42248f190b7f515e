"""GPU bring-up diagnostics: runs each kernel family in its own subprocess (a sticky CUDA fault in one
group cannot poison the others) and prints the error against torch / the oracle.  Development tool only.

    python tests/diag_gpu.py            # all groups
    python tests/diag_gpu.py gemm attn  # selected groups
"""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

GROUPS = ["eer", "gemm_simt", "gemm", "attn", "conv", "pos", "xf", "backend", "full_fp32", "full_bf16", "varlen"]


def rel(a, b):
    import torch
    a = a.float().cpu()
    b = b.float().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def make_engine(dtype, B, N, sd=None):
    import torch
    from occm_b200 import synth
    from occm_b200.models.sslassist import _Engine
    sd = sd or synth.make_state_dict(0)
    dev = torch.device("cuda", 0)
    t0 = time.time()
    e = _Engine({k: v for k, v in sd.items()}, dev, dtype, B, N)
    torch.cuda.synchronize()
    print(f"  engine built in {time.time() - t0:.1f}s")
    return e, sd


def g_eer():
    import numpy as np
    from occm_b200 import _lib
    from oracle import scoring as osc
    lib = _lib.load()
    kats = json.load(open(os.path.join(ROOT, "tests/golden/eer_kat.json")))
    for k in kats:
        if "target" in k:
            t, n = np.array(k["target"], dtype=np.float64), np.array(k["nontarget"], dtype=np.float64)
        elif k["name"] == "KAT3":
            rng = np.random.default_rng(1234)
            t = rng.normal(1, 1, 7355).astype(np.float32).astype(np.float64)
            n = rng.normal(-1, 1, 63882).astype(np.float32).astype(np.float64)
        else:
            r = np.random.default_rng(k["seed"])
            q = k["q"]
            t = np.round(r.normal(0.3, 1, k["nt"]) * q) / q
            n = np.round(r.normal(-0.3, 1, k["nn"]) * q) / q
        e, th = C.c_double(), C.c_double()
        rc = lib.occ_eer_host(t.ctypes.data, t.size, n.ctypes.data, n.size, C.byref(e), C.byref(th), 0)
        ok = rc == 0 and e.value == k["eer"] and th.value == k["thr"]
        print(f"  {k['name']}: rc={rc} got=({e.value!r},{th.value!r}) want=({k['eer']!r},{k['thr']!r}) {'OK' if ok else 'MISMATCH'}")


def _gemm(impl):
    import torch
    from occm_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    for (M, N, K, split, gelu, use_res) in [(128, 64, 64, 0, 0, 0), (128, 256, 64, 0, 0, 0), (300, 256, 128, 0, 0, 0),
                                            (402, 1024, 1024, 0, 0, 1), (402, 3072, 1024, 0, 0, 0), (402, 4096, 1024, 0, 1, 0),
                                            (402, 1024, 4096, 0, 0, 1), (402, 128, 1024, 1, 0, 0), (1000, 1024, 512, 1, 1, 1),
                                            (12864, 1024, 1024, 0, 0, 1)]:
        A = torch.randn(M, K, device=dev)
        W = torch.randn(N, K, device=dev) * 0.05
        bias = torch.randn(N, device=dev)
        res = torch.randn(M, N, device=dev) if use_res else None
        out = torch.full((M, N), float("nan"), device=dev)
        rc = lib.occ_debug_gemm(impl, M, N, K, split, A.data_ptr(), W.data_ptr(), bias.data_ptr(),
                                res.data_ptr() if res is not None else None, gelu, out.data_ptr(), None)
        if rc != 0:
            print(f"  M{M} N{N} K{K} split{split}: rc={rc} {_lib.last_error()}")
            continue
        if split:
            ref = A.double() @ W.double().t()
        else:
            ref = A.bfloat16().double() @ W.bfloat16().double().t()
        ref = ref + bias.double()
        if gelu:
            ref = torch.nn.functional.gelu(ref)
        if res is not None:
            ref = ref + res.double()
        print(f"  impl{impl} M{M} N{N} K{K} split{split} gelu{gelu} res{use_res}: rel={rel(out, ref):.3e} nan={int(torch.isnan(out).sum())}")


def g_gemm_simt():
    _gemm(0)


def g_gemm():
    _gemm(1)


def g_attn():
    import torch
    from occm_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    for B, T, masked in [(2, 49, False), (2, 201, False), (3, 201, True), (1, 499, False)]:
        qkv = torch.randn(B * T, 3072, device=dev)
        qkv[:, :1024] *= 0.125
        klen = torch.tensor([T - 7 * i for i in range(B)], dtype=torch.int32, device=dev) if masked else None
        q, k, v = (qkv[:, i * 1024:(i + 1) * 1024].reshape(B, T, 16, 64).transpose(1, 2).double() for i in range(3))
        s = q @ k.transpose(-1, -2)
        if masked:
            m = torch.arange(T, device=dev)[None, :] >= klen[:, None]
            s = s.masked_fill(m[:, None, None, :], float("-inf"))
        ref = (torch.softmax(s, -1) @ v).transpose(1, 2).reshape(B * T, 1024)
        for impl in (0, 1):
            ctx = torch.full((B * T, 1024), float("nan"), device=dev)
            rc = lib.occ_debug_attention(impl, qkv.data_ptr(), B, T, klen.data_ptr() if masked else None, ctx.data_ptr(), None)
            if rc != 0:
                print(f"  attn impl{impl} B{B} T{T}: rc={rc} {_lib.last_error()}")
                continue
            if masked:
                valid = (torch.arange(T, device=dev)[None, :] < klen[:, None]).reshape(-1)
                r = rel(ctx[valid], ref[valid])
            else:
                r = rel(ctx, ref)
            print(f"  attn impl{impl} B{B} T{T} masked={masked}: rel={r:.3e}")


def _stage_inputs(B=2, N=16000):
    import torch
    from occm_b200 import synth
    from oracle import frontend as ofe
    sd = synth.make_state_dict(0)
    wav = synth.make_waveforms(5, B, N)
    with torch.no_grad():
        feats = ofe.conv_feature_encoder(sd, wav)
        proj = ofe.feature_projection(sd, feats)
    return sd, wav, proj


def _run_stage(name, dtype):
    import torch
    from occm_b200 import _lib
    from oracle import frontend as ofe, backend as obe
    dev = torch.device("cuda", 0)
    B, N = 2, 16000
    sd, wav, proj = _stage_inputs(B, N)
    e, _ = make_engine(dtype, B, N, sd)
    lib, h = e.lib, e.handle
    T = proj.shape[1]
    tag = "fp32" if dtype == 0 else "bf16"
    with torch.no_grad():
        if name == "conv":
            out = torch.full((B, T, 1024), float("nan"), device=dev)
            rc = lib.occ_conv_encoder(h, wav.to(dev).data_ptr(), B, N, out.data_ptr(), None)
            torch.cuda.synchronize()
            print(f"  conv_encoder[{tag}] rc={rc} {_lib.last_error(h) if rc else ''} rel={rel(out, proj):.3e}")
        elif name == "pos":
            out = torch.full((B, T, 1024), float("nan"), device=dev)
            x = proj.to(dev).contiguous()
            rc = lib.occ_pos_conv(h, x.data_ptr(), B, T, out.data_ptr(), None)
            torch.cuda.synchronize()
            ref = proj + ofe.pos_conv(sd, proj)
            print(f"  pos_conv[{tag}] rc={rc} {_lib.last_error(h) if rc else ''} rel={rel(out, ref):.3e}  (pos term rel={rel(out.cpu() - proj, ref - proj):.3e})")
        elif name == "xf":
            x0 = proj + ofe.pos_conv(sd, proj)
            for nl in (1, 4):
                x = x0
                for i in range(nl):
                    x = ofe.encoder_layer(sd, x, i, None)
                ref = torch.nn.functional.layer_norm(x, (1024,), sd[ofe.P + "encoder.layer_norm.weight"], sd[ofe.P + "encoder.layer_norm.bias"])
                out = torch.full((B, T, 1024), float("nan"), device=dev)
                hid = torch.full((nl, B, T, 1024), float("nan"), device=dev)
                xin = x0.to(dev).contiguous()
                rc = lib.occ_transformer(h, xin.data_ptr(), B, T, nl, out.data_ptr(), hid.data_ptr(), None)
                torch.cuda.synchronize()
                print(f"  transformer[{tag}] layers={nl} rc={rc} {_lib.last_error(h) if rc else ''} rel(final LN)={rel(out, ref):.3e} rel(hidden last)={rel(hid[-1], x):.3e}")
        elif name == "backend":
            feat = ofe.xlsr_forward(sd, wav, n_layers=2)     # any [B,T,1024] tensor of the right scale works as backend input
            st = {}
            o_emb, o_out = obe.aasist_backend(sd, feat, st)
            emb = torch.full((B, 160), float("nan"), device=dev)
            out = torch.full((B, 2), float("nan"), device=dev)
            f = feat.to(dev).contiguous()
            rc = lib.occ_aasist_backend(h, f.data_ptr(), B, T, emb.data_ptr(), out.data_ptr(), None)
            torch.cuda.synchronize()
            print(f"  backend[{tag}] rc={rc} {_lib.last_error(h) if rc else ''} emb rel={rel(emb, o_emb):.3e} out rel={rel(out, o_out):.3e}")
    e.close()


def g_conv():
    _run_stage("conv", 0)
    _run_stage("conv", 1)


def g_pos():
    _run_stage("pos", 0)
    _run_stage("pos", 1)


def g_xf():
    _run_stage("xf", 0)
    _run_stage("xf", 1)


def g_backend():
    _run_stage("backend", 1)


def _full(dtype, B=2, N=16000, lengths=None):
    import torch
    from occm_b200 import synth
    from occm_b200.models.sslassist import AModel
    from oracle import scoring as osc
    dev = torch.device("cuda", 0)
    sd = synth.make_state_dict(0)
    wav = synth.make_waveforms(7, B, N)
    model = AModel(None, dev, dtype=dtype).to(dev).eval()
    t0 = time.time()
    emb, out = model(wav.to(dev), lengths=lengths)
    torch.cuda.synchronize()
    t1 = time.time()
    emb2, out2 = model(wav.to(dev), lengths=lengths)
    torch.cuda.synchronize()
    print(f"  first forward {t1 - t0:.2f}s, second {time.time() - t1:.3f}s, launches={model.kernel_launches}, repeatable={bool(torch.equal(emb, emb2))}")
    if lengths is None:
        o_emb, o_out = osc.amodel_forward(sd, wav)
    else:
        oe, oo = zip(*[osc.amodel_forward(sd, wav[i:i + 1, :lengths[i]]) for i in range(B)])
        o_emb, o_out = torch.cat(oe), torch.cat(oo)
    print(f"  full[{dtype}] emb rel={rel(emb, o_emb):.3e} out rel={rel(out, o_out):.3e}")
    per = ((emb.cpu() - o_emb).abs().amax(1) / o_emb.abs().amax(1)).tolist()
    print("  per-utt emb rel:", ["%.2e" % v for v in per])


def g_full_fp32():
    _full("fp32")


def g_full_bf16():
    _full("bf16")


def g_varlen():
    _full("fp32", B=3, N=40000, lengths=[40000, 16000, 23456])


def main():
    sel = sys.argv[1:] or GROUPS
    if len(sel) == 2 and sel[0] == "--group":
        print(f"== {sel[1]}", flush=True)
        try:
            globals()["g_" + sel[1]]()
        except Exception as ex:  # noqa: BLE001
            import traceback
            traceback.print_exc()
            print(f"  GROUP {sel[1]} FAILED: {ex}")
        return
    for g in sel:
        t0 = time.time()
        try:
            p = subprocess.run([sys.executable, os.path.abspath(__file__), "--group", g], timeout=600, capture_output=True, text=True)
            print(p.stdout, end="")
            if p.returncode != 0 or "Traceback" in p.stderr or "occ:" in p.stderr:
                print("  stderr tail:", p.stderr[-1500:])
        except subprocess.TimeoutExpired:
            print(f"== {g}: TIMEOUT")
        print(f"  ({g}: {time.time() - t0:.1f}s)", flush=True)


if __name__ == "__main__":
    main()
