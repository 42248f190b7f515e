"""GEMM pipeline probes on the B200 (development tool): times the transformer GEMM shapes with parts of the
kernel switched off to see which pipeline (TMA loads, MMA issue, epilogue) bounds the tile time.

    python tools/gemm_probe.py
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from occm_b200 import _lib  # noqa: E402

lib = _lib.load()
M = 64 * 201
SHAPES = [("qkv", 3072, 1024, 0, 0, 1), ("out", 1024, 1024, 0, 0, 1), ("fc1", 4096, 1024, 1, 0, 1), ("fc2", 1024, 4096, 0, 0, 1),
          ("outr", 1024, 1024, 0, 1, 0), ("fc2r", 1024, 4096, 0, 1, 0), ("out32", 1024, 1024, 0, 0, 0)]   # name, N, K, gelu, resid, bf16-out
PROBES = [(0, "full"), (4, "no-epi"), (1, "no-tma"), (5, "mma-only"), (3, "epi-only"),
          (5 + (1 << 12), "mma-only-BN128"), (5 + (2 << 12), "mma-only-BN64")]
ITERS = 20
if len(sys.argv) > 1:
    M = int(sys.argv[1])
if len(sys.argv) > 2:      # gemm_probe.py M shape[,shape] probe[,probe] iters   (for ncu captures)
    SHAPES = [x for x in SHAPES if x[0] in sys.argv[2].split(",")]
if len(sys.argv) > 3:
    PROBES = [x for x in PROBES if str(x[0]) in sys.argv[3].split(",")]
if len(sys.argv) > 4:
    ITERS = int(sys.argv[4])
if len(sys.argv) <= 1:
    import torch
    for name, N, K, gelu, resid, o16 in SHAPES:
        for fill in ("zeros", "randn"):
            a = getattr(torch, fill)(M, K, device="cuda", dtype=torch.bfloat16)
            w = getattr(torch, fill)(N, K, device="cuda", dtype=torch.bfloat16)
            for _ in range(3):
                torch.matmul(a, w.t())
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                torch.matmul(a, w.t())
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 20
            print(f"cublas {name} {fill}: {ms*1e3:7.1f}us ({2.0*M*N*K/ms/1e9:6.0f}TF)", flush=True)
for name, N, K, gelu, resid, o16 in SHAPES:
    fl = 2.0 * M * N * K
    row = []
    for pb, pn in PROBES:
        ms = C.c_double()
        _lib.check(lib.occ_debug_gemm_bench(M, N, K, gelu, resid, o16, pb, ITERS, C.byref(ms)), None, "bench")
        row.append(f"{pn}={ms.value*1e3:7.1f}us({fl/ms.value/1e9:6.0f}TF)")
    print(f"{name:4s} M={M} N={N} K={K}: " + "  ".join(row), flush=True)
