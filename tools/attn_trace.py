"""Development tool: clock-stamp timeline of attn_tc2 (OCC_ATTN_FLAGS bit 2 must be set in the environment).

    OCC_ATTN_FLAGS=7 python tools/attn_trace.py [--batch 64] [--frames 201]
Prints, per traced CTA, each work item's events in cycles relative to the CTA's first stamp.
"""
import argparse
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--frames", type=int, default=201)
    ap.add_argument("--ctas", default="0,73,147")
    a = ap.parse_args()
    import numpy as np
    import torch
    from occm_b200 import _lib
    lib = _lib.load()
    B, T = a.batch, a.frames
    qkv = torch.randn(B * T, 3072, device="cuda")
    qkv[:, :1024] *= 0.125
    ctx = torch.empty(B * T, 1024, device="cuda")
    for _ in range(3):
        assert lib.occ_debug_attention(1, qkv.data_ptr(), B, T, None, ctx.data_ptr(), None) == 0, _lib.last_error()
    n = 160 * 8 * 16
    buf = np.zeros(n, dtype=np.int64)
    assert lib.occ_debug_attention_trace(buf.ctypes.data, n) == 0, _lib.last_error()
    tr = buf.reshape(160, 8, 2, 8)
    names = ["S_iss", "Oh_iss", "O_iss", "s_full", "S_regs", "P_done", "o_full", "epi_end"]
    for cta in [int(c) for c in a.ctas.split(",")]:
        t = tr[cta]
        nz = t[t > 0]
        if nz.size == 0:
            print("cta", cta, "no stamps")
            continue
        t0 = nz.min()
        print(f"== cta {cta}  (cycles from the first stamp; item = pair, two slots)")
        print("item slot " + " ".join(f"{x:>8}" for x in names))
        for it in range(8):
            for sl in range(2):
                if (t[it, sl] > 0).any():
                    print(f"{it:4d} {sl:4d} " + " ".join(f"{(v - t0) if v > 0 else -1:8d}" for v in t[it, sl]))
    # steady-state period: S_iss of item 5 minus item 1, per slot, median over CTAs
    per = []
    for cta in range(148):
        t = tr[cta]
        if t[5, 0, 0] > 0 and t[1, 0, 0] > 0:
            per.append((t[5, 0, 0] - t[1, 0, 0]) / 4.0)
    if per:
        print("median cycles per item (S issue to S issue, slot 0, items 1..5):", float(np.median(per)))


if __name__ == "__main__":
    main()
