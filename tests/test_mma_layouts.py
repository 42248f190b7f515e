"""CPU restatement of the operand layouts of the two hand-written mma.sync convolutions (no GPU needed).

conv0 (occm_b200/csrc/frontend.cu: conv0_pack_frag_kernel / conv0_mma2_kernel) and the SE-ResNet34 3x3 convolution
(occm_b200/csrc/senet.cu: rs_pack_frag_kernel / rs_conv3_mma_kernel) put the conv TAPS into the K dimension of
mma.sync.m16n8k16 and use the spare K slots for the low-order bf16 parts of the operands.  This file restates, in numpy,
  * the lane <-> element maps of the m16n8k16 fragments (PTX ISA: A row-major 16x16, B col 16x8, C 16x8),
  * the slot plans of the two pack kernels and the per-lane operand selection of the two compute kernels,
and checks the product against a direct fp64 convolution: the plans must reproduce x*w to split-bf16 accuracy (~2^-17).
It pins the DESIGN of the layouts; the kernels themselves are checked against the oracle on the GPU (tests/test_gpu_*.py)."""
import numpy as np
import torch


def split(x):
    x = torch.as_tensor(np.asarray(x, dtype=np.float32))
    hi = x.to(torch.bfloat16).float()
    lo = (x - hi).to(torch.bfloat16).float()
    return hi.numpy(), lo.numpy()


def mma_16816(a_frag, b_frag, c_frag):
    """a_frag [32 lanes][4 regs][2], b_frag [32][2][2], c_frag [32][4] (accumulated in place)."""
    A = np.zeros((16, 16)); B = np.zeros((16, 8))
    for lane in range(32):
        g, t = lane >> 2, lane & 3
        for i in range(4):
            for e in range(2):
                A[g + 8 * (i & 1), 2 * t + 8 * (i >> 1) + e] = a_frag[lane, i, e]
        for i in range(2):
            for e in range(2):
                B[2 * t + 8 * i + e, g] = b_frag[lane, i, e]
    D = A @ B
    for lane in range(32):
        g, t = lane >> 2, lane & 3
        for i in range(4):
            c_frag[lane, i] += D[g + 8 * (i >> 1), 2 * t + (i & 1)]


def test_conv0_slot_plan_and_channel_map():
    """conv0: M = 16 frames, N = 8 channels per tile, K slots = [x_hi w_hi taps 0..9 | 1*b_hi | 1*b_lo | x_hi w_lo taps 6..9] and
    [x_lo w_hi taps 0..9 | x_hi w_lo taps 0..5]; weights centred over the channels; column j of tile n = 4q + i is channel
    32q + 8 (j / 2) + 2i + (j % 2)."""
    g0 = torch.Generator().manual_seed(0)
    C = 64                                        # the plan does not depend on the channel count: 8 tiles instead of 64
    w = (torch.randn(C, 10, generator=g0) * 0.3).numpy(); bias = (torch.randn(C, generator=g0) * 0.1 + 0.05).numpy()
    N = 5 * 21 + 10 + 2; L0 = (N - 10) // 5 + 1
    x = (torch.randn(N, generator=g0) * 0.1).numpy()
    mean = np.concatenate([w.astype(np.float64).mean(0), [bias.astype(np.float64).mean()]]).astype(np.float32)
    wc = np.concatenate([w - mean[None, :10], (bias - mean[10])[:, None]], 1).astype(np.float32)
    whi, wlo = split(wc)

    def chan(n, j):
        return 32 * (n >> 2) + 8 * (j >> 1) + 2 * (n & 3) + (j & 1)

    def slot_w(v, ch, s):
        if s < 10:
            return whi[ch, s]
        if v == 0:
            return (whi[ch, 10] if s == 10 else wlo[ch, 10]) if s < 12 else wlo[ch, s - 6]
        return wlo[ch, s - 10]

    NT = C // 8
    frag = np.zeros((NT, 2, 32, 2, 2), np.float32)
    for n in range(NT):
        for v in range(2):
            for lane in range(32):
                g, t = lane >> 2, lane & 3
                ch = chan(n, g)
                frag[n, v, lane, 0] = [slot_w(v, ch, 2 * t), slot_w(v, ch, 2 * t + 1)]
                frag[n, v, lane, 1] = [slot_w(v, ch, 2 * t + 8), slot_w(v, ch, 2 * t + 9)]
    out = np.zeros((L0, C))
    for f0 in range(0, L0, 16):
        A0 = np.zeros((32, 4, 2)); A1 = np.zeros((32, 4, 2))
        for lane in range(32):
            g, t = lane >> 2, lane & 3
            oB = 6 if t == 2 else 8
            oC = 8 if t == 0 else 2 * t - 2
            r = np.zeros((2, 4, 2))
            for h in range(2):
                base = (f0 + g + 8 * h) * 5
                ld = lambda j: x[base + j] if base + j < N else 0.0
                ha, la = split([ld(2 * t), ld(2 * t + 1)]); hb, _ = split([ld(oB), ld(oB + 1)]); hc, lc = split([ld(oC), ld(oC + 1)])
                r[h, 0] = ha; r[h, 1] = [1, 1] if t == 1 else hb; r[h, 2] = la; r[h, 3] = lc if t == 0 else hc
            A0[lane] = [r[0, 0], r[1, 0], r[0, 1], r[1, 1]]; A1[lane] = [r[0, 2], r[1, 2], r[0, 3], r[1, 3]]
        for n in range(NT):
            D = np.zeros((32, 4))
            mma_16816(A0, frag[n, 0], D); mma_16816(A1, frag[n, 1], D)
            for lane in range(32):
                g, t = lane >> 2, lane & 3
                ch = 32 * (n >> 2) + 8 * t + 2 * (n & 3)              # accumulator columns 2t, 2t+1 of tile n
                assert ch == chan(n, 2 * t) and ch + 1 == chan(n, 2 * t + 1)
                if f0 + g < L0:
                    out[f0 + g, ch], out[f0 + g, ch + 1] = D[lane, 0], D[lane, 1]
                if f0 + g + 8 < L0:
                    out[f0 + g + 8, ch], out[f0 + g + 8, ch + 1] = D[lane, 2], D[lane, 3]
    ref = np.zeros((L0, C))
    for f in range(L0):
        v = w.astype(np.float64) @ x[5 * f:5 * f + 10].astype(np.float64) + bias
        ref[f] = v - v.mean()                                          # the centred weights give LayerNorm's numerator directly
    assert np.abs(out - ref).max() < 4e-5 * max(1.0, np.abs(ref).max())


def test_resnet_conv3_slot_plan():
    """SE-ResNet34 3x3 / stride-1 conv: per input channel, M = 16 pixels of a row, N = 8 output channels, K slots =
    [x_hi w_hi taps 0..7 | x_lo w_hi taps 0..7] and [x_hi w_lo taps 0..7 | tap 8: hi*hi, lo*hi, hi*lo | 0...]."""
    g0 = torch.Generator().manual_seed(1)
    C, H, W = 8, 3, 32
    w = (torch.randn(C, C, 3, 3, generator=g0) * 0.2).numpy(); x = torch.randn(C, H, W, generator=g0).numpy()
    whi, wlo = split(w.reshape(C, C, 9))

    def slot_w(v, co, ci, sl):
        if v == 0:
            return whi[co, ci, sl & 7]
        if sl < 8:
            return wlo[co, ci, sl]
        if sl > 10:
            return 0.0
        return wlo[co, ci, 8] if sl == 10 else whi[co, ci, 8]

    NT = C // 8
    frag = np.zeros((C, NT, 2, 32, 2, 2), np.float32)
    for ci in range(C):
        for j in range(NT):
            for v in range(2):
                for lane in range(32):
                    g, t = lane >> 2, lane & 3
                    co = 8 * j + g
                    frag[ci, j, v, lane, 0] = [slot_w(v, co, ci, 2 * t), slot_w(v, co, ci, 2 * t + 1)]
                    frag[ci, j, v, lane, 1] = [slot_w(v, co, ci, 2 * t + 8), slot_w(v, co, ci, 2 * t + 9)]
    y = np.zeros((C, H, W)); xf = x.reshape(C, H * W)
    for h in range(H):
        for w0 in range(0, W, 16):
            acc = np.zeros((NT, 32, 4))
            for ci in range(C):
                A1 = np.zeros((32, 4, 2)); A2 = np.zeros((32, 4, 2))
                for lane in range(32):
                    g, t = lane >> 2, lane & 3
                    ky0, kx0, ky1, kx1 = (2 * t) // 3, (2 * t) % 3, (2 * t + 1) // 3, (2 * t + 1) % 3
                    pa = w0 + g; pb = pa + 8

                    def sample(p, ky, kx, on=True):
                        r, c = h + ky - 1, p + kx - 1
                        return xf[ci, r * W + c] if on and 0 <= r < H and 0 <= c < W else 0.0
                    v = [sample(pa, ky0, kx0), sample(pa, ky1, kx1), sample(pb, ky0, kx0), sample(pb, ky1, kx1),
                         sample(pa, 2, 2, t < 2), sample(pb, 2, 2, t < 2)]
                    hi, lo = split(v)
                    a8 = [hi[4], lo[4]] if t == 0 else [hi[4], 0]
                    b8 = [hi[5], lo[5]] if t == 0 else [hi[5], 0]
                    A1[lane] = [[hi[0], hi[1]], [hi[2], hi[3]], [lo[0], lo[1]], [lo[2], lo[3]]]
                    A2[lane] = [[hi[0], hi[1]], [hi[2], hi[3]], a8, b8]
                for j in range(NT):
                    mma_16816(A1, frag[ci, j, 0], acc[j]); mma_16816(A2, frag[ci, j, 1], acc[j])
            for lane in range(32):
                g, t = lane >> 2, lane & 3
                pa = w0 + g; pb = pa + 8
                for j in range(NT):
                    co = 8 * j + 2 * t
                    y[co, h, pa], y[co + 1, h, pa], y[co, h, pb], y[co + 1, h, pb] = acc[j, lane]
    ref = torch.nn.functional.conv2d(torch.tensor(x)[None].double(), torch.tensor(w).double(), padding=1)[0].numpy()
    assert np.abs(y - ref).max() < 4e-5 * max(1.0, np.abs(ref).max())
