// senet.cu -- the SE-ResNet34 backend on XLS-R features: /root/reference/models/senet.py
//   SELayer :13-28 (global average pool -> fc C -> C/16 -> ReLU -> fc -> sigmoid -> channel scale), SEBasicBlock :31-61
//   (conv3x3 + BN + ReLU, conv3x3 + BN, SE, + residual (1x1 stride-s conv + BN when the shape changes), ReLU),
//   ResNet.forward :121-143 (conv 7x7 s2 + BN + ReLU, maxpool 3x3 s2, layers [3,4,6,3] at 16/32/64/128 channels, global
//   average pool, embedding 128 -> 128 and classifier 128 -> 2), ssl_resnet34.forward :168-185.  Eval mode: BatchNorm folded.
// NCHW fp32 activations, H = frames, W = the 1024 feature channels; 2 GFLOP per 4 s utterance against the frontend's 149.  The 3x3
// convolutions run on mma.sync with split operands (rs_conv3_mma_kernel, fp32-class); the 7x7 stem, the 1x1 downsample convolutions
// and the other layers are fp32 SIMT kernels.
#include "common.cuh"
#include <cstdlib>
#include "kernels.h"

namespace occ {

// ---------------------------------------------------------------------------------------------- direct convolution
// y[b, co, ho, wo] = act(scale[co] * sum_{ci, ky, kx} w[co][ci][ky][kx] * x[b, ci, ho*S - P + ky, wo*S - P + kx] + shift[co])
// block = 256 threads x PX horizontally adjacent output pixels of one (b, group of 8 output channels).  The group's weights sit
// in shared memory as [ci][tap][8 channels], so one tap of all 8 channels is two 16 B broadcast loads feeding 8 * PX FMAs
// (the first form -- one scalar shared-memory load per FMA -- ran at 13 TFLOP/s: profiles/r2_n4_backends.json).
template <int KS, int S, int PX>
__global__ void __launch_bounds__(256)
rs_conv_kernel(const float* __restrict__ x, int Ci, int H, int W, const float* __restrict__ w, int Co, int Ho, int Wo,
               const float* __restrict__ scale, const float* __restrict__ shift, int relu, float* __restrict__ y) {
    extern __shared__ __align__(16) float sw[];       // [Ci][KS*KS][8]
    constexpr int P = KS / 2, KK = KS * KS, PW = KS + (PX - 1) * S;      // PW: patch width covering PX outputs
    const int cg = blockIdx.y, b = blockIdx.z;
    const int nw = 8 * Ci * KK;
    for (int i = threadIdx.x; i < nw; i += 256) {     // global [k][ci][t] -> shared [ci][t][k]
        const int k = i / (Ci * KK), r = i - k * (Ci * KK);
        sw[r * 8 + k] = w[(long long)cg * nw + i];
    }
    __syncthreads();
    const int Wq = (Wo + PX - 1) / PX;                // pixel groups per output row
    const int q = blockIdx.x * 256 + threadIdx.x;
    if (q >= Ho * Wq) return;
    const int ho = q / Wq, wo0 = (q - ho * Wq) * PX;
    float acc[PX][8];
#pragma unroll
    for (int px = 0; px < PX; ++px)
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[px][k] = 0.f;
    const float* xb = x + (long long)b * Ci * H * W;
    for (int ci = 0; ci < Ci; ++ci) {
        const float* xc = xb + (long long)ci * H * W;
        float p[KS][PW];
#pragma unroll
        for (int ky = 0; ky < KS; ++ky) {
            const int hh = ho * S - P + ky;
#pragma unroll
            for (int kx = 0; kx < PW; ++kx) {
                const int ww = wo0 * S - P + kx;
                p[ky][kx] = (hh >= 0 && hh < H && ww >= 0 && ww < W) ? __ldg(xc + (long long)hh * W + ww) : 0.f;
            }
        }
        const float4* wc = reinterpret_cast<const float4*>(sw + ci * KK * 8);
#pragma unroll
        for (int ky = 0; ky < KS; ++ky)
#pragma unroll
            for (int kx = 0; kx < KS; ++kx) {
                const float4 w0 = wc[(ky * KS + kx) * 2], w1 = wc[(ky * KS + kx) * 2 + 1];
#pragma unroll
                for (int px = 0; px < PX; ++px) {
                    const float v = p[ky][kx + px * S];
                    acc[px][0] = fmaf(w0.x, v, acc[px][0]); acc[px][1] = fmaf(w0.y, v, acc[px][1]);
                    acc[px][2] = fmaf(w0.z, v, acc[px][2]); acc[px][3] = fmaf(w0.w, v, acc[px][3]);
                    acc[px][4] = fmaf(w1.x, v, acc[px][4]); acc[px][5] = fmaf(w1.y, v, acc[px][5]);
                    acc[px][6] = fmaf(w1.z, v, acc[px][6]); acc[px][7] = fmaf(w1.w, v, acc[px][7]);
                }
            }
    }
#pragma unroll
    for (int px = 0; px < PX; ++px) {
        if (wo0 + px >= Wo) break;
        float* yb = y + ((long long)b * Co + cg * 8) * Ho * Wo + (long long)ho * Wo + wo0 + px;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            float v = fmaf(acc[px][k], __ldg(scale + cg * 8 + k), __ldg(shift + cg * 8 + k));
            if (relu) v = fmaxf(v, 0.f);
            yb[(long long)k * Ho * Wo] = v;
        }
    }
}

template <int KS, int S, int PX = (KS == 3 ? 4 : 1)>
static cudaError_t rs_conv(const float* x, int B, int Ci, int H, int W, const float* w, int Co, const float* scale, const float* shift,
                           int relu, float* y, cudaStream_t s) {
    const int P = KS / 2;
    const int Ho = (H + 2 * P - KS) / S + 1, Wo = (W + 2 * P - KS) / S + 1;
    if (Co % 8 != 0 || Ho < 1 || Wo < 1) return cudaErrorInvalidValue;
    const size_t smem = (size_t)8 * Ci * KS * KS * sizeof(float);
    cudaError_t e = ensure_dyn_smem(reinterpret_cast<const void*>(rs_conv_kernel<KS, S, PX>), smem);
    if (e != cudaSuccess) return e;
    const int groups = Ho * ((Wo + PX - 1) / PX);
    dim3 grid((unsigned)((groups + 255) / 256), (unsigned)(Co / 8), (unsigned)B);
    rs_conv_kernel<KS, S, PX><<<grid, 256, smem, s>>>(x, Ci, H, W, w, Co, Ho, Wo, scale, shift, relu, y);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------- 3x3 conv on mma.sync
// The 3x3 convolutions (29 at stride 1, C -> C channels, C = 16 / 32 / 64 / 128; 3 at stride 2, C -> 2C) are 90 % of the backend's time
// as SIMT kernels (27 TFLOP/s: issue 61 %, FMA pipe 43 %).  Here they run on the warp-level tensor path, per INPUT CHANNEL:
//   out[p, co] += sum_tap x[ci][p + tap] * w[co][ci][tap]     as  M = 16 pixels of a row, N = 8 output channels, K = 16 slots,
// the 9 taps and their low-order parts sharing the 16 slots of TWO m16n8k16 MMAs (fp32-class result, as conv0 in frontend.cu):
//   MMA 1: slots 0..7 = x_hi[tap s] * w_hi[tap s],  slots 8..15 = x_lo[tap s-8] * w_hi[tap s-8]
//   MMA 2: slots 0..7 = x_hi[tap s] * w_lo[tap s],  slot 8 = x_hi[8] w_hi[8],  9 = x_lo[8] w_hi[8],  10 = x_hi[8] w_lo[8],  11..15 = 0
// so a lane's A registers are uniform in kind -- (hi, hi) of taps 2t, 2t+1 for both MMAs, (lo, lo) of the same taps, and tap 8 for
// t < 2 -- and come straight from the NCHW planes (no transposition: K runs over taps, not channels).  Weights are pre-packed in
// B-fragment lane order ([ci][n-tile][mma][lane] uint2, rs_pack_frag_kernel) and staged through shared memory in chunks of input
// channels (cp.async, double-buffered when they do not fit at once); a warp keeps its 16 pixels x C accumulators in registers across
// the chunks (two warps share a pixel tile at 128 output channels).  Stride 2 only changes the sample offsets.  Epilogue: folded
// BatchNorm (+ ReLU), NCHW.
template <int CI, int CO> struct RsMma {
    static constexpr int NT = CO / 8;                                 // n-tiles of the layer
    static constexpr int NSPLIT = CO == 128 ? 2 : 1;                  // warps sharing a pixel tile (each takes NT / NSPLIT n-tiles)
    static constexpr int NTW = NT / NSPLIT;                           // n-tiles per warp
    static constexpr int WARPS = 8;
    static constexpr int ALL = CI * CO <= 1024;                       // all weights fit 64 KB: staged once per (persistent) CTA
    static constexpr int CIC = ALL ? CI : 512 / CO;                   // input channels per shared-memory chunk: all, or 32 KB worth
    static constexpr int NBUF = ALL ? 1 : 2;                          // multi-chunk sizes: double-buffered with cp.async
    static constexpr int CHUNK_U2 = CIC * NT * 2 * 32;                // uint2 per chunk
    static constexpr int SMEM = NBUF * CHUNK_U2 * 8;
    static constexpr int PD = CO == 16 ? 8 : 4;                      // input channels whose samples are in flight (the planes come from L2)
};

__global__ void rs_pack_frag_kernel(const float* __restrict__ w /*[Co][Ci][3][3]*/, int Ci, int Co, uint2* __restrict__ frag) {
    const int NT = Co / 8;
    const long long total = (long long)Ci * NT * 64;
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const int lane = (int)(idx & 31), v = (int)((idx >> 5) & 1), j = (int)((idx >> 6) % NT), ci = (int)(idx / (64 * NT));
    const int g = lane >> 2, t = lane & 3, co = 8 * j + g;
    const float* wp = w + ((long long)co * Ci + ci) * 9;
    auto slot = [&](int sl) -> uint32_t {         // bf16 bits of B(K slot sl, column g) of MMA v
        bf16 hi, lo;
        if (v == 0) { split_bf16(wp[sl & 7], hi, lo); return __bfloat16_as_ushort(hi); }
        if (sl < 8) { split_bf16(wp[sl], hi, lo); return __bfloat16_as_ushort(lo); }
        if (sl > 10) return 0u;
        split_bf16(wp[8], hi, lo);
        return __bfloat16_as_ushort(sl == 10 ? lo : hi);
    };
    frag[idx] = make_uint2(slot(2 * t) | (slot(2 * t + 1) << 16), slot(2 * t + 8) | (slot(2 * t + 9) << 16));
}

__device__ __forceinline__ void rs_mma_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// (lo16, hi16) = (bf16(a), bf16(b)) in one instruction, and the residuals' pair: x ~= hi + lo to ~2^-17
__device__ __forceinline__ void rs_split_pair(float a, float b, uint32_t& hi, uint32_t& lo) {
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(hi) : "f"(b), "f"(a));
    const float ra = a - __uint_as_float(hi << 16), rb = b - __uint_as_float(hi & 0xffff0000u);
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(lo) : "f"(rb), "f"(ra));
}
__device__ __forceinline__ void rs_cp_async16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}

template <int CI, int CO, int S>
__global__ void __launch_bounds__(RsMma<CI, CO>::WARPS * 32, 2)
rs_conv3_mma_kernel(const float* __restrict__ x, int Hi, int Wi, int H, int W, long long total_tiles, const uint2* __restrict__ frag_g,
                    const float* __restrict__ scale, const float* __restrict__ shift, int relu, float* __restrict__ y) {
    using K = RsMma<CI, CO>;
    extern __shared__ __align__(16) uint2 sfrag[];            // NBUF chunks: [ci][n-tile][mma][lane]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const int tpr = W >> 4;                                   // 16-pixel tiles per OUTPUT row (H x W outputs from Hi x Wi inputs, stride S)
    const int plane = Hi * Wi, oplane = H * W;
    constexpr int TPC = K::WARPS / K::NSPLIT;                 // pixel tiles per CTA pass
    const int nhalf = warp % K::NSPLIT;                       // which share of the n-tiles this warp computes
    // taps 2t, 2t+1 (and tap 8 for t < 2) of this lane: offsets from the (stride-scaled) pixel inside an input plane
    const int ky0 = (2 * t) / 3, kx0 = (2 * t) % 3, ky1 = (2 * t + 1) / 3, kx1 = (2 * t + 1) % 3;
    const int off0 = (ky0 - 1) * Wi + (kx0 - 1), off1 = (ky1 - 1) * Wi + (kx1 - 1), off8 = Wi + 1;
    auto stage = [&](int chunk, int buf) {                    // weights of input channels [chunk * CIC, +CIC) -> shared memory
        const uint4* src = reinterpret_cast<const uint4*>(frag_g + (long long)chunk * K::CHUNK_U2);
        uint4* dst = reinterpret_cast<uint4*>(sfrag + (long long)buf * K::CHUNK_U2);
        for (int i = threadIdx.x; i < K::CHUNK_U2 / 2; i += K::WARPS * 32) rs_cp_async16(dst + i, src + i);
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    bool loaded = false;
    for (long long tile0 = (long long)blockIdx.x * TPC; tile0 < total_tiles; tile0 += (long long)gridDim.x * TPC) {
        const long long tile = tile0 + warp / K::NSPLIT;
        const bool active = tile < total_tiles;
        int b = 0, h = 0, w0 = 0;
        if (active) {
            b = (int)(tile / ((long long)H * tpr));
            const int rem = (int)(tile - (long long)b * H * tpr);
            h = rem / tpr; w0 = (rem - h * tpr) << 4;
        }
        const int pa = w0 + g, pb = pa + 8;
        const int hi = S * h, ia = S * pa, ib = S * pb;       // input row / columns of the window centres
        // validity of the lane's six samples (zero padding of the conv)
        const bool r0 = hi + ky0 - 1 >= 0 && hi + ky0 - 1 < Hi, r1 = hi + ky1 - 1 >= 0 && hi + ky1 - 1 < Hi, r8 = hi + 1 < Hi;
        const bool va0 = active && r0 && ia + kx0 - 1 >= 0 && ia + kx0 - 1 < Wi, vb0 = active && r0 && ib + kx0 - 1 >= 0 && ib + kx0 - 1 < Wi;
        const bool va1 = active && r1 && ia + kx1 - 1 >= 0 && ia + kx1 - 1 < Wi, vb1 = active && r1 && ib + kx1 - 1 >= 0 && ib + kx1 - 1 < Wi;
        const bool va8 = active && t < 2 && r8 && ia + 1 < Wi, vb8 = active && t < 2 && r8 && ib + 1 < Wi;
        const float* xb = x + (long long)b * CI * plane;     // warp-uniform base of the utterance; per-lane 32-bit offsets below
        const int o0 = hi * Wi + ia + off0, o1 = hi * Wi + ia + off1, o8 = hi * Wi + ia + off8;
        float acc[K::NTW][4];
#pragma unroll
        for (int j = 0; j < K::NTW; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
        constexpr int PD = K::PD;
        float v[PD][6];
        auto fetch = [&](int ci, float (&d)[6]) {
            const float* xc = xb + ci * plane;
            d[0] = va0 ? __ldg(xc + o0) : 0.f; d[1] = va1 ? __ldg(xc + o1) : 0.f;
            d[2] = vb0 ? __ldg(xc + o0 + 8 * S) : 0.f; d[3] = vb1 ? __ldg(xc + o1 + 8 * S) : 0.f;
            d[4] = va8 ? __ldg(xc + o8) : 0.f; d[5] = vb8 ? __ldg(xc + o8 + 8 * S) : 0.f;
        };
#pragma unroll
        for (int i = 0; i < PD; ++i) fetch(i, v[i]);
        constexpr int NCHUNK = CI / K::CIC;
        if (K::NBUF == 2) stage(0, 0);
        for (int ch = 0; ch < NCHUNK; ++ch) {
            if (K::NBUF == 1) {
                if (!loaded) {                      // single-chunk sizes keep their weights across the tiles of the CTA
                    stage(0, 0);
                    asm volatile("cp.async.wait_group 0;" ::: "memory");
                    __syncthreads();
                    loaded = true;
                }
            } else {
                if (ch + 1 < NCHUNK) { stage(ch + 1, (ch + 1) & 1); asm volatile("cp.async.wait_group 1;" ::: "memory"); }
                else asm volatile("cp.async.wait_group 0;" ::: "memory");
                __syncthreads();                    // chunk ch has landed for every thread
            }
            const uint2* sbuf = sfrag + (long long)(K::NBUF == 2 ? (ch & 1) : 0) * K::CHUNK_U2 + (nhalf * K::NTW) * 64 + lane;
            const int c0 = ch * K::CIC;
#pragma unroll 1
            for (int cb = 0; cb < K::CIC; cb += PD) {
#pragma unroll
                for (int u = 0; u < PD; ++u) {
                    const int cc = cb + u;
                    uint32_t ahi, alo, bhi, blo, h8, l8;
                    rs_split_pair(v[u][0], v[u][1], ahi, alo);        // pixel pa: taps 2t, 2t+1
                    rs_split_pair(v[u][2], v[u][3], bhi, blo);        // pixel pb
                    rs_split_pair(v[u][4], v[u][5], h8, l8);          // tap 8: (pa, pb)
                    // MMA 2, slots 2t+8, 2t+9: (x_hi[8], x_lo[8]) for t = 0, (x_hi[8], 0) for t = 1, zeros above (the samples are 0 there)
                    const uint32_t a8 = t == 0 ? ((h8 & 0xffffu) | (l8 << 16)) : (h8 & 0xffffu);
                    const uint32_t b8 = t == 0 ? ((h8 >> 16) | (l8 & 0xffff0000u)) : (h8 >> 16);
                    if (c0 + cc + PD < CI) fetch(c0 + cc + PD, v[u]);
                    const uint2* fr = sbuf + cc * K::NT * 64;
#pragma unroll
                    for (int j = 0; j < K::NTW; ++j) {
                        const uint2 w1 = fr[(j * 2 + 0) * 32], w2 = fr[(j * 2 + 1) * 32];
                        rs_mma_16816(acc[j], ahi, bhi, alo, blo, w1.x, w1.y);
                        rs_mma_16816(acc[j], ahi, bhi, a8, b8, w2.x, w2.y);
                    }
                }
            }
            if (K::NBUF == 2) __syncthreads();     // every warp is done with this buffer before the next prefetch overwrites it
        }
        if (active) {
            float* yb = y + (long long)b * CO * oplane + h * W;
#pragma unroll
            for (int j = 0; j < K::NTW; ++j) {
                const int co = 8 * (nhalf * K::NTW + j) + 2 * t;
                const float s0 = __ldg(scale + co), s1 = __ldg(scale + co + 1), t0 = __ldg(shift + co), t1 = __ldg(shift + co + 1);
                float o0v = fmaf(acc[j][0], s0, t0), o1v = fmaf(acc[j][1], s1, t1), o2v = fmaf(acc[j][2], s0, t0), o3v = fmaf(acc[j][3], s1, t1);
                if (relu) { o0v = fmaxf(o0v, 0.f); o1v = fmaxf(o1v, 0.f); o2v = fmaxf(o2v, 0.f); o3v = fmaxf(o3v, 0.f); }
                yb[co * oplane + pa] = o0v; yb[(co + 1) * oplane + pa] = o1v;
                yb[co * oplane + pb] = o2v; yb[(co + 1) * oplane + pb] = o3v;
            }
        }
    }
}

template <int CI, int CO, int S>
static cudaError_t rs_conv3_mma_t(const float* x, int B, int Hi, int Wi, const uint2* frag, const float* scale, const float* shift, int relu,
                                  float* y, int num_sms, cudaStream_t s) {
    using K = RsMma<CI, CO>;
    cudaError_t e = ensure_dyn_smem(reinterpret_cast<const void*>(rs_conv3_mma_kernel<CI, CO, S>), K::SMEM);
    if (e != cudaSuccess) return e;
    const int H = (Hi + 2 - 3) / S + 1, W = (Wi + 2 - 3) / S + 1;
    if (W % 16 != 0) return cudaErrorInvalidValue;
    const long long total = (long long)B * H * (W >> 4);
    constexpr int TPC = K::WARPS / K::NSPLIT;
    long long blocks = (total + TPC - 1) / TPC;
    if (K::ALL && blocks > 2LL * num_sms) blocks = 2LL * num_sms;       // persistent CTAs, weights staged once
    rs_conv3_mma_kernel<CI, CO, S><<<(unsigned)blocks, K::WARPS * 32, K::SMEM, s>>>(x, Hi, Wi, H, W, total, frag, scale, shift, relu, y);
    return cudaGetLastError();
}
// x [B, Ci, Hi, Wi] -> y [B, Co, Ho, Wo], 3x3 / padding 1, stride 1 (Ci == Co in {16, 32, 64, 128}) or 2 (Co == 2 Ci); Wo % 16 == 0
static bool rs_conv3_mma_ok(int Ci, int Co, int stride, int Wi) {
    const int Wo = (Wi + 2 - 3) / stride + 1;
    if (Wo % 16 != 0) return false;
    if (stride == 1) return Ci == Co && (Co == 16 || Co == 32 || Co == 64 || Co == 128);
    return stride == 2 && Co == 2 * Ci && (Co == 32 || Co == 64 || Co == 128);
}
static cudaError_t rs_conv3_mma(const float* x, int B, int Ci, int Co, int stride, int Hi, int Wi, const void* frag, const float* scale,
                                const float* shift, int relu, float* y, int num_sms, cudaStream_t s) {
    const uint2* f = reinterpret_cast<const uint2*>(frag);
    if (stride == 1) {
        switch (Co) {
            case 16: return rs_conv3_mma_t<16, 16, 1>(x, B, Hi, Wi, f, scale, shift, relu, y, num_sms, s);
            case 32: return rs_conv3_mma_t<32, 32, 1>(x, B, Hi, Wi, f, scale, shift, relu, y, num_sms, s);
            case 64: return rs_conv3_mma_t<64, 64, 1>(x, B, Hi, Wi, f, scale, shift, relu, y, num_sms, s);
            case 128: return rs_conv3_mma_t<128, 128, 1>(x, B, Hi, Wi, f, scale, shift, relu, y, num_sms, s);
        }
    } else {
        switch (Co) {
            case 32: return rs_conv3_mma_t<16, 32, 2>(x, B, Hi, Wi, f, scale, shift, relu, y, num_sms, s);
            case 64: return rs_conv3_mma_t<32, 64, 2>(x, B, Hi, Wi, f, scale, shift, relu, y, num_sms, s);
            case 128: return rs_conv3_mma_t<64, 128, 2>(x, B, Hi, Wi, f, scale, shift, relu, y, num_sms, s);
        }
    }
    return cudaErrorInvalidValue;
}
cudaError_t launch_rs_pack_frag(const float* w, int Ci, int Co, void* frag, cudaStream_t s) {
    const long long total = (long long)Ci * (Co / 8) * 64;
    rs_pack_frag_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(w, Ci, Co, reinterpret_cast<uint2*>(frag));
    return cudaGetLastError();
}
size_t rs_frag_bytes(int Ci, int Co) { return (size_t)Ci * (Co / 8) * 64 * sizeof(uint2); }

// ---------------------------------------------------------------------------------------------- the 7x7 / stride-2 stem on mma.sync
// conv1 (senet.py:71: 1 -> 16 channels, 7x7, stride 2, padding 3) with its 49 taps in the K dimension, the same way: M = 16 output
// pixels of a row, N = 8 channels, ten m16n8k16 MMAs per n-tile --
//   MMA q (q = 0..5):   slots 0..7 = x_hi w_hi of taps 8q..8q+7,  slots 8..15 = x_lo w_hi of the same taps
//   MMA 6 + k (k < 3):  slots 0..15 = x_hi w_lo of taps 16k..16k+15
//   MMA 9:              slots 0, 1, 2 = tap 48 as x_hi w_hi, x_lo w_hi, x_hi w_lo
// so lane t needs taps 8q + 2t, 8q + 2t + 1 (q < 6) and tap 48 (t < 2) of its two pixels: 26 samples, 13 packed splits.
static constexpr int RS7_FRAG_U2 = 10 * 2 * 32;      // [mma][n-tile][lane] uint2
__global__ void rs_pack_stem_kernel(const float* __restrict__ w /*[16][1][7][7]*/, uint2* __restrict__ frag) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= RS7_FRAG_U2) return;
    const int lane = idx & 31, j = (idx >> 5) & 1, v = idx >> 6;
    const int g = lane >> 2, t = lane & 3, co = 8 * j + g;
    const float* wp = w + co * 49;
    auto slot = [&](int sl) -> uint32_t {
        bf16 hi, lo;
        if (v < 6) { split_bf16(wp[8 * v + (sl & 7)], hi, lo); return __bfloat16_as_ushort(hi); }
        if (v < 9) { split_bf16(wp[16 * (v - 6) + sl], hi, lo); return __bfloat16_as_ushort(lo); }
        if (sl > 2) return 0u;
        split_bf16(wp[48], hi, lo);
        return __bfloat16_as_ushort(sl == 2 ? lo : hi);
    };
    frag[idx] = make_uint2(slot(2 * t) | (slot(2 * t + 1) << 16), slot(2 * t + 8) | (slot(2 * t + 9) << 16));
}

__global__ void __launch_bounds__(256, 2)
rs_stem_mma_kernel(const float* __restrict__ x /*[B][Hi][Wi]*/, int Hi, int Wi, int H, int W, long long total_tiles,
                   const uint2* __restrict__ frag_g, const float* __restrict__ scale, const float* __restrict__ shift, float* __restrict__ y) {
    __shared__ uint2 sfrag[RS7_FRAG_U2];
    for (int i = threadIdx.x; i < RS7_FRAG_U2; i += blockDim.x) sfrag[i] = frag_g[i];
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const int tpr = W >> 4, oplane = H * W;
    // this lane's 13 taps: (ky, kx) and the offset inside the input plane
    int tky[13], tkx[13], toff[13];
#pragma unroll
    for (int i = 0; i < 13; ++i) {
        const int tap = i < 12 ? 8 * (i >> 1) + 2 * t + (i & 1) : 48;
        tky[i] = tap / 7; tkx[i] = tap - 7 * tky[i];
        toff[i] = (tky[i] - 3) * Wi + (tkx[i] - 3);
    }
    for (long long tile = (long long)blockIdx.x * 8 + warp; tile < total_tiles; tile += (long long)gridDim.x * 8) {
        const int b = (int)(tile / ((long long)H * tpr));
        const int rem = (int)(tile - (long long)b * H * tpr);
        const int h = rem / tpr, w0 = (rem - h * tpr) << 4;
        const int pa = w0 + g, pb = pa + 8;
        const int hi = 2 * h, ia = 2 * pa, ib = 2 * pb;
        const float* xp = x + (long long)b * Hi * Wi + (long long)hi * Wi + ia;      // window centre of pixel pa (pb: + 16)
        const bool inner = hi - 3 >= 0 && hi + 3 < Hi && 2 * w0 - 3 >= 0 && 2 * (w0 + 15) + 3 < Wi;      // warp-uniform: no padding in reach
        uint32_t ah[7], al[7], bh[7], bl[7];      // [q < 6]: taps 8q + 2t, +1 of pixels pa / pb as packed (hi, hi) and (lo, lo); [6]: tap 48 of (pa, pb)
        float sa[12], sb[12];                     // all 24 samples are requested before the first one is used (no branch between the loads)
        if (inner) {
#pragma unroll
            for (int i = 0; i < 12; ++i) { sa[i] = __ldg(xp + toff[i]); sb[i] = __ldg(xp + 16 + toff[i]); }
        } else {
#pragma unroll
            for (int i = 0; i < 12; ++i) {
                const bool r = hi + tky[i] - 3 >= 0 && hi + tky[i] - 3 < Hi;
                const int ca = ia + tkx[i] - 3, cb = ib + tkx[i] - 3;
                sa[i] = (r && ca >= 0 && ca < Wi) ? __ldg(xp + toff[i]) : 0.f;
                sb[i] = (r && cb >= 0 && cb < Wi) ? __ldg(xp + 16 + toff[i]) : 0.f;
            }
        }
#pragma unroll
        for (int q = 0; q < 6; ++q) {
            rs_split_pair(sa[2 * q], sa[2 * q + 1], ah[q], al[q]);
            rs_split_pair(sb[2 * q], sb[2 * q + 1], bh[q], bl[q]);
        }
        {
            float a48 = 0.f, b48 = 0.f;           // tap 48 = (ky 6, kx 6); lanes t < 2 only
            if (t < 2) {
                const bool r = hi + 3 < Hi;
                a48 = (r && ia + 3 < Wi) ? __ldg(xp + toff[12]) : 0.f;
                b48 = (r && ib + 3 < Wi) ? __ldg(xp + 16 + toff[12]) : 0.f;
            }
            uint32_t h8, l8;
            rs_split_pair(a48, b48, h8, l8);
            // MMA 9, slots 2t, 2t+1: (x_hi, x_lo) for t = 0, (x_hi, 0) for t = 1
            ah[6] = t == 0 ? ((h8 & 0xffffu) | (l8 << 16)) : (h8 & 0xffffu);
            bh[6] = t == 0 ? ((h8 >> 16) | (l8 & 0xffff0000u)) : (h8 >> 16);
        }
        float acc[2][4];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
            const uint2* fr = sfrag + j * 32 + lane;
#pragma unroll
            for (int q = 0; q < 6; ++q) { const uint2 wv = fr[q * 64]; rs_mma_16816(acc[j], ah[q], bh[q], al[q], bl[q], wv.x, wv.y); }
#pragma unroll
            for (int k = 0; k < 3; ++k) { const uint2 wv = fr[(6 + k) * 64]; rs_mma_16816(acc[j], ah[2 * k], bh[2 * k], ah[2 * k + 1], bh[2 * k + 1], wv.x, wv.y); }
            { const uint2 wv = fr[9 * 64]; rs_mma_16816(acc[j], ah[6], bh[6], 0u, 0u, wv.x, wv.y); }
        }
        float* yb = y + (long long)b * 16 * oplane + h * W;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int co = 8 * j + 2 * t;
            const float s0 = __ldg(scale + co), s1 = __ldg(scale + co + 1), t0 = __ldg(shift + co), t1 = __ldg(shift + co + 1);
            yb[co * oplane + pa] = fmaxf(fmaf(acc[j][0], s0, t0), 0.f); yb[(co + 1) * oplane + pa] = fmaxf(fmaf(acc[j][1], s1, t1), 0.f);
            yb[co * oplane + pb] = fmaxf(fmaf(acc[j][2], s0, t0), 0.f); yb[(co + 1) * oplane + pb] = fmaxf(fmaf(acc[j][3], s1, t1), 0.f);
        }
    }
}
cudaError_t launch_rs_pack_stem(const float* w, void* frag, cudaStream_t s) {
    rs_pack_stem_kernel<<<(RS7_FRAG_U2 + 255) / 256, 256, 0, s>>>(w, reinterpret_cast<uint2*>(frag));
    return cudaGetLastError();
}
size_t rs_stem_frag_bytes() { return (size_t)RS7_FRAG_U2 * sizeof(uint2); }
// feat [B, T, 1024] -> y [B, 16, H, 512] = relu(bn1(conv1(x)))
static cudaError_t rs_stem_mma(const float* x, int B, int Hi, int Wi, const void* frag, const float* scale, const float* shift, float* y,
                               int num_sms, cudaStream_t s) {
    const int H = (Hi + 6 - 7) / 2 + 1, W = (Wi + 6 - 7) / 2 + 1;
    if (W % 16 != 0) return cudaErrorInvalidValue;
    const long long total = (long long)B * H * (W >> 4);
    long long blocks = (total + 7) / 8;
    if (blocks > 2LL * num_sms) blocks = 2LL * num_sms;
    rs_stem_mma_kernel<<<(unsigned)blocks, 256, 0, s>>>(x, Hi, Wi, H, W, total, reinterpret_cast<const uint2*>(frag), scale, shift, y);
    return cudaGetLastError();
}

// MaxPool2d(kernel 3, stride 2, padding 1)   (senet.py:74)
__global__ void rs_maxpool_kernel(const float* __restrict__ x, long long planes, int H, int W, int Ho, int Wo, float* __restrict__ y) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= planes * Ho * Wo) return;
    const int wo = (int)(idx % Wo), ho = (int)((idx / Wo) % Ho);
    const long long pl = idx / ((long long)Wo * Ho);
    const float* xp = x + pl * H * W;
    float m = -INFINITY;
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
        const int hh = 2 * ho - 1 + ky;
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
            const int ww = 2 * wo - 1 + kx;
            if (hh >= 0 && hh < H && ww >= 0 && ww < W) m = fmaxf(m, xp[(long long)hh * W + ww]);
        }
    }
    y[idx] = m;
}

// ---------------------------------------------------------------------------------------------- squeeze / excite / merge
// mean over H x W of every (b, c) plane: one warp per plane, fixed summation order (lane-strided partials, shuffle tree)
__global__ void __launch_bounds__(256)
rs_avgpool_kernel(const float* __restrict__ x, long long planes, int hw, float* __restrict__ y) {
    const long long pl = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (pl >= planes) return;
    const float* xp = x + pl * hw;
    float s = 0.f;
    for (int i = lane; i < hw; i += 32) s += xp[i];
    s = warp_sum(s);
    if (lane == 0) y[pl] = s / (float)hw;
}

// SELayer.fc (senet.py:17-22): s = sigmoid(W2 relu(W1 y)); one CTA per utterance, C <= 128, C / 16 <= 8 hidden units
__global__ void __launch_bounds__(128)
rs_se_kernel(const float* __restrict__ y /*[B][C]*/, int C, const float* __restrict__ w1 /*[C/16][C]*/, const float* __restrict__ w2 /*[C][C/16]*/,
             float* __restrict__ sc /*[B][C]*/) {
    __shared__ float sy[128], sh[8];
    const int b = blockIdx.x, tid = threadIdx.x, R = C / 16;
    if (tid < C) sy[tid] = y[(long long)b * C + tid];
    __syncthreads();
    if (tid < R) {
        float s = 0.f;
        for (int i = 0; i < C; ++i) s = fmaf(w1[tid * C + i], sy[i], s);
        sh[tid] = fmaxf(s, 0.f);
    }
    __syncthreads();
    if (tid < C) {
        float s = 0.f;
        for (int i = 0; i < R; ++i) s = fmaf(w2[tid * R + i], sh[i], s);
        sc[(long long)b * C + tid] = 1.0f / (1.0f + expf(-s));
    }
}

// out = relu(t * s[b, c] + residual)   (senet.py:53-59)
__global__ void rs_scale_add_relu_kernel(const float* __restrict__ t, const float* __restrict__ sc, const float* __restrict__ res, long long total,
                                         int hw, float* __restrict__ out) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    out[idx] = fmaxf(fmaf(t[idx], __ldg(sc + idx / hw), res[idx]), 0.f);
}

// embedding (128 -> 128) and classifier (128 -> 2) on the pooled vector   (senet.py:136-139)
__global__ void __launch_bounds__(160)
rs_head_kernel(const float* __restrict__ v /*[B][128]*/, const float* __restrict__ we, const float* __restrict__ be, const float* __restrict__ wc,
               const float* __restrict__ bc, float* __restrict__ com, float* __restrict__ des) {
    __shared__ float sv[128];
    const int b = blockIdx.x, tid = threadIdx.x;
    if (tid < 128) sv[tid] = v[(long long)b * 128 + tid];
    __syncthreads();
    if (tid < 128) {
        float s = be[tid];
        for (int i = 0; i < 128; ++i) s = fmaf(we[tid * 128 + i], sv[i], s);
        com[(long long)b * 128 + tid] = s;
    } else if (tid < 130) {
        const int o = tid - 128;
        float s = bc[o];
        for (int i = 0; i < 128; ++i) s = fmaf(wc[o * 128 + i], sv[i], s);
        des[(long long)b * 2 + o] = s;
    }
}

// ---------------------------------------------------------------------------------------------- the network
// buf0: conv1 output [B,16,H1,512]; buf1..buf4: block activations (<= [B,16,H2,256] each); vec: [B,128] x 2
cudaError_t launch_resnet34(const ResnetW& w, const float* feat, int B, int T, float* buf0, float* buf1, float* buf2, float* buf3,
                            float* buf4, float* vec, float* com, float* des, cudaStream_t s, int* launches, int num_sms) {
    static const bool mma_env = [] { const char* e = getenv("OCC_RS_MMA"); return !(e && e[0] == '0'); }();     // 0: SIMT convs everywhere, for A/B runs
    const int H1 = (T + 6 - 7) / 2 + 1, W1 = 512;
    const int H2 = (H1 + 2 - 3) / 2 + 1, W2 = 256;
    if (B < 1 || T < 1) return cudaErrorInvalidValue;
    int n = 0;
    cudaError_t e = (w.conv1f && mma_env) ? rs_stem_mma(feat, B, T, 1024, w.conv1f, w.bn1s, w.bn1t, buf0, num_sms, s)
                                          : rs_conv<7, 2>(feat, B, 1, T, 1024, w.conv1, 16, w.bn1s, w.bn1t, 1, buf0, s);
    ++n;
    if (e != cudaSuccess) return e;
    {
        const long long total = (long long)B * 16 * H2 * W2;
        rs_maxpool_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(buf0, (long long)B * 16, H1, W1, H2, W2, buf1); ++n;
    }
    float *x = buf1, *t1 = buf2, *t2 = buf3, *ds = buf4;
    int C = 16, H = H2, W = W2;
    for (int bi = 0; bi < w.nblocks; ++bi) {
        const ResnetBlockW& k = w.blk[bi];
        const int st = k.stride, Co = k.planes;
        const int Ho = (H + 2 - 3) / st + 1, Wo = (W + 2 - 3) / st + 1;
        if (k.c1f && mma_env && rs_conv3_mma_ok(C, Co, st, W)) e = rs_conv3_mma(x, B, C, Co, st, H, W, k.c1f, k.b1s, k.b1t, 1, t1, num_sms, s);
        else e = st == 1 ? rs_conv<3, 1>(x, B, C, H, W, k.c1, Co, k.b1s, k.b1t, 1, t1, s) : rs_conv<3, 2>(x, B, C, H, W, k.c1, Co, k.b1s, k.b1t, 1, t1, s);
        if (e != cudaSuccess) return e;
        if (k.c2f && mma_env && rs_conv3_mma_ok(Co, Co, 1, Wo)) e = rs_conv3_mma(t1, B, Co, Co, 1, Ho, Wo, k.c2f, k.b2s, k.b2t, 0, t2, num_sms, s);
        else e = rs_conv<3, 1>(t1, B, Co, Ho, Wo, k.c2, Co, k.b2s, k.b2t, 0, t2, s);
        if (e != cudaSuccess) return e;
        const long long planes = (long long)B * Co;
        rs_avgpool_kernel<<<(unsigned)((planes + 7) / 8), 256, 0, s>>>(t2, planes, Ho * Wo, vec);
        rs_se_kernel<<<B, 128, 0, s>>>(vec, Co, k.se1, k.se2, vec + (long long)B * 128);
        const float* res = x;
        n += 4;
        if (k.dw != nullptr) {
            e = st == 1 ? rs_conv<1, 1>(x, B, C, H, W, k.dw, Co, k.dbs, k.dbt, 0, ds, s) : rs_conv<1, 2>(x, B, C, H, W, k.dw, Co, k.dbs, k.dbt, 0, ds, s);
            if (e != cudaSuccess) return e;
            res = ds; ++n;
        }
        const long long total = planes * Ho * Wo;
        rs_scale_add_relu_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(t2, vec + (long long)B * 128, res, total, Ho * Wo, t1); ++n;
        float* nx = t1; t1 = x; x = nx;             // block output becomes the next input; the old input buffer is scratch now
        C = Co; H = Ho; W = Wo;
    }
    {
        const long long planes = (long long)B * C;
        rs_avgpool_kernel<<<(unsigned)((planes + 7) / 8), 256, 0, s>>>(x, planes, H * W, vec);
        rs_head_kernel<<<B, 160, 0, s>>>(vec, w.emb_w, w.emb_b, w.cls_w, w.cls_b, com, des); n += 2;
    }
    if (launches) *launches = n;
    return cudaGetLastError();
}

}  // namespace occ
