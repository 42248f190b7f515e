// lcnn.cu -- the LCNN backend on XLS-R features: /root/reference/models/lcnn.py
//   mfm :123-137 (conv / linear to 2C channels, max of the two halves), group :140-150 (1x1 mfm -> 3x3 mfm; its BatchNorm is never
//   called), LCNN.forward :189-213, ssl_lcnn.forward :242-260.  Eval mode: Dropout = identity, BatchNorm folded to scale / shift.
// fp32 SIMT kernels (210 MFLOP per 4 s utterance against the frontend's 149 GFLOP: memory / latency-bound, no tensor cores):
//   lcnn_layer1   : conv 5x5 (1 -> 8) + mfm (-> 4) + maxpool 2x2 fused: thread = pooled pixel, 6x6 input patch in registers
//   lcnn_pointwise: the group's 1x1 conv + mfm, thread = pixel
//   lcnn_conv3    : conv 3x3 (C -> 2C') + mfm + maxpool 2x2 + BatchNorm fused: thread = (pooled pixel, 4 output channels)
//   lcnn_head     : AdaptiveAvgPool2d((1,64)) + fc0 / fc1 / fc2 (mfm) + fc3, one CTA per utterance
// Layout NCHW fp32, H = frames, W = the 1024 feature channels (the reference feeds [B,1,T,1024] "images").
#include "common.cuh"
#include "kernels.h"

namespace occ {

// ---------------------------------------------------------------------------------------------- layer 1
__global__ void __launch_bounds__(256)
lcnn_layer1_kernel(const float* __restrict__ feat, int B, int T, int H1, const float* __restrict__ w /*[8][25]*/,
                   const float* __restrict__ bias /*[8]*/, float* __restrict__ y /*[B][4][H1][512]*/) {
    __shared__ float sw[8 * 25], sb[8];
    for (int i = threadIdx.x; i < 200; i += blockDim.x) sw[i] = w[i];
    if (threadIdx.x < 8) sb[threadIdx.x] = bias[threadIdx.x];
    __syncthreads();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long total = (long long)B * H1 * 512;
    if (idx >= total) return;
    const int wp = (int)(idx & 511);
    const int hp = (int)((idx >> 9) % H1);
    const int b = (int)(idx / ((long long)H1 * 512));
    const float* x = feat + (long long)b * T * 1024;
    float p[6][6];                                   // rows 2hp-2 .. 2hp+3, columns 2wp-2 .. 2wp+3, zero padded
#pragma unroll
    for (int r = 0; r < 6; ++r) {
        const int hh = 2 * hp - 2 + r;
#pragma unroll
        for (int c = 0; c < 6; ++c) {
            const int ww = 2 * wp - 2 + c;
            p[r][c] = (hh >= 0 && hh < T && ww >= 0 && ww < 1024) ? __ldg(x + (long long)hh * 1024 + ww) : 0.f;
        }
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        float best = -INFINITY;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const int o = c + 4 * half;
#pragma unroll
            for (int dy = 0; dy < 2; ++dy)
#pragma unroll
                for (int dx = 0; dx < 2; ++dx) {
                    float acc = sb[o];
#pragma unroll
                    for (int ky = 0; ky < 5; ++ky)
#pragma unroll
                        for (int kx = 0; kx < 5; ++kx) acc = fmaf(sw[o * 25 + ky * 5 + kx], p[dy + ky][dx + kx], acc);
                    best = fmaxf(best, acc);
                }
        }
        y[(((long long)b * 4 + c) * H1 + hp) * 512 + wp] = best;
    }
}

// ---------------------------------------------------------------------------------------------- group: 1x1 mfm
template <int C>
__global__ void __launch_bounds__(256)
lcnn_pointwise_kernel(const float* __restrict__ x /*[B][C][H][W]*/, long long hw, long long total /*B*H*W*/,
                      const float* __restrict__ w /*[2C][C]*/, const float* __restrict__ bias, float* __restrict__ y) {
    __shared__ float sw[2 * C * C], sb[2 * C];
    for (int i = threadIdx.x; i < 2 * C * C; i += blockDim.x) sw[i] = w[i];
    if (threadIdx.x < 2 * C) sb[threadIdx.x] = bias[threadIdx.x];
    __syncthreads();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const long long b = idx / hw, pix = idx - b * hw;
    const float* xp = x + b * C * hw + pix;
    float v[C];
#pragma unroll
    for (int i = 0; i < C; ++i) v[i] = __ldg(xp + i * hw);
    float* yp = y + b * C * hw + pix;
#pragma unroll
    for (int c = 0; c < C; ++c) {
        float a0 = sb[c], a1 = sb[c + C];
#pragma unroll
        for (int i = 0; i < C; ++i) { a0 = fmaf(sw[c * C + i], v[i], a0); a1 = fmaf(sw[(c + C) * C + i], v[i], a1); }
        yp[c * hw] = fmaxf(a0, a1);
    }
}

// ---------------------------------------------------------------------------------------------- group: 3x3 mfm + pool + BN
// thread = (pooled pixel, group of 4 mfm output channels = 8 conv channels); input a [B][CI][H][W], output [B][CO][H/2][W/2]
template <int CI, int CO>
__global__ void __launch_bounds__(256)
lcnn_conv3_kernel(const float* __restrict__ a, int B, int H, int W, const float* __restrict__ w /*[2CO][CI][3][3]*/,
                  const float* __restrict__ bias /*[2CO]*/, const float* __restrict__ bns, const float* __restrict__ bnt,
                  float* __restrict__ y) {
    constexpr int G = CO / 4;
    __shared__ float sw[2 * CO * CI * 9], sb[2 * CO];
    for (int i = threadIdx.x; i < 2 * CO * CI * 9; i += blockDim.x) sw[i] = w[i];
    if (threadIdx.x < 2 * CO) sb[threadIdx.x] = bias[threadIdx.x];
    __syncthreads();
    const int H2 = H / 2, W2 = W / 2;
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long total = (long long)B * G * H2 * W2;
    if (idx >= total) return;
    const int wp = (int)(idx % W2);
    const int hp = (int)((idx / W2) % H2);
    const int g = (int)((idx / ((long long)W2 * H2)) % G);
    const int b = (int)(idx / ((long long)W2 * H2 * G));
    float acc[8][4];                                  // [conv channel: 4 x (c, c + CO)][pool position]
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int o = (k < 4) ? g * 4 + k : CO + g * 4 + (k - 4);
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[k][q] = sb[o];
    }
    const long long hw = (long long)H * W;
    for (int i = 0; i < CI; ++i) {
        const float* ap = a + ((long long)b * CI + i) * hw;
        float p[4][4];                                // rows 2hp-1 .. 2hp+2, columns 2wp-1 .. 2wp+2, zero padded
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int hh = 2 * hp - 1 + r;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int ww = 2 * wp - 1 + c;
                p[r][c] = (hh >= 0 && hh < H && ww >= 0 && ww < W) ? __ldg(ap + (long long)hh * W + ww) : 0.f;
            }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int o = (k < 4) ? g * 4 + k : CO + g * 4 + (k - 4);
            const float* wk = sw + (o * CI + i) * 9;
#pragma unroll
            for (int dy = 0; dy < 2; ++dy)
#pragma unroll
                for (int dx = 0; dx < 2; ++dx) {
                    float s = acc[k][dy * 2 + dx];
#pragma unroll
                    for (int ky = 0; ky < 3; ++ky)
#pragma unroll
                        for (int kx = 0; kx < 3; ++kx) s = fmaf(wk[ky * 3 + kx], p[dy + ky][dx + kx], s);
                    acc[k][dy * 2 + dx] = s;
                }
        }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        float best = -INFINITY;
#pragma unroll
        for (int q = 0; q < 4; ++q) best = fmaxf(best, fmaxf(acc[k][q], acc[k + 4][q]));      // mfm, then 2x2 max-pool
        const int c = g * 4 + k;
        y[(((long long)b * CO + c) * H2 + hp) * W2 + wp] = fmaf(best, __ldg(bns + c), __ldg(bnt + c));
    }
}

// ---------------------------------------------------------------------------------------------- head
// x3 [B][16][H3][128] -> adaptive average pool (1, 64) -> 1024 -> fc0 (64, mfm 32) -> fc1 (64, mfm 32) -> fc2 (16, mfm 8) -> fc3 (2)
__global__ void __launch_bounds__(256)
lcnn_head_kernel(const float* __restrict__ x3, int H3, const float* __restrict__ f0w, const float* __restrict__ f0b,
                 const float* __restrict__ f1w, const float* __restrict__ f1b, const float* __restrict__ f2w,
                 const float* __restrict__ f2b, const float* __restrict__ f3w, const float* __restrict__ f3b, float* __restrict__ out) {
    __shared__ float pooled[1024], h0[64], m0[32], h1[64], m1[32], h2[16], m2[8];
    const int b = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const float* x = x3 + (long long)b * 16 * H3 * 128;
    const float inv = 1.0f / (float)(2 * H3);
    for (int o = tid; o < 1024; o += 256) {          // o = c * 64 + bin; bin covers columns 2*bin, 2*bin + 1 (128 -> 64)
        const int c = o >> 6, bin = o & 63;
        const float* xc = x + (long long)c * H3 * 128 + 2 * bin;
        float s = 0.f;
        for (int h = 0; h < H3; ++h) s += xc[h * 128] + xc[h * 128 + 1];
        pooled[o] = s * inv;
    }
    __syncthreads();
    for (int o = warp; o < 64; o += 8) {              // fc0: 64 x 1024
        float s = 0.f;
        for (int i = lane; i < 1024; i += 32) s = fmaf(__ldg(f0w + o * 1024 + i), pooled[i], s);
        s = warp_sum(s);
        if (lane == 0) h0[o] = s + f0b[o];
    }
    __syncthreads();
    if (tid < 32) m0[tid] = fmaxf(h0[tid], h0[tid + 32]);
    __syncthreads();
    if (tid < 64) {                                   // fc1: 64 x 32
        float s = f1b[tid];
        for (int i = 0; i < 32; ++i) s = fmaf(f1w[tid * 32 + i], m0[i], s);
        h1[tid] = s;
    }
    __syncthreads();
    if (tid < 32) m1[tid] = fmaxf(h1[tid], h1[tid + 32]);
    __syncthreads();
    if (tid < 16) {                                   // fc2: 16 x 32
        float s = f2b[tid];
        for (int i = 0; i < 32; ++i) s = fmaf(f2w[tid * 32 + i], m1[i], s);
        h2[tid] = s;
    }
    __syncthreads();
    if (tid < 8) m2[tid] = fmaxf(h2[tid], h2[tid + 8]);
    __syncthreads();
    if (tid < 2) {                                    // fc3: 2 x 8
        float s = f3b[tid];
        for (int i = 0; i < 8; ++i) s = fmaf(f3w[tid * 8 + i], m2[i], s);
        out[b * 2 + tid] = s;
    }
}

cudaError_t launch_lcnn(const LcnnW& w, const float* feat, int B, int T, float* y1, float* a2, float* y2, float* a3, float* y3,
                        float* out, cudaStream_t s) {
    const int H1 = T / 2, H2 = H1 / 2, H3 = H2 / 2;
    if (B < 1 || H3 < 1) return cudaErrorInvalidValue;
    {
        const long long n = (long long)B * H1 * 512;
        lcnn_layer1_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(feat, B, T, H1, w.c1w, w.c1b, y1);
    }
    {
        const long long hw = (long long)H1 * 512, n = (long long)B * hw;
        lcnn_pointwise_kernel<4><<<(unsigned)((n + 255) / 256), 256, 0, s>>>(y1, hw, n, w.a2w, w.a2b, a2);
        const long long m = (long long)B * 2 * H2 * 256;
        lcnn_conv3_kernel<4, 8><<<(unsigned)((m + 255) / 256), 256, 0, s>>>(a2, B, H1, 512, w.k2w, w.k2b, w.bn2s, w.bn2t, y2);
    }
    {
        const long long hw = (long long)H2 * 256, n = (long long)B * hw;
        lcnn_pointwise_kernel<8><<<(unsigned)((n + 255) / 256), 256, 0, s>>>(y2, hw, n, w.a3w, w.a3b, a3);
        const long long m = (long long)B * 4 * H3 * 128;
        lcnn_conv3_kernel<8, 16><<<(unsigned)((m + 255) / 256), 256, 0, s>>>(a3, B, H2, 256, w.k3w, w.k3b, w.bn3s, w.bn3t, y3);
    }
    lcnn_head_kernel<<<B, 256, 0, s>>>(y3, H3, w.f0w, w.f0b, w.f1w, w.f1b, w.f2w, w.f2b, w.f3w, w.f3b, out);
    return cudaGetLastError();
}

}  // namespace occ
