// senet.cu -- the SE-ResNet34 backend on XLS-R features: /root/reference/models/senet.py
//   SELayer :13-28 (global average pool -> fc C -> C/16 -> ReLU -> fc -> sigmoid -> channel scale), SEBasicBlock :31-61
//   (conv3x3 + BN + ReLU, conv3x3 + BN, SE, + residual (1x1 stride-s conv + BN when the shape changes), ReLU),
//   ResNet.forward :121-143 (conv 7x7 s2 + BN + ReLU, maxpool 3x3 s2, layers [3,4,6,3] at 16/32/64/128 channels, global
//   average pool, embedding 128 -> 128 and classifier 128 -> 2), ssl_resnet34.forward :168-185.  Eval mode: BatchNorm folded.
// fp32 SIMT direct convolutions, NCHW, H = frames, W = the 1024 feature channels.  2 GFLOP per 4 s utterance against the
// frontend's 149: a first, parity-exact form of this "next" row -- not yet a tensor-core path.
#include "common.cuh"
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
                            float* buf4, float* vec, float* com, float* des, cudaStream_t s, int* launches) {
    const int H1 = (T + 6 - 7) / 2 + 1, W1 = 512;
    const int H2 = (H1 + 2 - 3) / 2 + 1, W2 = 256;
    if (B < 1 || T < 1) return cudaErrorInvalidValue;
    int n = 0;
    cudaError_t e = rs_conv<7, 2>(feat, B, 1, T, 1024, w.conv1, 16, w.bn1s, w.bn1t, 1, buf0, s); ++n;
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
        e = st == 1 ? rs_conv<3, 1>(x, B, C, H, W, k.c1, Co, k.b1s, k.b1t, 1, t1, s) : rs_conv<3, 2>(x, B, C, H, W, k.c1, Co, k.b1s, k.b1t, 1, t1, s);
        if (e != cudaSuccess) return e;
        e = rs_conv<3, 1>(t1, B, Co, Ho, Wo, k.c2, Co, k.b2s, k.b2t, 0, t2, s);
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
