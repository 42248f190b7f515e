// backend.cu -- AASIST graph-attention backend (fp32, CUDA cores; < 1 % of the FLOPs of the path but
// top-k node selection makes it precision sensitive, so nothing here is rounded to bf16).
// Follows /root/reference/models/sslassist.py:509-597 (AModel.forward after the frontend), :58-151
// (GraphAttentionLayer), :154-329 (HtrgGraphAttentionLayer), :332-368 (GraphPool), :373-429 (Residual_block).
// Every kernel takes per-utterance valid sizes so that one padded batch reproduces the reference's
// batch-1, un-padded behaviour (variable length).
#include "common.cuh"
#include "kernels.h"

namespace occ {

// =================================================================================================
// LL output [B,T,128] -> max_pool2d(3,3) over (feature, time) -> first_bn affine -> SELU -> x0 [B,1,42,Wm]
__global__ void ll_pool_kernel(const float* __restrict__ ll, int T, int Wm, const int* __restrict__ wlen,
                               float scale, float shift, float* __restrict__ x0) {
    const int b = blockIdx.y;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;       // f3 * Wm + w
    if (idx >= 42 * Wm) return;
    const int f3 = idx / Wm, w = idx - f3 * Wm;
    const int wl = wlen ? wlen[b] : Wm;
    float r = 0.f;
    if (w < wl) {
        float m = -INFINITY;
#pragma unroll
        for (int dt = 0; dt < 3; ++dt) {
            const float* p = ll + ((long long)b * T + 3 * w + dt) * 128 + 3 * f3;
            m = fmaxf(m, fmaxf(p[0], fmaxf(p[1], p[2])));
        }
        r = selu_f(m * scale + shift);
    }
    x0[(long long)b * 42 * Wm + idx] = r;
}
cudaError_t launch_ll_pool(const float* ll, int B, int T, int Wm, const int* wlen, float scale, float shift, float* x0, cudaStream_t s) {
    dim3 grid((42 * Wm + 127) / 128, B);
    ll_pool_kernel<<<grid, 128, 0, s>>>(ll, T, Wm, wlen, scale, shift, x0);
    return cudaGetLastError();
}

// =================================================================================================
// Direct conv2d, stride 1, kernel (KH, 3), zero padding (ph, 1); NCHW fp32; one thread = one output
// position x 16 output channels.  Epilogue: +bias; [BN affine + SELU]; [+resid]; [BN affine + SELU].
// wp: weights re-laid as [Cout/16][Cin][KH*3][16].
static constexpr int CV_CO = 16;
template <int KH>
__global__ void __launch_bounds__(128)
conv2d_kernel(const float* __restrict__ in, int Cin, int Hin, int Wm, const int* __restrict__ wlen,
              const float* __restrict__ wp, const float* __restrict__ bias, int Cout, int Hout, int ph,
              const float* __restrict__ s1, const float* __restrict__ t1, const float* __restrict__ resid,
              const float* __restrict__ s2, const float* __restrict__ t2, float* __restrict__ out) {
    extern __shared__ float wsm[];                      // [Cin][KH*3][16]
    const int cg = blockIdx.y, b = blockIdx.z;
    const int nw = Cin * KH * 3 * CV_CO;
    const float* wsrc = wp + (long long)cg * nw;
    for (int i = threadIdx.x; i < nw; i += blockDim.x) wsm[i] = wsrc[i];
    __syncthreads();
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= Hout * Wm) return;
    const int h = p / Wm, w = p - h * Wm;
    const int wl = wlen ? wlen[b] : Wm;
    float acc[CV_CO];
#pragma unroll
    for (int i = 0; i < CV_CO; ++i) acc[i] = 0.f;
    if (w < wl) {
        const float* inb = in + (long long)b * Cin * Hin * Wm;
        for (int ci = 0; ci < Cin; ++ci) {
            const float* inc = inb + (long long)ci * Hin * Wm;
#pragma unroll
            for (int kh = 0; kh < KH; ++kh) {
                const int hi = h + kh - ph;
                if (hi < 0 || hi >= Hin) continue;
#pragma unroll
                for (int kw = 0; kw < 3; ++kw) {
                    const int wi = w + kw - 1;
                    if (wi < 0 || wi >= wl) continue;
                    const float v = __ldg(inc + hi * Wm + wi);
                    const float4* wv = reinterpret_cast<const float4*>(wsm + ((ci * KH + kh) * 3 + kw) * CV_CO);
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const float4 ww = wv[q];
                        acc[4 * q] = fmaf(v, ww.x, acc[4 * q]);
                        acc[4 * q + 1] = fmaf(v, ww.y, acc[4 * q + 1]);
                        acc[4 * q + 2] = fmaf(v, ww.z, acc[4 * q + 2]);
                        acc[4 * q + 3] = fmaf(v, ww.w, acc[4 * q + 3]);
                    }
                }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < CV_CO; ++i) {
        const int co = cg * CV_CO + i;
        const long long o = (((long long)b * Cout + co) * Hout + h) * Wm + w;
        float v = 0.f;
        if (w < wl) {
            v = acc[i] + bias[co];
            if (s1) v = selu_f(v * s1[co] + t1[co]);
            if (resid) v += resid[o];
            if (s2) v = selu_f(v * s2[co] + t2[co]);
        }
        out[o] = v;
    }
}
cudaError_t launch_conv2d(int KH, const float* in, int B, int Cin, int Hin, int Wm, const int* wlen, const float* wp,
                          const float* bias, int Cout, int Hout, int ph, const float* s1, const float* t1, const float* resid,
                          const float* s2, const float* t2, float* out, cudaStream_t s) {
    dim3 grid((Hout * Wm + 127) / 128, Cout / CV_CO, B);
    const size_t smem = (size_t)Cin * KH * 3 * CV_CO * sizeof(float);
    if (KH == 2) conv2d_kernel<2><<<grid, 128, smem, s>>>(in, Cin, Hin, Wm, wlen, wp, bias, Cout, Hout, ph, s1, t1, resid, s2, t2, out);
    else if (KH == 1) conv2d_kernel<1><<<grid, 128, smem, s>>>(in, Cin, Hin, Wm, wlen, wp, bias, Cout, Hout, ph, s1, t1, resid, s2, t2, out);
    else return cudaErrorInvalidValue;
    return cudaGetLastError();
}

// =================================================================================================
// attention map (sslassist.py:465-471, 523): w = Conv1x1(128->64)( BN( SELU( Conv1x1(64->128)(x) ) ) )
// one thread per position; W1 [128][64] and W2^T [128][64] in shared memory (broadcast reads).
__global__ void __launch_bounds__(128)
att_map_kernel(const float* __restrict__ x, int HW, const float* __restrict__ W1, const float* __restrict__ b1,
               const float* __restrict__ bs, const float* __restrict__ bt, const float* __restrict__ W2T,
               const float* __restrict__ b2, float* __restrict__ wmap) {
    extern __shared__ float sm[];
    float* sW1 = sm;               // [128][64]
    float* sW2 = sm + 128 * 64;    // [128][64]  (W2T[j][c] = W2[c][j])
    for (int i = threadIdx.x; i < 128 * 64; i += blockDim.x) { sW1[i] = W1[i]; sW2[i] = W2T[i]; }
    __syncthreads();
    const int b = blockIdx.y;
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= HW) return;
    const float* xb = x + (long long)b * 64 * HW + p;
    float xv[64], wv[64];
#pragma unroll
    for (int c = 0; c < 64; ++c) { xv[c] = xb[(long long)c * HW]; wv[c] = __ldg(b2 + c); }
    for (int j = 0; j < 128; ++j) {
        float a = __ldg(b1 + j);
        const float4* r1 = reinterpret_cast<const float4*>(sW1 + j * 64);
#pragma unroll
        for (int q = 0; q < 16; ++q) {
            const float4 ww = r1[q];
            a = fmaf(ww.x, xv[4 * q], a); a = fmaf(ww.y, xv[4 * q + 1], a);
            a = fmaf(ww.z, xv[4 * q + 2], a); a = fmaf(ww.w, xv[4 * q + 3], a);
        }
        a = selu_f(a) * __ldg(bs + j) + __ldg(bt + j);
        const float4* r2 = reinterpret_cast<const float4*>(sW2 + j * 64);
#pragma unroll
        for (int q = 0; q < 16; ++q) {
            const float4 ww = r2[q];
            wv[4 * q] = fmaf(ww.x, a, wv[4 * q]); wv[4 * q + 1] = fmaf(ww.y, a, wv[4 * q + 1]);
            wv[4 * q + 2] = fmaf(ww.z, a, wv[4 * q + 2]); wv[4 * q + 3] = fmaf(ww.w, a, wv[4 * q + 3]);
        }
    }
    float* ob = wmap + (long long)b * 64 * HW + p;
#pragma unroll
    for (int c = 0; c < 64; ++c) ob[(long long)c * HW] = wv[c];
}
cudaError_t launch_att_map(const float* x, int B, int HW, const float* W1, const float* b1, const float* bs, const float* bt,
                           const float* W2T, const float* b2, float* wmap, cudaStream_t s) {
    static bool attr = false;
    const size_t smem = 2 * 128 * 64 * sizeof(float);
    if (!attr) {
        cudaError_t e = cudaFuncSetAttribute(att_map_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        attr = true;
    }
    dim3 grid((HW + 127) / 128, B);
    att_map_kernel<<<grid, 128, smem, s>>>(x, HW, W1, b1, bs, bt, W2T, b2, wmap);
    return cudaGetLastError();
}

// =================================================================================================
// spectro-temporal pooling (sslassist.py:526-538).  x, wmap: [B,64,42,Wm].
//   e_S[b][h][c] = sum_w x * softmax_w(wmap) + pos_S[h][c]       (one warp per (b,c,h))
//   e_T[b][w][c] = sum_h x * softmax_h(wmap)                     (one thread per (b,c,w))
__global__ void pool_S_kernel(const float* __restrict__ x, const float* __restrict__ wmap, int Wm, const int* __restrict__ wlen,
                              const float* __restrict__ posS, float* __restrict__ eS, int total_rows) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);   // (b*64 + c)*42 + h
    const int lane = threadIdx.x & 31;
    if (row >= total_rows) return;
    const int h = row % 42, c = (row / 42) % 64, b = row / (42 * 64);
    const int wl = wlen ? wlen[b] : Wm;
    const float* xr = x + (long long)row * Wm;
    const float* wr = wmap + (long long)row * Wm;
    float m = -INFINITY;
    for (int w = lane; w < wl; w += 32) m = fmaxf(m, wr[w]);
    m = warp_max(m);
    float den = 0.f, num = 0.f;
    for (int w = lane; w < wl; w += 32) {
        const float e = expf(wr[w] - m);
        den += e;
        num = fmaf(e, xr[w], num);
    }
    den = warp_sum(den);
    num = warp_sum(num);
    if (lane == 0) eS[((long long)b * 42 + h) * 64 + c] = num / den + posS[h * 64 + c];
}
__global__ void pool_T_kernel(const float* __restrict__ x, const float* __restrict__ wmap, int Wm, const int* __restrict__ wlen,
                              float* __restrict__ eT, int B) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;   // (b*64 + c)*Wm + w
    if (idx >= B * 64 * Wm) return;
    const int w = idx % Wm, c = (idx / Wm) % 64, b = idx / (Wm * 64);
    const int wl = wlen ? wlen[b] : Wm;
    float r = 0.f;
    if (w < wl) {
        const long long base = ((long long)(b * 64 + c) * 42) * Wm + w;
        float m = -INFINITY;
        for (int h = 0; h < 42; ++h) m = fmaxf(m, wmap[base + (long long)h * Wm]);
        float den = 0.f, num = 0.f;
        for (int h = 0; h < 42; ++h) {
            const float e = expf(wmap[base + (long long)h * Wm] - m);
            den += e;
            num = fmaf(e, x[base + (long long)h * Wm], num);
        }
        r = num / den;
    }
    eT[((long long)b * Wm + w) * 64 + c] = r;
}
cudaError_t launch_st_pool(const float* x, const float* wmap, int B, int Wm, const int* wlen, const float* posS, float* eS,
                           float* eT, cudaStream_t s) {
    const int rows = B * 64 * 42;
    pool_S_kernel<<<(rows + 7) / 8, 256, 0, s>>>(x, wmap, Wm, wlen, posS, eS, rows);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    const int n = B * 64 * Wm;
    pool_T_kernel<<<(n + 127) / 128, 128, 0, s>>>(x, wmap, Wm, wlen, eT, B);
    return cudaGetLastError();
}

// =================================================================================================
// Graph attention (homogeneous and heterogeneous) -- one CTA computes GA_ROWS complete output rows.
//   att[i][j] = sum_o aw_{type(i,j)}[o] * tanh( bp[o] + sum_d Wp[o][d] x_i[d] x_j[d] ) / temp
//   a = softmax_j(att[i][:]);  y_i = SELU(BN( Wa (sum_j a_j x_j) + ba + Wo x_i + bo ))
// Heterogeneous layers (n1 nodes of type 1 then n2 of type 2) pick aw11 / aw22 / aw12 per block
// (sslassist.py:284-291; aw12 serves both off-diagonal blocks).  blockIdx.x == gridDim.x-1 with
// `master` set computes the master-node update instead (sslassist.py:234-239, 255-270, 310-316).
static constexpr int GA_ROWS = 8;

template <int DIN, int DOUT>
__global__ void __launch_bounds__(256)
gat_kernel(const GatArgs a) {
    extern __shared__ float sm[];
    const int z = blockIdx.z, b = blockIdx.y;
    const GatW& W = a.w[z];
    const int nmax = a.nmax[z];
    const int n1 = a.n1[z] ? a.n1[z][b] : a.n1c[z];
    const int n2 = a.n2[z] ? a.n2[z][b] : a.n2c[z];
    const int n = n1 + n2;
    float* sx = sm;                              // [nmax][DIN]
    float* sWp = sx + nmax * DIN;                // [DOUT][DIN]
    float* satt = sWp + DOUT * DIN;              // [GA_ROWS][nmax]
    float* sagg = satt + GA_ROWS * nmax;         // [GA_ROWS][DIN]
    const float* xb = a.x[z] + (long long)b * nmax * DIN;
    const int tid = threadIdx.x;
    const bool master_block = a.has_master && (blockIdx.x == gridDim.x - 1);
    const int i0 = blockIdx.x * GA_ROWS;
    if (!master_block && i0 >= n) return;
    for (int i = tid; i < n * DIN; i += blockDim.x) sx[i] = xb[i];
    const float* Wsel = master_block ? W.WpM : W.Wp;
    for (int i = tid; i < DOUT * DIN; i += blockDim.x) sWp[i] = Wsel[i];
    __syncthreads();

    if (master_block) {
        // ---- master node: att_j = awM . tanh(WpM (x_j * m) + bpM) / temp ; softmax over j
        const float* mi = W.master_in + (long long)b * W.master_stride;
        for (int j = tid; j < n; j += blockDim.x) {
            float zv[DIN];
#pragma unroll
            for (int d = 0; d < DIN; ++d) zv[d] = sx[j * DIN + d] * __ldg(mi + d);
            float acc = 0.f;
            for (int o = 0; o < DOUT; ++o) {
                float t = __ldg(W.bpM + o);
                const float4* wr = reinterpret_cast<const float4*>(sWp + o * DIN);
#pragma unroll
                for (int q = 0; q < DIN / 4; ++q) {
                    const float4 ww = wr[q];
                    t = fmaf(ww.x, zv[4 * q], t); t = fmaf(ww.y, zv[4 * q + 1], t);
                    t = fmaf(ww.z, zv[4 * q + 2], t); t = fmaf(ww.w, zv[4 * q + 3], t);
                }
                acc = fmaf(__ldg(W.awM + o), tanhf(t), acc);
            }
            satt[j] = acc / W.temp;
        }
        __syncthreads();
        if (tid < 32) {
            float m = -INFINITY;
            for (int j = tid; j < n; j += 32) m = fmaxf(m, satt[j]);
            m = warp_max(m);
            float den = 0.f;
            for (int j = tid; j < n; j += 32) { const float e = expf(satt[j] - m); satt[j] = e; den += e; }
            den = warp_sum(den);
            const float inv = 1.0f / den;
            for (int j = tid; j < n; j += 32) satt[j] *= inv;
        }
        __syncthreads();
        if (tid < DIN) {
            float acc = 0.f;
            for (int j = 0; j < n; ++j) acc = fmaf(satt[j], sx[j * DIN + tid], acc);
            sagg[tid] = acc;
        }
        __syncthreads();
        if (tid < DOUT) {
            float v = __ldg(W.baM + tid) + __ldg(W.boM + tid);
            for (int d = 0; d < DIN; ++d) {
                v = fmaf(__ldg(W.WaM + tid * DIN + d), sagg[d], v);
                v = fmaf(__ldg(W.WoM + tid * DIN + d), __ldg(mi + d), v);
            }
            W.master_out[(long long)b * DOUT + tid] = v;
        }
        return;
    }

    const int rows = min(GA_ROWS, n - i0);
    // ---- pairwise attention logits
    for (int pid = tid; pid < rows * n; pid += blockDim.x) {
        const int il = pid / n, j = pid - il * n;
        const int i = i0 + il;
        float zv[DIN];
#pragma unroll
        for (int d = 0; d < DIN; ++d) zv[d] = sx[i * DIN + d] * sx[j * DIN + d];
        const float* aw = (i < n1) ? ((j < n1) ? W.aw11 : W.aw12) : ((j < n1) ? W.aw12 : W.aw22);
        float acc = 0.f;
        for (int o = 0; o < DOUT; ++o) {
            float t = __ldg(W.bp + o);
            const float4* wr = reinterpret_cast<const float4*>(sWp + o * DIN);
#pragma unroll
            for (int q = 0; q < DIN / 4; ++q) {
                const float4 ww = wr[q];
                t = fmaf(ww.x, zv[4 * q], t); t = fmaf(ww.y, zv[4 * q + 1], t);
                t = fmaf(ww.z, zv[4 * q + 2], t); t = fmaf(ww.w, zv[4 * q + 3], t);
            }
            acc = fmaf(__ldg(aw + o), tanhf(t), acc);
        }
        satt[il * nmax + j] = acc / W.temp;
    }
    __syncthreads();
    // ---- softmax over j (one warp per row)
    {
        const int warp = tid >> 5, lane = tid & 31;
        if (warp < rows) {
            float* r = satt + warp * nmax;
            float m = -INFINITY;
            for (int j = lane; j < n; j += 32) m = fmaxf(m, r[j]);
            m = warp_max(m);
            float den = 0.f;
            for (int j = lane; j < n; j += 32) { const float e = expf(r[j] - m); r[j] = e; den += e; }
            den = warp_sum(den);
            const float inv = 1.0f / den;
            for (int j = lane; j < n; j += 32) r[j] *= inv;
        }
    }
    __syncthreads();
    // ---- aggregate
    for (int k = tid; k < rows * DIN; k += blockDim.x) {
        const int il = k / DIN, d = k - il * DIN;
        const float* r = satt + il * nmax;
        float acc = 0.f;
        for (int j = 0; j < n; ++j) acc = fmaf(r[j], sx[j * DIN + d], acc);
        sagg[k] = acc;
    }
    __syncthreads();
    // ---- project + BN + SELU
    for (int k = tid; k < rows * DOUT; k += blockDim.x) {
        const int il = k / DOUT, o = k - il * DOUT;
        const int i = i0 + il;
        float v = __ldg(W.ba + o) + __ldg(W.bo + o);
        const float* wa = W.Wa + o * DIN;
        const float* wo = W.Wo + o * DIN;
        for (int d = 0; d < DIN; ++d) {
            v = fmaf(__ldg(wa + d), sagg[il * DIN + d], v);
            v = fmaf(__ldg(wo + d), sx[i * DIN + d], v);
        }
        v = selu_f(v * __ldg(W.bns + o) + __ldg(W.bnt + o));
        a.y[z][((long long)b * nmax + i) * DOUT + o] = v;
    }
}

template <int DIN, int DOUT>
static cudaError_t launch_gat_t(const GatArgs& a, int B, int nz, cudaStream_t s) {
    int nmax = a.nmax[0];
    for (int z = 1; z < nz; ++z) nmax = a.nmax[z] > nmax ? a.nmax[z] : nmax;
    const size_t smem = ((size_t)nmax * DIN + DOUT * DIN + GA_ROWS * nmax + GA_ROWS * DIN) * sizeof(float);
    static size_t cur = 0;
    if (smem > cur) {
        cudaError_t e = cudaFuncSetAttribute(gat_kernel<DIN, DOUT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        cur = smem;
    }
    dim3 grid((nmax + GA_ROWS - 1) / GA_ROWS + (a.has_master ? 1 : 0), B, nz);
    gat_kernel<DIN, DOUT><<<grid, 256, smem, s>>>(a);
    return cudaGetLastError();
}
cudaError_t launch_gat(const GatArgs& a, int B, int nz, int din, int dout, cudaStream_t s) {
    if (din == 64 && dout == 64) return launch_gat_t<64, 64>(a, B, nz, s);
    if (din == 64 && dout == 32) return launch_gat_t<64, 32>(a, B, nz, s);
    if (din == 32 && dout == 32) return launch_gat_t<32, 32>(a, B, nz, s);
    return cudaErrorInvalidValue;
}

// =================================================================================================
// GraphPool (sslassist.py:341-368): s = sigmoid(w.h + b); keep k = max(int(n*0.5),1) best nodes in
// descending-score order (torch.topk sorted=True; ties: lower index first); out rows = h * s.
// blockIdx.z selects one of up to 4 pools sharing a launch.
template <int D>
__global__ void __launch_bounds__(128)
graph_pool_kernel(const PoolArgs a) {
    extern __shared__ float sm[];   // scores [nmax]
    const int z = blockIdx.z, b = blockIdx.y;
    const int n = a.n[z] ? a.n[z][b] : a.nc[z];
    const int off = a.noff[z] ? a.noff[z][b] : a.noffc[z];
    const float* hb = a.h[z] + (long long)b * a.h_bstride[z] + (long long)off * D;
    const int k = max(n / 2, 1);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = warp; i < n; i += 4) {
        float acc = 0.f;
        for (int d = lane; d < D; d += 32) acc = fmaf(hb[i * D + d], __ldg(a.w[z] + d), acc);
        acc = warp_sum(acc);
        if (lane == 0) sm[i] = 1.0f / (1.0f + expf(-(acc + __ldg(a.bias[z]))));
    }
    __syncthreads();
    for (int i = warp; i < n; i += 4) {
        const float si = sm[i];
        int cnt = 0;
        for (int j = lane; j < n; j += 32) {
            const float sj = sm[j];
            cnt += (sj > si) || (sj == si && j < i);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        if (cnt < k) {
            float* ob = a.out[z] + ((long long)b * a.kmax[z] + cnt) * D;
            for (int d = lane; d < D; d += 32) ob[d] = hb[i * D + d] * si;
        }
    }
    if (tid == 0 && a.kout[z]) a.kout[z][b] = k;
}
cudaError_t launch_graph_pool(const PoolArgs& a, int B, int nz, int D, int nmax, cudaStream_t s) {
    dim3 grid(1, B, nz);
    const size_t smem = (size_t)nmax * sizeof(float);
    if (D == 64) graph_pool_kernel<64><<<grid, 128, smem, s>>>(a);
    else if (D == 32) graph_pool_kernel<32><<<grid, 128, smem, s>>>(a);
    else return cudaErrorInvalidValue;
    return cudaGetLastError();
}

// =================================================================================================
// HS-GAL input: xc = cat( proj_type1(x1), proj_type2(x2) )   (sslassist.py:201-206), both branches per launch.
template <int D>
__global__ void __launch_bounds__(128)
hproj_kernel(const HprojArgs a) {
    const int z = blockIdx.z, b = blockIdx.y;
    const int n1 = a.n1[z] ? a.n1[z][b] : a.n1c[z];
    const int n2 = a.n2[z] ? a.n2[z][b] : a.n2c[z];
    const int nm = a.n1max[z] + a.n2max[z];
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;    // node * D + o
    if (idx >= (n1 + n2) * D) return;
    const int i = idx / D, o = idx - i * D;
    const bool t1 = i < n1;
    const float* xr = t1 ? a.x1[z] + ((long long)b * a.n1max[z] + i) * D : a.x2[z] + ((long long)b * a.n2max[z] + (i - n1)) * D;
    const float* wr = (t1 ? a.W1[z] : a.W2[z]) + o * D;
    float v = __ldg((t1 ? a.b1[z] : a.b2[z]) + o);
#pragma unroll 8
    for (int d = 0; d < D; ++d) v = fmaf(__ldg(wr + d), xr[d], v);
    a.xc[z][((long long)b * nm + i) * D + o] = v;
}
cudaError_t launch_hproj(const HprojArgs& a, int B, int nz, int D, cudaStream_t s) {
    int nm = 0;
    for (int z = 0; z < nz; ++z) nm = max(nm, a.n1max[z] + a.n2max[z]);
    dim3 grid((nm * D + 127) / 128, B, nz);
    if (D == 64) hproj_kernel<64><<<grid, 128, 0, s>>>(a);
    else if (D == 32) hproj_kernel<32><<<grid, 128, 0, s>>>(a);
    else return cudaErrorInvalidValue;
    return cudaGetLastError();
}

// =================================================================================================
// Readout (sslassist.py:557-597).  Per branch z: T_z = T1_z + Taug_z, S_z = S1_z + Saug_z, M_z = m_z + maug_z;
// T = max(T_1, T_2) element-wise (row i of branch 1 pairs with row i of branch 2), same for S, M;
// emb = [max_i |T|, mean_i T, max_i |S|, mean_i S, M]  (5 x 32);  out = W emb + b.
__global__ void __launch_bounds__(32)
readout_kernel(const ReadoutArgs a) {
    __shared__ float se[160];
    const int b = blockIdx.x, d = threadIdx.x;   // 32 threads = 32 feature dims
    const int kT = a.kT ? a.kT[b] : a.kTc;
    const int kS = a.kS ? a.kS[b] : a.kSc;
    const int nm = a.kTmax + a.kSmax;
    float tmax = 0.f, tsum = 0.f, smax = 0.f, ssum = 0.f;
    for (int i = 0; i < kT; ++i) {
        float v[2];
#pragma unroll
        for (int z = 0; z < 2; ++z)
            v[z] = a.T1[z][((long long)b * a.kTmax + i) * 32 + d] + a.y2[z][((long long)b * nm + i) * 32 + d];
        const float t = fmaxf(v[0], v[1]);
        tmax = fmaxf(tmax, fabsf(t));
        tsum += t;
    }
    for (int i = 0; i < kS; ++i) {
        float v[2];
#pragma unroll
        for (int z = 0; z < 2; ++z)
            v[z] = a.S1[z][((long long)b * a.kSmax + i) * 32 + d] + a.y2[z][((long long)b * nm + kT + i) * 32 + d];
        const float t = fmaxf(v[0], v[1]);
        smax = fmaxf(smax, fabsf(t));
        ssum += t;
    }
    const float m0 = a.m[0][(long long)b * 32 + d] + a.maug[0][(long long)b * 32 + d];
    const float m1 = a.m[1][(long long)b * 32 + d] + a.maug[1][(long long)b * 32 + d];
    se[d] = tmax;
    se[32 + d] = tsum / (float)kT;
    se[64 + d] = smax;
    se[96 + d] = ssum / (float)kS;
    se[128 + d] = fmaxf(m0, m1);
    __syncwarp();
    if (a.emb) for (int i = d; i < 160; i += 32) a.emb[(long long)b * 160 + i] = se[i];
    if (a.out) {
        for (int c = 0; c < 2; ++c) {
            float acc = 0.f;
            for (int i = d; i < 160; i += 32) acc = fmaf(__ldg(a.Wout + c * 160 + i), se[i], acc);
            acc = warp_sum(acc);
            if (d == 0) a.out[(long long)b * 2 + c] = acc + __ldg(a.bout + c);
        }
    }
}
cudaError_t launch_readout(const ReadoutArgs& a, int B, cudaStream_t s) {
    readout_kernel<<<B, 32, 0, s>>>(a);
    return cudaGetLastError();
}

}  // namespace occ
