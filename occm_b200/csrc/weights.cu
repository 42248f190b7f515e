// weights.cu -- one-time weight preparation kernels (occ_finalize_weights): bf16 conversion with an optional
// low-order plane, K-major re-layout of the conv weights, weight-norm and eval-BatchNorm folding.
#include "common.cuh"
#include "kernels.h"

namespace occ {

// out[i] = bf16(in[i] * (row < scale_rows ? scale : 1)) (+ lo plane); in is [rows][cols]
__global__ void pack_linear_kernel(const float* __restrict__ in, long long rows, long long cols, long long scale_rows, float scale,
                                   bf16* __restrict__ out, long long plane) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * cols) return;
    const long long r = i / cols;
    const float v = in[i] * (r < scale_rows ? scale : 1.0f);
    bf16 h, l;
    split_bf16(v, h, l);
    out[i] = h;
    if (plane) out[plane + i] = l;
}
cudaError_t launch_pack_linear(const float* in, long long rows, long long cols, long long scale_rows, float scale, bf16* out,
                               long long plane, cudaStream_t s) {
    const long long n = rows * cols;
    pack_linear_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(in, rows, cols, scale_rows, scale, out, plane);
    return cudaGetLastError();
}

// LayerNorm folded into a linear layer (GemmDesc::ln_stat_in): one CTA per output row n.
//   w = in[n][k] * scale;  W'[n][k] = bf16(w * gamma[k]);  s[n] = sum_k float(W'[n][k])  (the ROUNDED weights: a constant row then cancels exactly);
//   c[n] = bias[n] * scale + sum_k w * beta[k].   Fixed reduction order (thread-strided partials, tree over the block).
__global__ void __launch_bounds__(256)
pack_linear_ln_kernel(const float* __restrict__ in, long long cols, long long scale_rows, float scale, const float* __restrict__ gamma,
                      const float* __restrict__ beta, const float* __restrict__ bias, bf16* __restrict__ out, float* __restrict__ colsum,
                      float* __restrict__ cbias) {
    __shared__ float ss[256], sc[256];
    const long long n = blockIdx.x;
    const float f = n < scale_rows ? scale : 1.0f;
    float as = 0.f, ac = 0.f;
    for (long long k = threadIdx.x; k < cols; k += 256) {
        const float w = in[n * cols + k] * f;
        const bf16 h = __float2bfloat16_rn(w * gamma[k]);
        out[n * cols + k] = h;
        as += __bfloat162float(h);
        ac = fmaf(w, beta[k], ac);
    }
    ss[threadIdx.x] = as; sc[threadIdx.x] = ac;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) { ss[threadIdx.x] += ss[threadIdx.x + o]; sc[threadIdx.x] += sc[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { colsum[n] = ss[0]; cbias[n] = bias[n] * f + sc[0]; }
}
cudaError_t launch_pack_linear_ln(const float* in, long long rows, long long cols, long long scale_rows, float scale, const float* gamma,
                                  const float* beta, const float* bias, bf16* out, float* colsum, float* cbias, cudaStream_t s) {
    pack_linear_ln_kernel<<<(unsigned)rows, 256, 0, s>>>(in, cols, scale_rows, scale, gamma, beta, bias, out, colsum, cbias);
    return cudaGetLastError();
}

// Conv1d weight [Cout][Cin][k] -> [Cout][k*Cin] with K index j*Cin + ci (matches channel-last activation rows)
__global__ void pack_conv1d_kernel(const float* __restrict__ in, int Cout, int Cin, int k, bf16* __restrict__ out, long long plane) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long n = (long long)Cout * Cin * k;
    if (i >= n) return;
    const int ci = (int)(i % Cin);
    const int j = (int)((i / Cin) % k);
    const int co = (int)(i / ((long long)Cin * k));
    const float v = in[((long long)co * Cin + ci) * k + j];
    bf16 h, l;
    split_bf16(v, h, l);
    out[i] = h;
    if (plane) out[plane + i] = l;
}
cudaError_t launch_pack_conv1d(const float* in, int Cout, int Cin, int k, bf16* out, long long plane, cudaStream_t s) {
    const long long n = (long long)Cout * Cin * k;
    pack_conv1d_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(in, Cout, Cin, k, out, plane);
    return cudaGetLastError();
}

// pos-conv: weight_norm(dim=2): W[co][ci][j] = g[j] * v[co][ci][j] / ||v[:,:,j]||_F
// (torchaudio components.py:211).  Step 1: norms[j];  step 2: pack to [co][j*64 + ci].
__global__ void posconv_norm_kernel(const float* __restrict__ v, int n_rows /*1024*64*/, int k, float* __restrict__ norms) {
    __shared__ double red[256];
    const int j = blockIdx.x;
    double acc = 0.0;
    for (int r = threadIdx.x; r < n_rows; r += blockDim.x) {
        const double x = (double)v[(long long)r * k + j];
        acc += x * x;
    }
    red[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) norms[j] = (float)sqrt(red[0]);
}
__global__ void pack_posconv_kernel(const float* __restrict__ v, const float* __restrict__ g, const float* __restrict__ norms,
                                    int Cout, int Cg, int k, bf16* __restrict__ out, long long plane) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;     // co*(k*Cg) + j*Cg + ci
    const long long n = (long long)Cout * Cg * k;
    if (i >= n) return;
    const int ci = (int)(i % Cg);
    const int j = (int)((i / Cg) % k);
    const int co = (int)(i / ((long long)Cg * k));
    const float w = v[((long long)co * Cg + ci) * k + j] * (g[j] / norms[j]);
    bf16 h, l;
    split_bf16(w, h, l);
    out[i] = h;
    if (plane) out[plane + i] = l;
}
cudaError_t launch_pack_posconv(const float* v, const float* g, float* norms_tmp, bf16* out, long long plane, cudaStream_t s) {
    posconv_norm_kernel<<<128, 256, 0, s>>>(v, 1024 * 64, 128, norms_tmp);
    const long long n = 1024LL * 64 * 128;
    pack_posconv_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(v, g, norms_tmp, 1024, 64, 128, out, plane);
    return cudaGetLastError();
}

// Toeplitz-expanded pos-conv weight (GemmDesc::toep_T): four adjacent output frames share one N = 256 tile.  Row
// (grp*256 + jj*64 + n), column (tap'*64 + ci), tap' in [0, 131): w[grp*64+n][ci][tap' - jj] when 0 <= tap' - jj < 128, else 0
// (out[4r + jj] = sum_j w[j] x[4r + jj + j - 64] = sum_tap' w[tap' - jj] x[4r + tap' - 64]).
__global__ void pack_posconv_toeplitz_kernel(const float* __restrict__ v, const float* __restrict__ g, const float* __restrict__ norms,
                                             bf16* __restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long K = 131LL * 64;
    if (i >= 4096LL * K) return;
    const int ci = (int)(i % 64);
    const int tp = (int)((i / 64) % 131);
    const int row = (int)(i / K);
    const int n = row & 63, jj = (row >> 6) & 3, grp = row >> 8;
    const int j = tp - jj;
    float w = 0.f;
    if (j >= 0 && j < 128) w = v[((long long)(grp * 64 + n) * 64 + ci) * 128 + j] * (g[j] / norms[j]);
    out[i] = __float2bfloat16_rn(w);
}
cudaError_t launch_pack_posconv_toeplitz(const float* v, const float* g, float* norms_tmp, bf16* out, cudaStream_t s) {
    posconv_norm_kernel<<<128, 256, 0, s>>>(v, 1024 * 64, 128, norms_tmp);
    const long long n = 4096LL * 131 * 64;
    pack_posconv_toeplitz_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(v, g, norms_tmp, out);
    return cudaGetLastError();
}

// eval BatchNorm -> y = x*scale + shift
__global__ void fold_bn_kernel(const float* __restrict__ w, const float* __restrict__ b, const float* __restrict__ mean,
                               const float* __restrict__ var, int C, float eps, float* __restrict__ scale, float* __restrict__ shift) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float sc = w[c] / sqrtf(var[c] + eps);
    scale[c] = sc;
    shift[c] = b[c] - mean[c] * sc;
}
cudaError_t launch_fold_bn(const float* w, const float* b, const float* mean, const float* var, int C, float* scale, float* shift, cudaStream_t s) {
    fold_bn_kernel<<<(C + 127) / 128, 128, 0, s>>>(w, b, mean, var, C, 1e-5f, scale, shift);
    return cudaGetLastError();
}

// Conv2d weight [Cout][Cin][KH][3] -> [Cout/16][Cin][KH*3][16]
__global__ void pack_conv2d_kernel(const float* __restrict__ in, int Cout, int Cin, int taps, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = Cout * Cin * taps;
    if (i >= n) return;
    const int l = i % 16;
    const int tap = (i / 16) % taps;
    const int ci = (i / (16 * taps)) % Cin;
    const int cg = i / (16 * taps * Cin);
    out[i] = in[((cg * 16 + l) * Cin + ci) * taps + tap];
}
cudaError_t launch_pack_conv2d(const float* in, int Cout, int Cin, int taps, float* out, cudaStream_t s) {
    const int n = Cout * Cin * taps;
    pack_conv2d_kernel<<<(n + 255) / 256, 256, 0, s>>>(in, Cout, Cin, taps, out);
    return cudaGetLastError();
}

// conv2d weight [Cout][Cin][taps] fp32 -> tcgen05 B operand bf16 hi/lo [64][taps*64] (K-major, k = tap*64 + ci), zero padded
// to 64 output / 64 input channels (AASIST residual blocks as shift-GEMMs, occ_api.cu stage_backend).
__global__ void pack_conv2d_tc_kernel(const float* __restrict__ in, int Cout, int Cin, int taps, bf16* __restrict__ out, long long plane) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int K = taps * 64;
    if (i >= 64 * K) return;
    const int n = i / K, k = i - n * K;
    const int tap = k / 64, ci = k - tap * 64;
    const float v = (n < Cout && ci < Cin) ? in[((long long)n * Cin + ci) * taps + tap] : 0.f;
    bf16 h, l;
    split_bf16(v, h, l);
    out[i] = h;
    out[plane + i] = l;
}
cudaError_t launch_pack_conv2d_tc(const float* in, int Cout, int Cin, int taps, bf16* out, long long plane, cudaStream_t s) {
    const int n = 64 * taps * 64;
    pack_conv2d_tc_kernel<<<(n + 255) / 256, 256, 0, s>>>(in, Cout, Cin, taps, out, plane);
    return cudaGetLastError();
}
// out[0..total) = in[i] for i < n, else fill
__global__ void pad_vec_kernel(const float* __restrict__ in, int n, float fill, float* __restrict__ out, int total) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < total) out[i] = i < n ? in[i] : fill;
}
cudaError_t launch_pad_vec(const float* in, int n, float fill, float* out, int total, cudaStream_t s) {
    pad_vec_kernel<<<(total + 127) / 128, 128, 0, s>>>(in, n, fill, out, total);
    return cudaGetLastError();
}

// out[c][r] = in[r][c], scaled vector copy helpers
__global__ void transpose_kernel(const float* __restrict__ in, int R, int C, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R * C) return;
    const int r = i / C, c = i - r * C;
    out[c * R + r] = in[i];
}
cudaError_t launch_transpose(const float* in, int R, int C, float* out, cudaStream_t s) {
    transpose_kernel<<<(R * C + 255) / 256, 256, 0, s>>>(in, R, C, out);
    return cudaGetLastError();
}
__global__ void scale_copy_kernel(const float* __restrict__ in, int n, int scale_n, float scale, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[i] * (i < scale_n ? scale : 1.0f);
}
cudaError_t launch_scale_copy(const float* in, int n, int scale_n, float scale, float* out, cudaStream_t s) {
    scale_copy_kernel<<<(n + 255) / 256, 256, 0, s>>>(in, n, scale_n, scale, out);
    return cudaGetLastError();
}

}  // namespace occ
