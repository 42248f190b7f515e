// occ_internal.h -- host-side structures shared by the translation units of libocc_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <string>

namespace occ {

typedef __nv_bfloat16 bf16;

// One "shift-GEMM" on the tcgen05 tensor cores:
//   out[b, t, n] = epi( sum_{tap} sum_{c < Ct} A[b, row(t, tap), sub(tap), col0(n) + c] * W[n, tap*Ct + c] )
// A is addressed through a 5-D TMA map {cols, sub, rows, batch, plane}.  Tap tau has the linear offset
//   lin = shift0 + (tau / ntaps_w) * shift_h + (tau % ntaps_w) * shift_w,
// and reads A row  t + floor(lin / a_sub)  at sub-position  lin mod a_sub.  Out-of-range rows read as zeros
// (TMA fill), which is how conv padding is implemented.  Covers nn.Linear (1 tap), the stride-2 wav2vec2 conv
// layers (a_sub = 2: frame = 2*row + sub), and the grouped positional conv (128 taps, N-tile = group).
struct GemmDesc {
    // ---- A operand, bf16
    const bf16* A = nullptr;
    int a_cols = 0;                 // addressable columns (channels) of one A (row, sub) position
    int a_sub = 1;                  // sub-positions per row (conv stride), 1 for plain matrices
    int64_t a_sub_stride = 0;       // elements between sub-positions
    int64_t a_row_stride = 0;       // elements between consecutive rows
    int a_rows = 0;                 // rows per batch item visible to TMA (bounds for zero fill)
    int64_t a_batch_stride = 0;     // elements
    int nbatch = 1;
    int64_t a_plane_stride = 0;     // elements between the hi and lo planes (split mode)
    // ---- W operand, bf16 [N, ktot] row-major (K-major), ktot = nkb*64
    const bf16* W = nullptr;
    int N = 0;
    int64_t w_plane_stride = 0;
    // ---- K loop
    int nkb = 0;                    // 64-wide k-blocks per pass
    int kb_per_tap = 0;             // k-blocks per tap (= nkb when there is a single tap)
    int ntaps_w = 1;
    int shift0 = 0, shift_h = 0, shift_w = 0;
    int a_col_group_stride = 0;     // A column origin of N-tile j = j * a_col_group_stride (grouped conv), else 0
    int split = 0;                  // 1: three passes hi*hi + hi*lo + lo*hi (fp32-accurate)
    int k_real = 0;                 // when > 0: only the first k_real columns of every 64-wide k-block are non-zero (in A and W), so
                                    // the K = 16 MMA steps beyond them are skipped (AASIST blocks with 32 input channels)
    // ---- M tiling / output rows
    int rows_per_batch = 0;         // valid output rows per batch item; output row = b*rows_per_batch + t
    // ---- epilogue: v = acc + bias[n]; if gelu v = gelu(v); v += resid[row, n]
    const float* bias = nullptr;
    const float* resid = nullptr;   // fp32 [rows, ld_resid]
    int64_t ld_resid = 0;
    float* out32 = nullptr;         // optional fp32 output
    int64_t ld32 = 0;
    float* out32_copy = nullptr;    // optional second fp32 destination of the same values (same ld32): the per-layer hidden-state export
                                    // of the FC2 epilogue (models/xlsr.py:39-52 all-layers form) -- no separate device-to-device copy
    bf16* out16 = nullptr;          // optional bf16 output (hi plane; lo plane at +out16_plane_stride when != 0)
    int64_t ld16 = 0;
    int64_t out16_plane_stride = 0;
    int gelu = 0;
    // Fused per-row LayerNorm + erf-GELU (the wav2vec 2.0 conv block, torchaudio components.py:90-93): when ln_gamma != nullptr
    // the epilogue computes v = acc + bias over the WHOLE output row (N must be 512 = both 256-column accumulator stages of a
    // 256-row tile), mean and variance of the row in fp32 (two passes over tensor memory, as F.layer_norm: eps 1e-5, biased
    // variance), y = gelu((v - mean) * rstd * gamma + beta), and writes bf16 once -- no fp32 round trip through HBM and no
    // separate LayerNorm launch.  bf16 output only (out16, no lo plane), BN = 256, no split.
    const float* ln_gamma = nullptr;
    const float* ln_beta = nullptr;
    // LayerNorm folded into a linear layer (transformer pre-LN blocks, torchaudio components.py:345-401).  LN(x) W^T + b =
    //   rstd * (x W'^T - mean * s) + c   with  W' = W diag(gamma),  s[n] = sum_k W'[n,k],  c[n] = b[n] + sum_k beta[k] W[n,k],
    // so the GEMM reads the RAW residual stream (bf16) and the per-row statistics enter in the epilogue.
    //   producer side (out-proj / FC2, fp32 output + residual): ln_x16_out receives bf16(out) and ln_stat_out the per-row partial
    //     sums {sum x, sum x^2} of this launch's output, one float2 per (row, 256-column tile, epilogue column group): [rows][12][2]
    //     for N = 1024 -- fixed slots, summed in a fixed order by the consumer (deterministic, batch-invariant);
    //   consumer side (QKV / FC1, bf16 output): ln_stat_in = that array, ln_colsum = s, bias = c.
    bf16* ln_x16_out = nullptr;
    float* ln_stat_out = nullptr;
    const float* ln_stat_in = nullptr;
    const float* ln_colsum = nullptr;
    // Toeplitz pos-conv output mapping (toep_T > 0; BN = 256, bf16 output, GELU): the GEMM runs over the FLAT padded frame sequence
    // (utterance b occupies frames [b*toep_Tp, b*toep_Tp + toep_T), zeros between), one M row per 4 frames (a_sub = 4); column
    // (jj*64 + n) of N-tile grp is frame 4*row + jj, channel grp*64 + n.  The epilogue writes y16[(b*toep_T + t)*1024 + channel]
    // for t < toep_T and drops the gap frames.
    int toep_T = 0, toep_Tp = 0, toep_B = 0;
    const int32_t* row_valid = nullptr;  // optional [nbatch]: rows t >= row_valid[b] are written as zeros
    int BN = 256;                   // N tile: 64, 128 or 256
    // ---- conv2d epilogue of the AASIST residual blocks (conv_mode = 1; N = 64, channel-last padded activations):
    //   v = acc + bias[n]; [v = selu(v*cs1[n] + ct1[n])]; [v += resid_hi + resid_lo]; [v = selu(v*cs2[n] + ct2[n])];
    //   positions outside rows [conv_h0, conv_h1) x columns [1, wl] of the padded (Hp x conv_wp) grid are written as 0
    //   (they are the zero padding the next conv reads).  Output: out16 hi/lo planes, or out32.
    int conv_mode = 0;
    const float *cs1 = nullptr, *ct1 = nullptr, *cs2 = nullptr, *ct2 = nullptr;
    const bf16* cresid = nullptr;       // hi plane, lo plane at + cresid_plane
    int64_t cresid_plane = 0;
    int conv_wp = 0, conv_h0 = 0, conv_h1 = 0, conv_wm = 0;
    const int32_t* conv_wlen = nullptr; // optional per-utterance valid width (else conv_wm)
};

// Per-(device, kernel) opt-in to more than 48 KB of dynamic shared memory.  cudaFuncSetAttribute is a per-device setting, so
// the cache is keyed by the CURRENT device (two handles on two GPUs in one process each get their own opt-in); thread-safe.
cudaError_t ensure_dyn_smem(const void* func, size_t bytes);      // gemm_tc.cu

// gemm_tc.cu
cudaError_t launch_gemm_tc(const GemmDesc& d, cudaStream_t s, int num_sms, std::string* err, int probe = 0);
cudaError_t launch_gemm_simt(const GemmDesc& d, cudaStream_t s);

}  // namespace occ
