// backend_attmap.cu -- spectro-temporal attention map and pooling of AASIST (models/sslassist.py:465-471, 520-538).
//   w = Conv1x1(128->64)( BN( SELU( Conv1x1(64->128)(x) ) ) )  per (frequency, time) position   -> att_map_mma_kernel
//   e_S[b][h][c] = sum_w x * softmax_w(w) + pos_S ;  e_T[b][w][c] = sum_h x * softmax_h(w)     -> pool_S2 / pool_T2
// The two 1x1 convolutions are a [P x 64][64 x 128] and a [P x 128][128 x 64] product per position tile: they run on the
// warp-level tensor-core path (mma.sync m16n8k16, bf16 operands split hi + lo, three MMAs hi*hi + hi*lo + lo*hi with fp32
// accumulation ~ 2^-17 relative, same scheme as the residual-encoder GEMMs) with the 128-wide hidden activations kept in
// registers (the accumulator fragment of the first product is the A fragment of the second).  The former SIMT kernel was
// bound by broadcast shared-memory loads (profiles/r1c_backend_ncu_full.txt: 303 us, L1 72 %).
// x is the channel-last fp32 residual-encoder output; the map is written channel-last [B][42*Wm][64] as well.
#include "common.cuh"
#include "kernels.h"

namespace occ {

static constexpr int AM_LD1 = 72;     // bf16 row pitch of W1 [128][64] in shared memory: 36 words -> bank = 4 g + q, conflict-free
static constexpr int AM_LD2 = 136;    // bf16 row pitch of W2 [64][128]: 68 words -> bank = 4 g + q
static constexpr int AM_SMEM = 2 * (128 * AM_LD1 + 64 * AM_LD2) * 2;

__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// (x, y) -> packed bf16 pair of the high parts and of the residuals
__device__ __forceinline__ void split_pack(float x, float y, uint32_t& hi, uint32_t& lo) {
    hi = pack_bf16(x, y);
    const __nv_bfloat162 h2 = *reinterpret_cast<const __nv_bfloat162*>(&hi);
    lo = pack_bf16(x - __low2float(h2), y - __high2float(h2));
}

__global__ void __launch_bounds__(256)
att_map_mma_kernel(XView xv, int Wm, int B, const float* __restrict__ W1, const float* __restrict__ b1,
                   const float* __restrict__ bs, const float* __restrict__ bt, const float* __restrict__ W2,
                   const float* __restrict__ b2, float* __restrict__ wmap) {
    pdl_wait();       // launched with the PDL attribute: nothing above this line touches global memory
    extern __shared__ __align__(16) uint8_t am_smem[];
    bf16* sW1h = reinterpret_cast<bf16*>(am_smem);          // [128][AM_LD1]
    bf16* sW1l = sW1h + 128 * AM_LD1;
    bf16* sW2h = sW1l + 128 * AM_LD1;                       // [64][AM_LD2]
    bf16* sW2l = sW2h + 64 * AM_LD2;
    for (int i = threadIdx.x; i < 128 * 64; i += 256) {
        bf16 hi, lo;
        split_bf16(__ldg(W1 + i), hi, lo);
        const int n = i >> 6, k = i & 63;
        sW1h[n * AM_LD1 + k] = hi; sW1l[n * AM_LD1 + k] = lo;
        split_bf16(__ldg(W2 + i), hi, lo);
        const int n2 = i >> 7, k2 = i & 127;
        sW2h[n2 * AM_LD2 + k2] = hi; sW2l[n2 * AM_LD2 + k2] = lo;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, q = lane & 3;
    const int HW = 42 * Wm;
    const int tiles_per_b = (HW + 15) >> 4;
    const int ntiles = B * tiles_per_b;
    for (int tile = blockIdx.x * 8 + warp; tile < ntiles; tile += gridDim.x * 8) {
        const int b = tile / tiles_per_b;
        const int p0 = (tile - b * tiles_per_b) << 4;
        const int pA = p0 + g, pB = p0 + g + 8;
        const int cA = min(pA, HW - 1), cB = min(pB, HW - 1);
        const float* xa = xv.x + (long long)b * xv.xb + (long long)(cA / Wm) * xv.xh + (long long)(cA % Wm) * xv.xw;
        const float* xb = xv.x + (long long)b * xv.xb + (long long)(cB / Wm) * xv.xh + (long long)(cB % Wm) * xv.xw;
        // ---- A fragments of x (16 positions x 64 channels), hi / lo
        uint32_t ah[4][4], al[4][4];
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
            const float2 v0 = *reinterpret_cast<const float2*>(xa + ks * 16 + 2 * q);
            const float2 v1 = *reinterpret_cast<const float2*>(xb + ks * 16 + 2 * q);
            const float2 v2 = *reinterpret_cast<const float2*>(xa + ks * 16 + 2 * q + 8);
            const float2 v3 = *reinterpret_cast<const float2*>(xb + ks * 16 + 2 * q + 8);
            split_pack(v0.x, v0.y, ah[ks][0], al[ks][0]);
            split_pack(v1.x, v1.y, ah[ks][1], al[ks][1]);
            split_pack(v2.x, v2.y, ah[ks][2], al[ks][2]);
            split_pack(v3.x, v3.y, ah[ks][3], al[ks][3]);
        }
        // ---- hidden = BN(SELU(x W1^T + b1)): 16 n-tiles of 8; accumulator pairs become the A fragments of the second product
        uint32_t hh[8][4], hl[8][4];
#pragma unroll
        for (int nt = 0; nt < 16; ++nt) {
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            const bf16* wh = sW1h + (nt * 8 + g) * AM_LD1 + 2 * q;
            const bf16* wl = sW1l + (nt * 8 + g) * AM_LD1 + 2 * q;
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                const uint32_t bh0 = *reinterpret_cast<const uint32_t*>(wh + ks * 16), bh1 = *reinterpret_cast<const uint32_t*>(wh + ks * 16 + 8);
                const uint32_t bl0 = *reinterpret_cast<const uint32_t*>(wl + ks * 16), bl1 = *reinterpret_cast<const uint32_t*>(wl + ks * 16 + 8);
                mma16816(acc, ah[ks], bh0, bh1);
                mma16816(acc, ah[ks], bl0, bl1);
                mma16816(acc, al[ks], bh0, bh1);
            }
            const int j = nt * 8 + 2 * q;
            const float2 bb = __ldg(reinterpret_cast<const float2*>(b1 + j));
            const float2 s2 = __ldg(reinterpret_cast<const float2*>(bs + j));
            const float2 t2 = __ldg(reinterpret_cast<const float2*>(bt + j));
            const float v0 = fmaf(selu_f(acc[0] + bb.x), s2.x, t2.x), v1 = fmaf(selu_f(acc[1] + bb.y), s2.y, t2.y);
            const float v2 = fmaf(selu_f(acc[2] + bb.x), s2.x, t2.x), v3 = fmaf(selu_f(acc[3] + bb.y), s2.y, t2.y);
            const int ks2 = nt >> 1, o = (nt & 1) * 2;          // even n-tile: k 2q..2q+1 (regs 0,1); odd: k 2q+8.. (regs 2,3)
            split_pack(v0, v1, hh[ks2][o], hl[ks2][o]);
            split_pack(v2, v3, hh[ks2][o + 1], hl[ks2][o + 1]);
        }
        // ---- w = hidden W2^T + b2: 8 n-tiles of 8 output channels
        float* oa = wmap + ((long long)b * HW + pA) * 64;
        float* ob = wmap + ((long long)b * HW + pB) * 64;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            const bf16* wh = sW2h + (nt * 8 + g) * AM_LD2 + 2 * q;
            const bf16* wl = sW2l + (nt * 8 + g) * AM_LD2 + 2 * q;
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
                const uint32_t bh0 = *reinterpret_cast<const uint32_t*>(wh + ks * 16), bh1 = *reinterpret_cast<const uint32_t*>(wh + ks * 16 + 8);
                const uint32_t bl0 = *reinterpret_cast<const uint32_t*>(wl + ks * 16), bl1 = *reinterpret_cast<const uint32_t*>(wl + ks * 16 + 8);
                mma16816(acc, hh[ks], bh0, bh1);
                mma16816(acc, hh[ks], bl0, bl1);
                mma16816(acc, hl[ks], bh0, bh1);
            }
            const int c = nt * 8 + 2 * q;
            const float2 bb = __ldg(reinterpret_cast<const float2*>(b2 + c));
            if (pA < HW) *reinterpret_cast<float2*>(oa + c) = make_float2(acc[0] + bb.x, acc[1] + bb.y);
            if (pB < HW) *reinterpret_cast<float2*>(ob + c) = make_float2(acc[2] + bb.x, acc[3] + bb.y);
        }
    }
}

cudaError_t launch_att_map(XView x, int B, int Wm, const float* W1, const float* b1, const float* bs, const float* bt,
                           const float* W2, const float* b2, float* wmap, int num_sms, cudaStream_t s) {
    if (x.xc != 1) return cudaErrorInvalidValue;          // channel-last input only
    {
        cudaError_t e = ensure_dyn_smem(reinterpret_cast<const void*>(att_map_mma_kernel), AM_SMEM);
        if (e != cudaSuccess) return e;
    }
    const int ntiles = B * ((42 * Wm + 15) / 16);
    int grid = (ntiles + 7) / 8;
    if (grid > 2 * num_sms) grid = 2 * num_sms;
    (void)launch_pdl(att_map_mma_kernel, dim3(grid), dim3((unsigned)(256)), (size_t)(AM_SMEM), s, x, Wm, B, W1, b1, bs, bt, W2, b2, wmap);
    return cudaGetLastError();
}

// =================================================================================================
// Pooling over channel-last x and wmap: thread = channel, so every load is a coalesced 256 B row.
//   pool_S2: one 64-thread group per (b, h): softmax over the valid widths  w < wl
//   pool_T2: one 64-thread group per (b, w): softmax over the 42 frequencies; w >= wl -> 0
__global__ void __launch_bounds__(256)
pool_S2_kernel(XView xv, const float* __restrict__ wmap, int Wm, const int* __restrict__ wlen, const float* __restrict__ posS,
               float* __restrict__ eS, int B) {
    pdl_wait();       // launched with the PDL attribute: nothing above this line touches global memory
    const int row = blockIdx.x * 4 + (threadIdx.x >> 6);         // b * 42 + h
    const int c = threadIdx.x & 63;
    if (row >= B * 42) return;
    const int b = row / 42, h = row - b * 42;
    const int wl = wlen ? wlen[b] : Wm;
    const float* wr = wmap + ((long long)b * 42 * Wm + (long long)h * Wm) * 64 + c;
    const float* xr = xv.x + (long long)b * xv.xb + (long long)h * xv.xh + c;
    float m = -INFINITY;
    for (int w = 0; w < wl; ++w) m = fmaxf(m, wr[(long long)w * 64]);
    float den0 = 0.f, num0 = 0.f, den1 = 0.f, num1 = 0.f;
    int w = 0;
    for (; w + 1 < wl; w += 2) {
        const float e0 = expf(wr[(long long)w * 64] - m), e1 = expf(wr[(long long)(w + 1) * 64] - m);
        den0 += e0; den1 += e1;
        num0 = fmaf(e0, xr[(long long)w * xv.xw], num0);
        num1 = fmaf(e1, xr[(long long)(w + 1) * xv.xw], num1);
    }
    if (w < wl) { const float e0 = expf(wr[(long long)w * 64] - m); den0 += e0; num0 = fmaf(e0, xr[(long long)w * xv.xw], num0); }
    eS[(long long)row * 64 + c] = (num0 + num1) / (den0 + den1) + posS[h * 64 + c];
}
__global__ void __launch_bounds__(256)
pool_T2_kernel(XView xv, const float* __restrict__ wmap, int Wm, const int* __restrict__ wlen, float* __restrict__ eT, int B) {
    pdl_wait();       // launched with the PDL attribute: nothing above this line touches global memory
    const int row = blockIdx.x * 4 + (threadIdx.x >> 6);         // b * Wm + w
    const int c = threadIdx.x & 63;
    if (row >= B * Wm) return;
    const int b = row / Wm, w = row - b * Wm;
    const int wl = wlen ? wlen[b] : Wm;
    float r = 0.f;
    if (w < wl) {
        const float* wr = wmap + ((long long)b * 42 * Wm + w) * 64 + c;
        const float* xr = xv.x + (long long)b * xv.xb + (long long)w * xv.xw + c;
        float m = -INFINITY;
#pragma unroll 6
        for (int h = 0; h < 42; ++h) m = fmaxf(m, wr[(long long)h * Wm * 64]);
        float den0 = 0.f, num0 = 0.f, den1 = 0.f, num1 = 0.f;
#pragma unroll 3
        for (int h = 0; h < 42; h += 2) {
            const float e0 = expf(wr[(long long)h * Wm * 64] - m), e1 = expf(wr[(long long)(h + 1) * Wm * 64] - m);
            den0 += e0; den1 += e1;
            num0 = fmaf(e0, xr[(long long)h * xv.xh], num0);
            num1 = fmaf(e1, xr[(long long)(h + 1) * xv.xh], num1);
        }
        r = (num0 + num1) / (den0 + den1);
    }
    eT[(long long)row * 64 + c] = r;
}
cudaError_t launch_st_pool(XView x, const float* wmap, int B, int Wm, const int* wlen, const float* posS, float* eS,
                           float* eT, cudaStream_t s) {
    if (x.xc != 1) return cudaErrorInvalidValue;
    (void)launch_pdl(pool_S2_kernel, dim3((unsigned)((B * 42 + 3) / 4)), dim3((unsigned)(256)), (size_t)(0), s, x, wmap, Wm, wlen, posS, eS, B);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    (void)launch_pdl(pool_T2_kernel, dim3((unsigned)((B * Wm + 3) / 4)), dim3((unsigned)(256)), (size_t)(0), s, x, wmap, Wm, wlen, eT, B);
    return cudaGetLastError();
}

}  // namespace occ
