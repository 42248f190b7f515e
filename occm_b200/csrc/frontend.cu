// frontend.cu -- the memory-bound / small kernels of the XLS-R frontend:
//   conv0_ln_gelu   : Conv1d(1->512,k10,s5)+bias -> LayerNorm(512) -> erf-GELU, channel-last bf16 out
//   rows_ln         : per-row LayerNorm (+GELU) (+second LayerNorm) fp32 -> bf16 (hi/lo) / fp32
//   attn_bf16       : fused multi-head attention, mma.sync bf16, online softmax (T-tiled, key mask)
//   attn_f32        : fp32 SIMT attention for the fp32-accurate mode
// Reference arithmetic: torchaudio/models/wav2vec2/components.py:53-143 (conv blocks), :146-182 (feature
// projection LN), :237-310 (attention), :345-401 (pre-LN encoder layer) -- the in-container statement of
// fairseq wav2vec 2.0 that /root/reference/models/sslassist.py:48 calls.
#include "common.cuh"
#include <cstdlib>
#include <string>
#include "kernels.h"

namespace occ {

// =================================================================================================
// conv0: wav [B, N] fp32 -> out bf16 [B, L0, 512] (hi plane; lo plane at +plane_stride when != 0)
// One warp computes FR consecutive frames; lane owns channel pairs c = i*64 + 2*lane (+1), i < 8.
static constexpr int C0_FR = 4;
static constexpr int C0_WARPS = 8;

__global__ void __launch_bounds__(C0_WARPS * 32)
conv0_ln_gelu_kernel(const float* __restrict__ wav, int N, int L0, int B,
                     const float* __restrict__ w /*[512][10]*/, const float* __restrict__ bias,
                     const float* __restrict__ gamma, const float* __restrict__ beta,
                     bf16* __restrict__ out, long long plane_stride) {
    __shared__ float2 ws[10][256];     // ws[j][p] = (w[2p][j], w[2p+1][j]) with p = i*32+lane <-> c = i*64 + 2*lane
    __shared__ float2 bs[256], gs[256], es[256];
    for (int idx = threadIdx.x; idx < 256; idx += blockDim.x) {
        const int i = idx >> 5, l = idx & 31;
        const int c = i * 64 + 2 * l;
        for (int j = 0; j < 10; ++j) ws[j][idx] = make_float2(w[c * 10 + j], w[(c + 1) * 10 + j]);
        bs[idx] = make_float2(bias[c], bias[c + 1]);
        gs[idx] = make_float2(gamma[c], gamma[c + 1]);
        es[idx] = make_float2(beta[c], beta[c + 1]);
    }
    __syncthreads();
    pdl_wait();          // the weights staged above are immutable after occ_finalize_weights; the waveform may come from the previous kernel
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long groups_per_b = (L0 + C0_FR - 1) / C0_FR;
    const long long total = groups_per_b * B;
    for (long long g = (long long)blockIdx.x * C0_WARPS + warp; g < total; g += (long long)gridDim.x * C0_WARPS) {
        const int b = (int)(g / groups_per_b);
        const int f0 = (int)(g - (long long)b * groups_per_b) * C0_FR;
        const float* x = wav + (long long)b * N + (long long)f0 * 5;
        // samples needed: 5*(FR-1) + 10 = 25
        float xs[5 * (C0_FR - 1) + 10];
#pragma unroll
        for (int j = 0; j < 5 * (C0_FR - 1) + 10; ++j) {
            const long long pos = (long long)f0 * 5 + j;
            xs[j] = pos < N ? __ldg(x + j) : 0.f;
        }
        // channel pairs live packed (fp32x2): the conv FMAs, the LayerNorm and the GELU all run on the FFMA2/FMUL2/FADD2 pipe
        uint64_t acc[C0_FR][8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float2 bb = bs[i * 32 + lane];
            const uint64_t bp = f2_pack(bb.x, bb.y);
#pragma unroll
            for (int f = 0; f < C0_FR; ++f) acc[f][i] = bp;
        }
#pragma unroll
        for (int j = 0; j < 10; ++j) {
            uint64_t xp[C0_FR];
#pragma unroll
            for (int f = 0; f < C0_FR; ++f) xp[f] = f2_splat(xs[5 * f + j]);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float2 ww = ws[j][i * 32 + lane];
                const uint64_t wp = f2_pack(ww.x, ww.y);
#pragma unroll
                for (int f = 0; f < C0_FR; ++f) acc[f][i] = f2_fma(wp, xp[f], acc[f][i]);
            }
        }
#pragma unroll
        for (int f = 0; f < C0_FR; ++f) {
            if (f0 + f >= L0) break;
            uint64_t s2 = acc[f][0];
#pragma unroll
            for (int i = 1; i < 8; ++i) s2 = f2_add(s2, acc[f][i]);
            float sa, sb;
            f2_unpack(s2, sa, sb);
            const float mean = warp_sum(sa + sb) * (1.0f / 512.0f);
            const uint64_t nm = f2_splat(-mean);
            uint64_t q2 = f2_splat(0.f);
#pragma unroll
            for (int i = 0; i < 8; ++i) { const uint64_t d = f2_add(acc[f][i], nm); q2 = f2_fma(d, d, q2); }
            float qa, qb;
            f2_unpack(q2, qa, qb);
            const float rstd = rsqrtf(warp_sum(qa + qb) * (1.0f / 512.0f) + 1e-5f);
            const uint64_t rs = f2_splat(rstd);
            bf16* o = out + ((long long)b * L0 + f0 + f) * 512;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float2 gg = gs[i * 32 + lane], ee = es[i * 32 + lane];
                float y0, y1;
                f2_unpack(f2_fma(f2_mul(f2_add(acc[f][i], nm), rs), f2_pack(gg.x, gg.y), f2_pack(ee.x, ee.y)), y0, y1);
                gelu_erf_fast2(y0, y1);
                const uint32_t h = pack_bf16(y0, y1);
                *reinterpret_cast<uint32_t*>(o + i * 64 + 2 * lane) = h;
                if (plane_stride != 0) {
                    const __nv_bfloat162 hh = *reinterpret_cast<const __nv_bfloat162*>(&h);
                    *reinterpret_cast<uint32_t*>(o + plane_stride + i * 64 + 2 * lane) =
                        pack_bf16(y0 - __low2float(hh), y1 - __high2float(hh));
                }
            }
        }
    }
}

// -------------------------------------------------------------------------------------------------
// conv0 on the warp-level tensor path (mma.sync m16n8k16 bf16, fp32 accumulate) -- the product path.  The conv is K = 10, far
// too thin for a tcgen05 tile, and the SIMT kernel above is bound by the FMA pipe (LayerNorm + erf-GELU + the 10 FMAs per
// element; FFMA2 halves the issue slots, not the pipe time: 2.0 cycles per warp-instruction, tools/ubench/fma_rate.cu), so the
// 10 FMAs and the mean move onto the tensor pipe:
//   * M = 16 frames per warp, N = 8 channels per tile (64 tiles), K = 16 slots;
//   * fp32-class accuracy from TWO MMAs: slots 0..9 carry x_hi * w_hi (first MMA) and x_lo * w_hi (second); the 6 + 6 spare
//     slots carry x_hi * w_lo (taps 6..9 in the first, 0..5 in the second) and the bias as 1 * b_hi + 1 * b_lo (x_lo * w_lo ~ 2^-18
//     relative is dropped, as in every split GEMM of the fp32 mode);
//   * the weights and the bias are CENTRED over the 512 channels at pack time (w'[c][j] = w[c][j] - mean_c w[.][j]): the conv is
//     linear, so the MMA result IS v_c - mean_c(v), LayerNorm's numerator, and only sum d^2 remains to be reduced;
//   * column j of tile n = 4q + i is channel 32q + 8 (j / 2) + 2i + (j % 2): the accumulator columns (2t, 2t + 1) of four consecutive
//     tiles are the 8 ADJACENT channels 32q + 8t .. + 7 of a frame -> one 16-byte bf16 store; channel pairs feed the fp32x2 pipe.
// conv0_pack_frag_kernel writes the weights in B-fragment lane order ([tile][mma][lane] uint2) and (gamma, gamma', beta, beta') per
// channel pair ([q][i][t] float4); both are staged in shared memory once per CTA.
static constexpr int C0M2_FRAG_U2 = 64 * 2 * 32;     // uint2
static constexpr int C0M2_GB_F4 = 256;
__device__ __forceinline__ int c0m2_channel(int n, int j) { return 32 * (n >> 2) + 8 * (j >> 1) + 2 * (n & 3) + (j & 1); }

__global__ void __launch_bounds__(512)
conv0_pack_frag_kernel(const float* __restrict__ w /*[512][10]*/, const float* __restrict__ bias, const float* __restrict__ gamma,
                       const float* __restrict__ beta, uint2* __restrict__ fragb, float4* __restrict__ gb2) {
    __shared__ float mean_s[11];
    __shared__ float wc[512][11];      // centred weights, [c][10] = centred bias
    const int c = threadIdx.x;
    if (c < 11) {                      // fixed summation order (deterministic), fp64
        double acc = 0.0;
        for (int i = 0; i < 512; ++i) acc += (double)(c < 10 ? w[i * 10 + c] : bias[i]);
        mean_s[c] = (float)(acc / 512.0);
    }
    __syncthreads();
    for (int j = 0; j < 10; ++j) wc[c][j] = w[c * 10 + j] - mean_s[j];
    wc[c][10] = bias[c] - mean_s[10];
    __syncthreads();
    for (int idx = threadIdx.x; idx < C0M2_FRAG_U2; idx += blockDim.x) {
        const int lane = idx & 31, v = (idx >> 5) & 1, n = idx >> 6;
        const int g = lane >> 2, t = lane & 3;
        const int ch = c0m2_channel(n, g);
        auto slot = [&](int s) -> uint32_t {      // bf16 bits of B(K slot s, column g) of MMA v
            bf16 hi, lo;
            if (s < 10) { split_bf16(wc[ch][s], hi, lo); return __bfloat16_as_ushort(hi); }
            if (v == 0) {
                if (s < 12) { split_bf16(wc[ch][10], hi, lo); return __bfloat16_as_ushort(s == 10 ? hi : lo); }
                split_bf16(wc[ch][s - 6], hi, lo); return __bfloat16_as_ushort(lo);
            }
            split_bf16(wc[ch][s - 10], hi, lo); return __bfloat16_as_ushort(lo);
        };
        fragb[idx] = make_uint2(slot(2 * t) | (slot(2 * t + 1) << 16), slot(2 * t + 8) | (slot(2 * t + 9) << 16));
    }
    if (c < C0M2_GB_F4) {              // gamma / beta of the channel pair (32q + 8t + 2i, +1) at [(q*4 + i)*4 + t]
        const int t = c & 3, i = (c >> 2) & 3, q = c >> 4;
        const int ch = 32 * q + 8 * t + 2 * i;
        gb2[c] = make_float4(gamma[ch], gamma[ch + 1], beta[ch], beta[ch + 1]);
    }
}

__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint4& a, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b0), "r"(b1));
}

// The kernel: two passes, software-pipelined.  A one-pass form (M = channels, all 128 results of a thread in registers, 168
// registers, 12 warps per SM) measured 0.36 ms: its warps run in lock step through an HMMA phase and an FMA / MUFU phase, so the
// pipes take turns instead of overlapping (ncu: no pipe above 57 %; numbers in DESIGN.md section 4).  Here the MMAs are
// simply issued TWICE (the tensor pipe idles at 14 % otherwise) and the two passes of CONSECUTIVE items are interleaved:
//   pass 1 (item i+1): M = 16 frames per warp, N = 8 channels per tile, 64 tiles x 2 MMAs -> only sum d^2 per frame is kept;
//   pass 2 (item i):   the same MMAs again, four tiles (= 8 adjacent channels per thread and frame) at a time -> normalise, affine,
//                      GELU on packed chains, one 16-byte bf16 store per frame;
// every iteration of the 16-step loop carries 8 MMAs of each pass, so tensor, FMA and MUFU work is mixed at all times.
// Weights are the B operand ([tile][mma][lane] uint2, conv0_pack_frag_kernel); the samples of item i+2 are fetched under item i.
static constexpr int C0M2_WARPS = 8;
template <bool LO>
__global__ void __launch_bounds__(C0M2_WARPS * 32, 2)
conv0_mma2_kernel(const float* __restrict__ wav, int N, int L0, int B, const uint2* __restrict__ frag_g,
                  const float4* __restrict__ gb_g, bf16* __restrict__ out, long long plane_stride) {
    __shared__ uint2 frag[C0M2_FRAG_U2];    // 32 KB
    __shared__ float4 gbs[C0M2_GB_F4];      //  4 KB
    for (int i = threadIdx.x; i < C0M2_FRAG_U2; i += blockDim.x) frag[i] = frag_g[i];
    for (int i = threadIdx.x; i < C0M2_GB_F4; i += blockDim.x) gbs[i] = gb_g[i];
    __syncthreads();
    pdl_wait();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const long long groups_per_b = (L0 + 15) / 16;
    const long long total = groups_per_b * B;
    const long long stride = (long long)gridDim.x * C0M2_WARPS;
    const int oB = (t == 2) ? 6 : 8;
    const int oC = (t == 0) ? 8 : 2 * t - 2;
    float xr[2][6];                          // the lane's six samples of frames g and g + 8 of an item
    auto fetch = [&](long long it) {
#pragma unroll
        for (int h = 0; h < 2; ++h)
#pragma unroll
            for (int j = 0; j < 6; ++j) xr[h][j] = 0.f;
        if (it < total) {
            const int b = (int)(it / groups_per_b);
            const int f0 = (int)(it - (long long)b * groups_per_b) * 16;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const long long base = (long long)(f0 + g + 8 * h) * 5;
                const float* x = wav + (long long)b * N + base;
                auto ld = [&](int j) -> float { return base + j < N ? __ldg(x + j) : 0.f; };
                xr[h][0] = ld(2 * t); xr[h][1] = ld(2 * t + 1); xr[h][2] = ld(oB); xr[h][3] = ld(oB + 1); xr[h][4] = ld(oC); xr[h][5] = ld(oC + 1);
            }
        }
    };
    // A fragments (frames x K slots) of the two MMAs from xr: {a0, a1, a2, a3} = {(g, 2t..), (g+8, 2t..), (g, 2t+8..), (g+8, 2t+8..)}
    auto build = [&](uint4& A0, uint4& A1) {
        uint32_t r[2][4];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            bf16 ha0, la0, ha1, la1, hb0, lb0, hb1, lb1, hc0, lc0, hc1, lc1;
            split_bf16(xr[h][0], ha0, la0); split_bf16(xr[h][1], ha1, la1);
            split_bf16(xr[h][2], hb0, lb0); split_bf16(xr[h][3], hb1, lb1);
            split_bf16(xr[h][4], hc0, lc0); split_bf16(xr[h][5], hc1, lc1);
            auto pk = [](bf16 a, bf16 c) -> uint32_t { return (uint32_t)__bfloat16_as_ushort(a) | ((uint32_t)__bfloat16_as_ushort(c) << 16); };
            r[h][0] = pk(ha0, ha1);
            r[h][1] = (t == 1) ? 0x3F803F80u /* (1.0, 1.0): the bias slots */ : pk(hb0, hb1);
            r[h][2] = pk(la0, la1);
            r[h][3] = (t == 0) ? pk(lc0, lc1) : pk(hc0, hc1);
        }
        A0 = make_uint4(r[0][0], r[1][0], r[0][1], r[1][1]);
        A1 = make_uint4(r[0][2], r[1][2], r[0][3], r[1][3]);
    };
    // sum of d^2 over tile n for frames g (columns 2t, 2t+1 packed) and g + 8
    auto pass1_tile = [&](int n, const uint4& A0, const uint4& A1, uint64_t& qg, uint64_t& qh) {
        float d[4] = {0.f, 0.f, 0.f, 0.f};
        const uint2 w0 = frag[(n * 2 + 0) * 32 + lane], w1 = frag[(n * 2 + 1) * 32 + lane];
        mma_bf16_16816(d, A0, w0.x, w0.y);
        mma_bf16_16816(d, A1, w1.x, w1.y);
        const uint64_t u = f2_pack(d[0], d[1]), v = f2_pack(d[2], d[3]);
        qg = f2_fma(u, u, qg);
        qh = f2_fma(v, v, qh);
    };
    auto rstd_of = [&](uint64_t qa, uint64_t qb) -> float {
        float sa, sb;
        f2_unpack(f2_add(qa, qb), sa, sb);
        float q = sa + sb;
        q += __shfl_xor_sync(0xffffffffu, q, 1);
        q += __shfl_xor_sync(0xffffffffu, q, 2);
        return rsqrtf(q * (1.0f / 512.0f) + 1e-5f);
    };

    long long it = (long long)blockIdx.x * C0M2_WARPS + warp;
    if (it >= total) return;
    uint4 A0, A1;
    fetch(it);
    build(A0, A1);
    fetch(it + stride);
    float rg, rh;
    {
        uint64_t qg[2] = {f2_splat(0.f), f2_splat(0.f)}, qh[2] = {f2_splat(0.f), f2_splat(0.f)};
#pragma unroll 8
        for (int n = 0; n < 64; ++n) pass1_tile(n, A0, A1, qg[n & 1], qh[n & 1]);
        rg = rstd_of(qg[0], qg[1]);
        rh = rstd_of(qh[0], qh[1]);
    }
    for (; it < total; it += stride) {
        const int b = (int)(it / groups_per_b);
        const int f0 = (int)(it - (long long)b * groups_per_b) * 16;
        uint4 N0, N1;                        // the next item (all-zero samples beyond the last one: harmless work, nothing stored)
        build(N0, N1);
        fetch(it + 2 * stride);
        const uint64_t rsg = f2_splat(rg), rsh = f2_splat(rh);
        const bool vg = f0 + g < L0, vh = f0 + g + 8 < L0;
        bf16* og = out + ((long long)b * L0 + f0 + g) * 512 + 8 * t;
        uint64_t qg[2] = {f2_splat(0.f), f2_splat(0.f)}, qh[2] = {f2_splat(0.f), f2_splat(0.f)};
#pragma unroll 2
        for (int q = 0; q < 16; ++q) {
            uint64_t y[8];                   // [2i]: frame g, channels 32q + 8t + 2i, +1;  [2i + 1]: frame g + 8
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int n = 4 * q + i;
                float d[4] = {0.f, 0.f, 0.f, 0.f};
                const uint2 w0 = frag[(n * 2 + 0) * 32 + lane], w1 = frag[(n * 2 + 1) * 32 + lane];
                mma_bf16_16816(d, A0, w0.x, w0.y);
                mma_bf16_16816(d, A1, w1.x, w1.y);
                {   // pass 1 of the next item on the same weight fragments
                    float e[4] = {0.f, 0.f, 0.f, 0.f};
                    mma_bf16_16816(e, N0, w0.x, w0.y);
                    mma_bf16_16816(e, N1, w1.x, w1.y);
                    const uint64_t u = f2_pack(e[0], e[1]), v = f2_pack(e[2], e[3]);
                    qg[i & 1] = f2_fma(u, u, qg[i & 1]);
                    qh[i & 1] = f2_fma(v, v, qh[i & 1]);
                }
                const float4 gg = gbs[(q * 4 + i) * 4 + t];
                const uint64_t ga = f2_pack(gg.x, gg.y), be = f2_pack(gg.z, gg.w);
                y[2 * i] = f2_fma(f2_mul(f2_pack(d[0], d[1]), rsg), ga, be);
                y[2 * i + 1] = f2_fma(f2_mul(f2_pack(d[2], d[3]), rsh), ga, be);
            }
            gelu_erf_pairs<8>(y);
            uint32_t hg[4], hh[4];
            float yg[8], yh[8];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                f2_unpack(y[2 * i], yg[2 * i], yg[2 * i + 1]);
                f2_unpack(y[2 * i + 1], yh[2 * i], yh[2 * i + 1]);
                hg[i] = pack_bf16(yg[2 * i], yg[2 * i + 1]);
                hh[i] = pack_bf16(yh[2 * i], yh[2 * i + 1]);
            }
            if (vg) *reinterpret_cast<uint4*>(og + q * 32) = make_uint4(hg[0], hg[1], hg[2], hg[3]);
            if (vh) *reinterpret_cast<uint4*>(og + 8 * 512 + q * 32) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
            if (LO) {
                auto lo2 = [](float y0, float y1, uint32_t h) -> uint32_t {
                    const __nv_bfloat162 b2 = *reinterpret_cast<const __nv_bfloat162*>(&h);
                    return pack_bf16(y0 - __low2float(b2), y1 - __high2float(b2));
                };
                uint32_t lg[4], lh[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) { lg[i] = lo2(yg[2 * i], yg[2 * i + 1], hg[i]); lh[i] = lo2(yh[2 * i], yh[2 * i + 1], hh[i]); }
                if (vg) *reinterpret_cast<uint4*>(og + plane_stride + q * 32) = make_uint4(lg[0], lg[1], lg[2], lg[3]);
                if (vh) *reinterpret_cast<uint4*>(og + plane_stride + 8 * 512 + q * 32) = make_uint4(lh[0], lh[1], lh[2], lh[3]);
            }
        }
        rg = rstd_of(qg[0], qg[1]);
        rh = rstd_of(qh[0], qh[1]);
        A0 = N0; A1 = N1;
    }
}

cudaError_t launch_conv0_pack_frag(const float* w, const float* bias, const float* gamma, const float* beta, void* frag, cudaStream_t s) {
    uint2* f = reinterpret_cast<uint2*>(frag);
    conv0_pack_frag_kernel<<<1, 512, 0, s>>>(w, bias, gamma, beta, f, reinterpret_cast<float4*>(f + C0M2_FRAG_U2));
    return cudaGetLastError();
}
size_t conv0_frag_bytes() { return (size_t)C0M2_FRAG_U2 * sizeof(uint2) + (size_t)C0M2_GB_F4 * sizeof(float4); }

cudaError_t launch_conv0(const float* wav, int B, int N, int L0, const float* w, const float* bias, const float* gamma,
                         const float* beta, bf16* out, long long plane_stride, cudaStream_t s, const void* frag, int num_sms) {
    static const bool mma_env = [] { const char* e = getenv("OCC_CONV0_MMA"); return !(e && e[0] == '0'); }();     // 0: the SIMT form, for A/B runs
    if (frag != nullptr && mma_env) {
        const long long items = (long long)((L0 + 15) / 16) * B;
        const int sms = num_sms > 0 ? num_sms : 148;
        long long blocks = (items + C0M2_WARPS - 1) / C0M2_WARPS;
        if (blocks > 2LL * sms) blocks = 2LL * sms;
        if (blocks < 1) blocks = 1;
        const uint2* f2 = reinterpret_cast<const uint2*>(frag);
        const float4* gb2 = reinterpret_cast<const float4*>(f2 + C0M2_FRAG_U2);
        const dim3 gr((unsigned)blocks), bl(C0M2_WARPS * 32);
        if (plane_stride != 0) return launch_pdl(conv0_mma2_kernel<true>, gr, bl, (size_t)0, s, wav, N, L0, B, f2, gb2, out, plane_stride);
        return launch_pdl(conv0_mma2_kernel<false>, gr, bl, (size_t)0, s, wav, N, L0, B, f2, gb2, out, plane_stride);
    }
    const long long groups = (long long)((L0 + C0_FR - 1) / C0_FR) * B;
    long long blocks = (groups + C0_WARPS - 1) / C0_WARPS;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    return launch_pdl(conv0_ln_gelu_kernel, dim3((unsigned)blocks), dim3(C0_WARPS * 32), (size_t)0, s, wav, N, L0, B, w, bias, gamma, beta, out, plane_stride);
}

// =================================================================================================
// rows_ln: one warp per row of width W (512 or 1024).
//   y = LN(x; g1, b1);  if GELU: y = gelu(y);  if g2: y = LN(y; g2, b2)
//   writes any of: out32 fp32, out16 bf16 (+lo plane).   Rows with (row % rows_per_batch) >= valid[b] -> zeros.
//   Residual form (transformer): x <- x + add (add16 bf16 or add32 fp32: the previous sub-block's GEMM output), the sum is
//   written back to xsum (the fp32 residual stream, in place) and then normalised.  g1 == nullptr: sum only.
template <int W, bool GELU>
__global__ void __launch_bounds__(256)
rows_ln_kernel(const float* x /* may alias xsum (in place) */, long long M, const float* __restrict__ g1, const float* __restrict__ b1,
               const float* __restrict__ g2, const float* __restrict__ b2, float* __restrict__ out32,
               bf16* __restrict__ out16, long long plane_stride, const bf16* __restrict__ add16,
               const float* __restrict__ add32, float* xsum, const bf16* __restrict__ x16) {
    constexpr int V = W / 128;   // float4 per lane
    pdl_wait();
    const int lane = threadIdx.x & 31;
    const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= M) return;
    float v[V * 4];
    if (x16 != nullptr) {        // bf16 input (bf16 mode: the conv GEMM writes its pre-LN output in bf16)
        const uint2* xr = reinterpret_cast<const uint2*>(x16 + row * W);
#pragma unroll
        for (int i = 0; i < V; ++i) {
            const uint2 t = xr[i * 32 + lane];
            const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&t.x);
            const __nv_bfloat162 c = *reinterpret_cast<const __nv_bfloat162*>(&t.y);
            v[4 * i] = __low2float(a); v[4 * i + 1] = __high2float(a); v[4 * i + 2] = __low2float(c); v[4 * i + 3] = __high2float(c);
        }
    } else {
        const float4* xr = reinterpret_cast<const float4*>(x + row * W);
#pragma unroll
        for (int i = 0; i < V; ++i) {
            const float4 t = xr[i * 32 + lane];
            v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
        }
    }
    if (add16 != nullptr || add32 != nullptr) {
        if (add16 != nullptr) {
            const uint2* ar = reinterpret_cast<const uint2*>(add16 + row * W);
#pragma unroll
            for (int i = 0; i < V; ++i) {
                const uint2 t = ar[i * 32 + lane];
                const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&t.x);
                const __nv_bfloat162 c = *reinterpret_cast<const __nv_bfloat162*>(&t.y);
                v[4 * i] += __low2float(a); v[4 * i + 1] += __high2float(a);
                v[4 * i + 2] += __low2float(c); v[4 * i + 3] += __high2float(c);
            }
        } else {
            const float4* ar = reinterpret_cast<const float4*>(add32 + row * W);
#pragma unroll
            for (int i = 0; i < V; ++i) {
                const float4 t = ar[i * 32 + lane];
                v[4 * i] += t.x; v[4 * i + 1] += t.y; v[4 * i + 2] += t.z; v[4 * i + 3] += t.w;
            }
        }
        float4* xo = reinterpret_cast<float4*>(xsum + row * W);
#pragma unroll
        for (int i = 0; i < V; ++i) xo[i * 32 + lane] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        if (g1 == nullptr) return;
    }
    auto norm = [&](const float* __restrict__ g, const float* __restrict__ b) {      // packed fp32x2 arithmetic (FADD2/FFMA2/FMUL2)
        uint64_t s2 = f2_pack(v[0], v[1]);
#pragma unroll
        for (int i = 1; i < V * 2; ++i) s2 = f2_add(s2, f2_pack(v[2 * i], v[2 * i + 1]));
        float sa, sb;
        f2_unpack(s2, sa, sb);
        const float mean = warp_sum(sa + sb) * (1.0f / W);
        const uint64_t nm = f2_splat(-mean);
        uint64_t q2 = f2_splat(0.f);
#pragma unroll
        for (int i = 0; i < V * 2; ++i) { const uint64_t d = f2_add(f2_pack(v[2 * i], v[2 * i + 1]), nm); q2 = f2_fma(d, d, q2); }
        float qa, qb;
        f2_unpack(q2, qa, qb);
        const float rstd = rsqrtf(warp_sum(qa + qb) * (1.0f / W) + 1e-5f);
        const uint64_t rs = f2_splat(rstd);
#pragma unroll
        for (int i = 0; i < V; ++i) {
            const float4 gg = __ldg(reinterpret_cast<const float4*>(g) + i * 32 + lane);
            const float4 bb = __ldg(reinterpret_cast<const float4*>(b) + i * 32 + lane);
            f2_unpack(f2_fma(f2_mul(f2_add(f2_pack(v[4 * i], v[4 * i + 1]), nm), rs), f2_pack(gg.x, gg.y), f2_pack(bb.x, bb.y)), v[4 * i], v[4 * i + 1]);
            f2_unpack(f2_fma(f2_mul(f2_add(f2_pack(v[4 * i + 2], v[4 * i + 3]), nm), rs), f2_pack(gg.z, gg.w), f2_pack(bb.z, bb.w)), v[4 * i + 2], v[4 * i + 3]);
        }
    };
    norm(g1, b1);
    if (GELU) {
#pragma unroll
        for (int i = 0; i < V * 4; i += 2) gelu_erf_fast2(v[i], v[i + 1]);
    }
    if (g2 != nullptr) norm(g2, b2);
    if (out32 != nullptr) {
        float4* o = reinterpret_cast<float4*>(out32 + row * W);
#pragma unroll
        for (int i = 0; i < V; ++i) o[i * 32 + lane] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    }
    if (out16 != nullptr) {
        uint2* o = reinterpret_cast<uint2*>(out16 + row * W);
#pragma unroll
        for (int i = 0; i < V; ++i) {
            const uint32_t h0 = pack_bf16(v[4 * i], v[4 * i + 1]), h1 = pack_bf16(v[4 * i + 2], v[4 * i + 3]);
            o[i * 32 + lane] = make_uint2(h0, h1);
            if (plane_stride != 0) {
                const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&h0);
                const __nv_bfloat162 c = *reinterpret_cast<const __nv_bfloat162*>(&h1);
                uint2* lo = reinterpret_cast<uint2*>(out16 + plane_stride + row * W);
                lo[i * 32 + lane] = make_uint2(pack_bf16(v[4 * i] - __low2float(a), v[4 * i + 1] - __high2float(a)),
                                               pack_bf16(v[4 * i + 2] - __low2float(c), v[4 * i + 3] - __high2float(c)));
            }
        }
    }
}

cudaError_t launch_rows_ln(const float* x, long long M, int W, bool gelu, const float* g1, const float* b1, const float* g2,
                           const float* b2, float* out32, bf16* out16, long long plane_stride, cudaStream_t s,
                           const bf16* add16, const float* add32, float* xsum, const bf16* x16) {
    if (M <= 0) return cudaSuccess;
    const int wpb = 8;
    const unsigned blocks = (unsigned)((M + wpb - 1) / wpb);
    if ((add16 || add32) && !xsum) return cudaErrorInvalidValue;
    if (W == 512 && gelu) (void)launch_pdl(rows_ln_kernel<512, true>, dim3((unsigned)blocks), dim3((unsigned)(wpb * 32)), (size_t)0, s, x, M, g1, b1, g2, b2, out32, out16, plane_stride, add16, add32, xsum, x16);
    else if (W == 512) (void)launch_pdl(rows_ln_kernel<512, false>, dim3((unsigned)blocks), dim3((unsigned)(wpb * 32)), (size_t)0, s, x, M, g1, b1, g2, b2, out32, out16, plane_stride, add16, add32, xsum, x16);
    else if (W == 1024 && !gelu) (void)launch_pdl(rows_ln_kernel<1024, false>, dim3((unsigned)blocks), dim3((unsigned)(wpb * 32)), (size_t)0, s, x, M, g1, b1, g2, b2, out32, out16, plane_stride, add16, add32, xsum, x16);
    else return cudaErrorInvalidValue;
    return cudaGetLastError();
}

// =================================================================================================
// convert fp32 -> bf16 hi (+lo) planes, elementwise (used to stage fp32 inputs of the stage entry points).
__global__ void f32_to_bf16_kernel(const float* __restrict__ x, long long n, bf16* __restrict__ hi, long long plane_stride) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    bf16 h, l;
    split_bf16(x[i], h, l);
    hi[i] = h;
    if (plane_stride != 0) hi[plane_stride + i] = l;
}
cudaError_t launch_f32_to_bf16(const float* x, long long n, bf16* hi, long long plane_stride, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    f32_to_bf16_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(x, n, hi, plane_stride);
    return cudaGetLastError();
}

__global__ void bf16_planes_to_f32_kernel(const bf16* __restrict__ hi, long long plane_stride, long long n, float* __restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float v = __bfloat162float(hi[i]);
    if (plane_stride != 0) v += __bfloat162float(hi[plane_stride + i]);
    out[i] = v;
}
cudaError_t launch_bf16_planes_to_f32(const bf16* hi, long long plane_stride, long long n, float* out, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    bf16_planes_to_f32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(hi, plane_stride, n, out);
    return cudaGetLastError();
}

// =================================================================================================
// attn_bf16: qkv bf16 [B*T, 3072] (q pre-scaled by 1/8 through the weights) -> ctx bf16 [B*T, 1024].
// grid (ceil(T/128), 16, B); 8 warps x 16 query rows; keys in chunks of 64, double-buffered cp.async.
static constexpr int AT_LD = 72;   // padded row (bf16) -> conflict-free ldmatrix

__device__ __forceinline__ void cp_async16(void* dst, const void* src, bool pred) {
    const uint32_t d = smem_u32(dst);
    const int sz = pred ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void ldsm_x4(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, const void* p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(smem_u32(p)) : "memory");
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, const void* p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(smem_u32(p)) : "memory");
}
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

static constexpr int AT_WARPS = 8;                 // 16 query rows per warp -> 128-row query tile
static constexpr int AT_BQ = AT_WARPS * 16;
static constexpr int AT_SMEM = (AT_BQ + 4 * 64) * AT_LD * 2;

__device__ __forceinline__ float ex2_fast(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// Work that is pure padding is skipped: warps whose 16 query rows lie beyond the utterance only help with the loads, and in
// the last key chunk only the 8-key score tiles / 16-key PV steps that contain a valid key are issued (T = 201 -> 13/16 of
// the query tiles x 13/16 of the key tiles of the padded 256 x 256 problem).
__global__ void __launch_bounds__(AT_WARPS * 32)
attn_bf16_kernel(const bf16* __restrict__ qkv, int T, const int* __restrict__ klen, bf16* __restrict__ ctx) {
    extern __shared__ __align__(16) uint8_t at_smem[];
    bf16* sQ = reinterpret_cast<bf16*>(at_smem);            // [AT_BQ][AT_LD]
    bf16* sKb = sQ + AT_BQ * AT_LD;                         // [2][64][AT_LD]
    bf16* sVb = sKb + 2 * 64 * AT_LD;                       // [2][64][AT_LD]
    const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int kl = klen ? klen[b] : T;
    if (qt * AT_BQ >= kl) return;                    // whole query tile is padding
    const long long row0 = (long long)b * T;
    const bf16* qbase = qkv + row0 * 3072 + h * 64;
    const bf16* kbase = qbase + 1024;
    const bf16* vbase = qbase + 2048;
    const int nchunks = (kl + 63) / 64;
    const bool active = qt * AT_BQ + warp * 16 < kl; // this warp owns at least one real query row

    auto load_tile = [&](bf16* dst, const bf16* src, int r0, int rows) {
        // rows x 8 chunks of 16 B, zero-filled beyond the utterance
        for (int idx = tid; idx < rows * 8; idx += AT_WARPS * 32) {
            const int r = idx >> 3, c = idx & 7;
            const bool ok = (r0 + r) < kl;
            cp_async16(dst + r * AT_LD + c * 8, src + (long long)(ok ? (r0 + r) : 0) * 3072 + c * 8, ok);
        }
    };
    load_tile(sQ, qbase, qt * AT_BQ, AT_BQ);
    load_tile(sKb, kbase, 0, 64);
    load_tile(sVb, vbase, 0, 64);
    cp_async_commit();

    uint32_t qf[4][4];
    float o[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) { o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f; }
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;     // m*: running row maxima, already multiplied by log2(e)
    const float LOG2E = 1.4426950408889634f;

    for (int kc = 0; kc < nchunks; ++kc) {
        const int buf = kc & 1;
        const bf16* sK = sKb + buf * 64 * AT_LD;
        const bf16* sV = sVb + buf * 64 * AT_LD;
        if (kc + 1 < nchunks) {
            load_tile(sKb + (buf ^ 1) * 64 * AT_LD, kbase, (kc + 1) * 64, 64);
            load_tile(sVb + (buf ^ 1) * 64 * AT_LD, vbase, (kc + 1) * 64, 64);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        if (active) {
            if (kc == 0) {
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {
                    const bf16* p = sQ + (warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * AT_LD + ks * 16 + (lane >> 4) * 8;
                    ldsm_x4(qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3], p);
                }
            }
            const int nvalid = min(64, kl - kc * 64);
            const int nj = (nvalid + 7) >> 3;            // 8-key score tiles holding a valid key
            const int nks = (nvalid + 15) >> 4;          // 16-key PV steps holding a valid key
            // ---- S = Q K^T (16 x 64 per warp)
            float s[8][4];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
                if (j < nj) {
#pragma unroll
                    for (int i = 0; i < 2; ++i) {
                        uint32_t b0, b1, b2, b3;
                        const bf16* p = sK + (j * 8 + (lane & 7)) * AT_LD + i * 32 + (lane >> 3) * 8;
                        ldsm_x4(b0, b1, b2, b3, p);
                        mma_bf16_16816(s[j], qf[2 * i], b0, b1);
                        mma_bf16_16816(s[j], qf[2 * i + 1], b2, b3);
                    }
                }
            }
            // ---- mask keys >= kl, online softmax (rows g = lane/4 and g+8)
            const int kbase_idx = kc * 64 + 2 * (lane & 3);
            float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int k0 = kbase_idx + j * 8;
                if (k0 >= kl) { s[j][0] = -INFINITY; s[j][2] = -INFINITY; }
                if (k0 + 1 >= kl) { s[j][1] = -INFINITY; s[j][3] = -INFINITY; }
                mx0 = fmaxf(mx0, fmaxf(s[j][0], s[j][1]));
                mx1 = fmaxf(mx1, fmaxf(s[j][2], s[j][3]));
            }
            mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
            mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
            const float mn0 = fmaxf(m0, mx0 * LOG2E), mn1 = fmaxf(m1, mx1 * LOG2E);   // finite: every chunk has >= 1 valid key
            const float sc0 = ex2_fast(m0 - mn0), sc1 = ex2_fast(m1 - mn1);
            m0 = mn0; m1 = mn1;
            l0 *= sc0; l1 *= sc1;
#pragma unroll
            for (int j = 0; j < 8; ++j) { o[j][0] *= sc0; o[j][1] *= sc0; o[j][2] *= sc1; o[j][3] *= sc1; }
            uint32_t pf[4][4];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float p0 = ex2_fast(fmaf(s[j][0], LOG2E, -mn0)), p1 = ex2_fast(fmaf(s[j][1], LOG2E, -mn0));
                const float p2 = ex2_fast(fmaf(s[j][2], LOG2E, -mn1)), p3 = ex2_fast(fmaf(s[j][3], LOG2E, -mn1));
                l0 += p0 + p1; l1 += p2 + p3;
                pf[j >> 1][(j & 1) * 2 + 0] = pack_bf16(p0, p1);
                pf[j >> 1][(j & 1) * 2 + 1] = pack_bf16(p2, p3);
            }
            // ---- O += P V
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                if (ks < nks) {
#pragma unroll
                    for (int jd = 0; jd < 8; jd += 2) {
                        uint32_t b0, b1, b2, b3;
                        const bf16* p = sV + (ks * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * AT_LD + jd * 8 + (lane >> 4) * 8;
                        ldsm_x4_t(b0, b1, b2, b3, p);
                        mma_bf16_16816(o[jd], pf[ks], b0, b1);
                        mma_bf16_16816(o[jd + 1], pf[ks], b2, b3);
                    }
                }
            }
        }
        __syncthreads();   // all warps done with buf before it is refilled
    }
    if (!active) return;
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = 1.0f / l0, i1 = 1.0f / l1;
    const int r0 = qt * AT_BQ + warp * 16 + (lane >> 2);
    const int r1 = r0 + 8;
#pragma unroll
    for (int jd = 0; jd < 8; ++jd) {
        const int d = h * 64 + jd * 8 + 2 * (lane & 3);
        if (r0 < kl) *reinterpret_cast<uint32_t*>(ctx + (row0 + r0) * 1024 + d) = pack_bf16(o[jd][0] * i0, o[jd][1] * i0);
        if (r1 < kl) *reinterpret_cast<uint32_t*>(ctx + (row0 + r1) * 1024 + d) = pack_bf16(o[jd][2] * i1, o[jd][3] * i1);
    }
}

cudaError_t launch_attn_tc2(const bf16* qkv, int B, int T, const int* klen, bf16* ctx, int num_sms, cudaStream_t s, std::string* err);   // attn_tc2.cu
cudaError_t launch_attn_tc3(const bf16* qkv, int B, int T, const int* klen, bf16* ctx, int num_sms, cudaStream_t s, std::string* err);   // attn_tc3.cu

cudaError_t launch_attn_bf16(const bf16* qkv, int B, int T, const int* klen, bf16* ctx, int num_sms, cudaStream_t s) {
    // T <= 256: whole score rows in TMEM / registers (attn_tc2.cu); longer utterances: key chunks of 128 with an online softmax,
    // still tcgen05 (attn_tc3.cu).  OCC_ATTN_LEGACY=1 selects the mma.sync kernel above for A/B runs; it is not a product path.
    static const bool legacy = [] { const char* e = getenv("OCC_ATTN_LEGACY"); return e && e[0] == '1'; }();
    if (!legacy) {
        if (T <= 256) return launch_attn_tc2(qkv, B, T, klen, ctx, num_sms, s, nullptr);
        return launch_attn_tc3(qkv, B, T, klen, ctx, num_sms, s, nullptr);
    }
    {
        cudaError_t e = ensure_dyn_smem(reinterpret_cast<const void*>(attn_bf16_kernel), AT_SMEM);
        if (e != cudaSuccess) return e;
    }
    dim3 grid((T + AT_BQ - 1) / AT_BQ, 16, B);
    attn_bf16_kernel<<<grid, AT_WARPS * 32, AT_SMEM, s>>>(qkv, T, klen, ctx);
    return cudaGetLastError();
}

// =================================================================================================
// attn_f32: qkv fp32 [B*T, 3072] -> ctx bf16 hi/lo planes [B*T, 1024].  One warp per query row,
// keys in chunks of 64 held in shared memory (padded rows), online softmax in fp32.
static constexpr int AF_KC = 64;
static constexpr int AF_WARPS = 8;

__global__ void __launch_bounds__(AF_WARPS * 32)
attn_f32_kernel(const float* __restrict__ qkv, int T, const int* __restrict__ klen, bf16* __restrict__ ctx, long long plane_stride) {
    __shared__ float sK[AF_KC][65];
    __shared__ float sV[AF_KC][64];
    __shared__ float sQ[AF_WARPS][64];
    __shared__ float sP[AF_WARPS][AF_KC];
    const int h = blockIdx.y, b = blockIdx.z;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int kl = klen ? klen[b] : T;
    const int q = blockIdx.x * AF_WARPS + warp;
    const long long row0 = (long long)b * T;
    const bool active = q < kl;
    if (blockIdx.x * AF_WARPS >= kl) return;
    if (active) {
        const float* qp = qkv + (row0 + q) * 3072 + h * 64;
        sQ[warp][lane] = qp[lane];
        sQ[warp][lane + 32] = qp[lane + 32];
    }
    float m = -INFINITY, l = 0.f, o0 = 0.f, o1 = 0.f;
    for (int k0 = 0; k0 < kl; k0 += AF_KC) {
        __syncthreads();
        for (int idx = threadIdx.x; idx < AF_KC * 64; idx += blockDim.x) {
            const int r = idx >> 6, c = idx & 63;
            const bool ok = (k0 + r) < kl;
            const float* base = qkv + (row0 + (ok ? k0 + r : 0)) * 3072 + h * 64 + c;
            sK[r][c] = ok ? base[1024] : 0.f;
            sV[r][c] = ok ? base[2048] : 0.f;
        }
        __syncthreads();
        if (active) {
            float s0 = 0.f, s1 = 0.f;
#pragma unroll 16
            for (int d = 0; d < 64; ++d) {
                const float qd = sQ[warp][d];
                s0 = fmaf(qd, sK[lane][d], s0);
                s1 = fmaf(qd, sK[lane + 32][d], s1);
            }
            if (k0 + lane >= kl) s0 = -INFINITY;
            if (k0 + lane + 32 >= kl) s1 = -INFINITY;
            const float mx = warp_max(fmaxf(s0, s1));
            const float mn = fmaxf(m, mx);
            const float sc = expf(m - mn);
            const float p0 = expf(s0 - mn), p1 = expf(s1 - mn);
            l = l * sc + warp_sum(p0 + p1);
            o0 *= sc; o1 *= sc;
            m = mn;
            sP[warp][lane] = p0;
            sP[warp][lane + 32] = p1;
            __syncwarp();
#pragma unroll 16
            for (int k = 0; k < AF_KC; ++k) {
                const float pk = sP[warp][k];
                o0 = fmaf(pk, sV[k][lane], o0);
                o1 = fmaf(pk, sV[k][lane + 32], o1);
            }
            __syncwarp();
        }
    }
    if (active) {
        const float inv = 1.0f / l;
        bf16* op = ctx + (row0 + q) * 1024 + h * 64;
        bf16 hi, lo;
        split_bf16(o0 * inv, hi, lo);
        op[lane] = hi;
        if (plane_stride) op[plane_stride + lane] = lo;
        split_bf16(o1 * inv, hi, lo);
        op[lane + 32] = hi;
        if (plane_stride) op[plane_stride + lane + 32] = lo;
    }
}

cudaError_t launch_attn_f32(const float* qkv, int B, int T, const int* klen, bf16* ctx, long long plane_stride, cudaStream_t s) {
    dim3 grid((T + AF_WARPS - 1) / AF_WARPS, 16, B);
    attn_f32_kernel<<<grid, AF_WARPS * 32, 0, s>>>(qkv, T, klen, ctx, plane_stride);
    return cudaGetLastError();
}

}  // namespace occ
