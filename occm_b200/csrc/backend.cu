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
    pdl_wait();       // launched with the PDL attribute: nothing above this line touches global memory
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
    (void)launch_pdl(ll_pool_kernel, dim3(grid), dim3((unsigned)(128)), (size_t)(0), s, ll, T, Wm, wlen, scale, shift, x0);
    return cudaGetLastError();
}

// =================================================================================================
// Residual blocks (models/sslassist.py:373-429).  The 11 convolutions with >= 32 input channels run as tcgen05 shift-GEMMs
// (gemm_tc.cu, GemmDesc::conv_mode) over channel-last, zero-padded bf16 hi/lo activations [B][Hp][Wp][64]; only the two
// single-input-channel convolutions of block 0 are computed here:
//   A = selu(bn2(conv1(x0)))  conv1 1->32, (2,3), pad (1,1): 43 rows   (bn1 + selu on the input are dead code in the reference,
//   I = conv_downsample(x0)   1->32, (1,3), pad (0,1): 42 rows          sslassist.py:408-415: conv1 sees the raw block input)
// One thread per padded-grid position; the border, columns >= wl and channels 32..63 are written as zeros.
__global__ void __launch_bounds__(128)
res0_front_kernel(const float* __restrict__ x0, int Wm, const int* __restrict__ wlen, const float* __restrict__ w1,
                  const float* __restrict__ b1, const float* __restrict__ s2, const float* __restrict__ t2,
                  const float* __restrict__ wd, const float* __restrict__ bd, int Hp, bf16* __restrict__ A,
                  bf16* __restrict__ I, long long plane) {
    pdl_wait();       // launched with the PDL attribute: nothing above this line touches global memory
    __shared__ float sw1[32 * 6], swd[32 * 3], sb1[32], ss2[32], st2[32], sbd[32];
    for (int i = threadIdx.x; i < 32 * 6; i += blockDim.x) sw1[i] = w1[i];
    for (int i = threadIdx.x; i < 32 * 3; i += blockDim.x) swd[i] = wd[i];
    if (threadIdx.x < 32) { sb1[threadIdx.x] = b1[threadIdx.x]; ss2[threadIdx.x] = s2[threadIdx.x]; st2[threadIdx.x] = t2[threadIdx.x]; sbd[threadIdx.x] = bd[threadIdx.x]; }
    __syncthreads();
    const int b = blockIdx.y, Wp = Wm + 2;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;       // (positions >= Hp * Wp of the last block compute zeros and store nothing)
    const int hp = t / Wp, wp = t - hp * Wp;
    const int wl = wlen ? wlen[b] : Wm;
    const bool col_ok = wp >= 1 && wp <= wl;
    const bool a_ok = col_ok && hp >= 1 && hp < 44, i_ok = col_ok && hp >= 1 && hp < 43;
    // x0 window: rows ho-1, ho (ho = hp-1), columns wo-1..wo+1 (wo = wp-1); zero outside [0,42) x [0,wl)
    float xv[2][3];
    const float* xb = x0 + (long long)b * 42 * Wm;
#pragma unroll
    for (int kh = 0; kh < 2; ++kh)
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) {
            const int hi = hp - 2 + kh, wi = wp - 2 + kw;
            xv[kh][kw] = (hi >= 0 && hi < 42 && wi >= 0 && wi < wl) ? xb[hi * Wm + wi] : 0.f;
        }
    uint32_t ah[16], al[16], ih[16], il[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) {
        float va[2], vi[2];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const int co = 2 * q + e;
            float a = sb1[co];
#pragma unroll
            for (int kh = 0; kh < 2; ++kh)
#pragma unroll
                for (int kw = 0; kw < 3; ++kw) a = fmaf(sw1[co * 6 + kh * 3 + kw], xv[kh][kw], a);
            va[e] = a_ok ? selu_f(fmaf(a, ss2[co], st2[co])) : 0.f;
            float d = sbd[co];
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) d = fmaf(swd[co * 3 + kw], xv[1][kw], d);     // row h = hp-1 is window row kh = 1
            vi[e] = i_ok ? d : 0.f;
        }
        ah[q] = pack_bf16(va[0], va[1]);
        ih[q] = pack_bf16(vi[0], vi[1]);
        const __nv_bfloat162 x = *reinterpret_cast<const __nv_bfloat162*>(&ah[q]);
        const __nv_bfloat162 y = *reinterpret_cast<const __nv_bfloat162*>(&ih[q]);
        al[q] = pack_bf16(va[0] - __low2float(x), va[1] - __high2float(x));
        il[q] = pack_bf16(vi[0] - __low2float(y), vi[1] - __high2float(y));
    }
    // The 32 positions of a warp are one contiguous 4 KB block per tensor and plane: each 128 B row is staged in shared memory
    // (16 B pieces XOR-swizzled by row) and the block goes out as 8 fully coalesced 512 B store instructions (the thread = row
    // stores touched 32 lines per instruction: 67 us at B = 64).
    __shared__ uint4 stg[4][32 * 8];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint4* sw = stg[warp];
    const int t0w = t - lane;                            // first position of this warp
    const int nvalid = Hp * Wp - t0w;                    // rows r < nvalid exist
    const long long ow = ((long long)b * Hp * Wp + t0w) * 64;
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    auto flush = [&](bf16* dst, const uint32_t* v) {
#pragma unroll
        for (int q = 0; q < 4; ++q) sw[lane * 8 + (q ^ (lane & 7))] = make_uint4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
#pragma unroll
        for (int q = 4; q < 8; ++q) sw[lane * 8 + (q ^ (lane & 7))] = z;          // channels 32..63
        __syncwarp();
        uint4* gp = reinterpret_cast<uint4*>(dst + ow);
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int idx = it * 32 + lane, r = idx >> 3, q = idx & 7;
            if (r < nvalid) gp[idx] = sw[r * 8 + (q ^ (r & 7))];
        }
        __syncwarp();
    };
    flush(A, ah); flush(A + plane, al); flush(I, ih); flush(I + plane, il);
}
cudaError_t launch_res0_front(const float* x0, int B, int Wm, const int* wlen, const float* w1, const float* b1, const float* s2,
                              const float* t2, const float* wd, const float* bd, int Hp, bf16* A, bf16* I, long long plane, cudaStream_t s) {
    dim3 grid((Hp * (Wm + 2) + 127) / 128, B);
    (void)launch_pdl(res0_front_kernel, dim3(grid), dim3((unsigned)(128)), (size_t)(0), s, x0, Wm, wlen, w1, b1, s2, t2, wd, bd, Hp, A, I, plane);
    return cudaGetLastError();
}

// attention map and spectro-temporal pooling: backend_attmap.cu

// =================================================================================================
// Graph attention (homogeneous and heterogeneous) -- one CTA computes GA_ROWS complete output rows.
//   att[i][j] = sum_o aw_{type(i,j)}[o] * tanh( bp[o] + sum_d Wp[o][d] x_i[d] x_j[d] ) / temp
//   a = softmax_j(att[i][:]);  y_i = SELU(BN( Wa (sum_j a_j x_j) + ba + Wo x_i + bo ))
// Heterogeneous layers (n1 nodes of type 1 then n2 of type 2) pick aw11 / aw22 / aw12 per block
// (sslassist.py:284-291; aw12 serves both off-diagonal blocks).  blockIdx.x == gridDim.x-1 with
// `master` set computes the master-node update instead (sslassist.py:234-239, 255-270, 310-316).
// One CTA per (utterance, graph): x_i * x_j is symmetric and the off-diagonal type blocks share aw12, so att[i][j] ==
// att[j][i] and only the n(n+1)/2 pairs i <= j are evaluated (the row-block form recomputed every pair twice and reloaded
// x and Wp nine times per graph).  The whole n x n logit matrix stays in shared memory for the row softmax.
static constexpr int GA_THREADS = 512;

__device__ __forceinline__ float tanh_fast(float t) {
    // 1 - 2 / (1 + e^{2t}) with MUFU exp/rcp: |abs error| ~ 2e-7 (the logits sum 32-64 of these against O(0.1) weights)
    const float e = __expf(2.0f * fabsf(t));
    const float r = 1.0f - __fdividef(2.0f, 1.0f + e);
    return copysignf(r, t);
}

__device__ __forceinline__ void ga_mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// (x, y) -> packed bf16 pair of the high parts and of the residuals
__device__ __forceinline__ void ga_split_pack(float x, float y, uint32_t& hi, uint32_t& lo) {
    hi = pack_bf16(x, y);
    const __nv_bfloat162 h2 = *reinterpret_cast<const __nv_bfloat162*>(&hi);
    lo = pack_bf16(x - __low2float(h2), y - __high2float(h2));
}

template <int DIN, int DOUT>
__global__ void __launch_bounds__(GA_THREADS, DOUT == 32 ? 2 : 1)      // HS-GAL launches: 256 CTAs (graph + master), two per SM = one wave
gat_kernel(const GatArgs a) {
    pdl_wait();       // launched with the PDL attribute: nothing above this line touches global memory
    extern __shared__ __align__(16) float sm[];
    constexpr int LDX = DIN + 4;                 // padded node row: conflict-free 16 B reads of distinct rows
    constexpr int LDW = DIN + 1;                 // padded weight row for the lane = output-channel reads of the projection
    const int z = blockIdx.z, b = blockIdx.y;
    const GatW& W = a.w[z];
    const int nmax = a.nmax[z];
    const int n1 = a.n1[z] ? a.n1[z][b] : a.n1c[z];
    const int n2 = a.n2[z] ? a.n2[z][b] : a.n2c[z];
    const int n = n1 + n2;
    const int lda = nmax + 1;
    float* sx = sm;                              // [nmax][LDX]
    float* sWp = sx + nmax * LDX;                // [DOUT][DIN]  (later: Wa [DOUT][LDW])
    float* sWo = sWp + DOUT * LDW;               // [DOUT][LDW]
    float* satt = sWo + DOUT * LDW;              // [nmax][lda]
    float* sagg = satt + ((nmax * lda + 3) & ~3);    // [nmax][DIN], 16 B aligned (float4 rows; the launch adds 16 B of slack)
    unsigned short* spair = reinterpret_cast<unsigned short*>(sagg + nmax * DIN);   // [nmax (nmax+1) / 2]
    const float* xb = a.x[z] + (long long)b * nmax * DIN;
    const int tid = threadIdx.x;
    const bool master_block = blockIdx.x == 1;
    if (n <= 0) return;
    for (int i = tid; i < n * (DIN / 4); i += GA_THREADS) {
        const int r = i / (DIN / 4), q = i - r * (DIN / 4);
        *reinterpret_cast<float4*>(sx + r * LDX + 4 * q) = *reinterpret_cast<const float4*>(xb + r * DIN + 4 * q);
    }
    // pair blocks: Wp as bf16 hi / lo planes [DOUT][PW] (B operand of the tensor-core logits; spills into the sWo region, which is
    // only filled after the pair phase); master block: fp32 WpM
    constexpr int PW = DIN + 8;                  // bf16 row pitch: (PW / 2) words -> bank = 4 g + q (DIN 64) / 20 g + q (DIN 32): conflict-free
    bf16* sWh = reinterpret_cast<bf16*>(sWp);
    bf16* sWl = sWh + DOUT * PW;
    static_assert(2 * DOUT * PW * 2 <= 2 * DOUT * LDW * 4, "bf16 Wp planes must fit in the sWp + sWo regions");
    static_assert(DOUT <= DIN, "master block keeps its n x DOUT tanh values in the sagg region");
    if (master_block) {
        for (int i = tid; i < DOUT * DIN; i += GA_THREADS) sWp[(i / DIN) * LDW + (i % DIN)] = __ldg(W.WpM + i);     // padded rows: lane = o reads
    } else {
        for (int i = tid; i < DOUT * DIN; i += GA_THREADS) {
            bf16 hi, lo;
            split_bf16(__ldg(W.Wp + i), hi, lo);
            const int o = i / DIN, d = i - o * DIN;
            sWh[o * PW + d] = hi; sWl[o * PW + d] = lo;
        }
    }
    if (!master_block) {
        for (int i = tid; i < n; i += GA_THREADS) {
            const int base = i * n - (i * (i - 1)) / 2;          // pairs (i, i..n-1) start here
            for (int j = i; j < n; ++j) spair[base + j - i] = (unsigned short)((i << 8) | j);
        }
    }
    __syncthreads();

    if (master_block) {
        // ---- master node: att_j = awM . tanh(WpM (x_j * m) + bpM) / temp ; softmax over j
        const float* mi = W.master_in + (long long)b * W.master_stride;
        // thread = (node j, output o): the DIN-long chain of one output; then thread = node sums its DOUT terms in order (the
        // chains and their order are those of the former thread = node loop, which kept this CTA alive for ~10 us on 54 threads)
        for (int e = tid; e < n * DOUT; e += GA_THREADS) {
            const int j = e / DOUT, o = e - j * DOUT;
            float t = __ldg(W.bpM + o);
            const float* wr = sWp + o * LDW;
            const float* xr = sx + j * LDX;
#pragma unroll 8
            for (int d = 0; d < DIN; ++d) t = fmaf(wr[d], xr[d] * __ldg(mi + d), t);
            sagg[e] = tanhf(t);
        }
        __syncthreads();
        for (int j = tid; j < n; j += GA_THREADS) {
            float acc = 0.f;
            for (int o = 0; o < DOUT; ++o) acc = fmaf(__ldg(W.awM + o), sagg[j * DOUT + o], acc);
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
            for (int j = 0; j < n; ++j) acc = fmaf(satt[j], sx[j * LDX + tid], acc);
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

    // ---- pairwise attention logits, upper triangle.  Per warp a tile of 16 pairs: Z = x_i * x_j [16 x DIN] is formed in the A-fragment
    // layout (fp32 product, split into bf16 hi + lo), T = Z Wp^T [16 x DOUT] on mma.sync m16n8k16 as hi*hi + hi*lo + lo*hi with fp32
    // accumulation (~2^-17 relative, the scheme of the attention map and the residual-encoder GEMMs), then aw . tanh(T + bp) per row.
    // (The SIMT form -- thread = pair, 2 * DIN * DOUT FMAs against broadcast LDS -- ran at 40 % issue with 4 warps per scheduler:
    // 129 us for the two 64-wide graphs at B = 64, profiles/r1l.)
    const int npairs = n * (n + 1) / 2;
    const float inv_temp = 1.0f / W.temp;
    {
        const int warp = tid >> 5, lane = tid & 31, g = lane >> 2, q = lane & 3;
        const int ntile = (npairs + 15) >> 4;
        for (int tile = warp; tile < ntile; tile += GA_THREADS / 32) {
            const int pA = tile * 16 + g, pB = pA + 8;
            const int ijA = spair[min(pA, npairs - 1)], ijB = spair[min(pB, npairs - 1)];
            const int iA = ijA >> 8, jA = ijA & 255, iB = ijB >> 8, jB = ijB & 255;
            const float* xiA = sx + iA * LDX + 2 * q; const float* xjA = sx + jA * LDX + 2 * q;
            const float* xiB = sx + iB * LDX + 2 * q; const float* xjB = sx + jB * LDX + 2 * q;
            uint32_t ah[DIN / 16][4], al[DIN / 16][4];
#pragma unroll
            for (int ks = 0; ks < DIN / 16; ++ks) {
                const float2 u0 = *reinterpret_cast<const float2*>(xiA + ks * 16), v0 = *reinterpret_cast<const float2*>(xjA + ks * 16);
                const float2 u1 = *reinterpret_cast<const float2*>(xiB + ks * 16), v1 = *reinterpret_cast<const float2*>(xjB + ks * 16);
                const float2 u2 = *reinterpret_cast<const float2*>(xiA + ks * 16 + 8), v2 = *reinterpret_cast<const float2*>(xjA + ks * 16 + 8);
                const float2 u3 = *reinterpret_cast<const float2*>(xiB + ks * 16 + 8), v3 = *reinterpret_cast<const float2*>(xjB + ks * 16 + 8);
                ga_split_pack(u0.x * v0.x, u0.y * v0.y, ah[ks][0], al[ks][0]);
                ga_split_pack(u1.x * v1.x, u1.y * v1.y, ah[ks][1], al[ks][1]);
                ga_split_pack(u2.x * v2.x, u2.y * v2.y, ah[ks][2], al[ks][2]);
                ga_split_pack(u3.x * v3.x, u3.y * v3.y, ah[ks][3], al[ks][3]);
            }
            const float* awA = (iA < n1) ? ((jA < n1) ? W.aw11 : W.aw12) : ((jA < n1) ? W.aw12 : W.aw22);
            const float* awB = (iB < n1) ? ((jB < n1) ? W.aw11 : W.aw12) : ((jB < n1) ? W.aw12 : W.aw22);
            float accA = 0.f, accB = 0.f;
#pragma unroll 2
            for (int nt = 0; nt < DOUT / 8; ++nt) {
                float c[4] = {0.f, 0.f, 0.f, 0.f};
                const bf16* wh = sWh + (nt * 8 + g) * PW + 2 * q;
                const bf16* wl = sWl + (nt * 8 + g) * PW + 2 * q;
#pragma unroll
                for (int ks = 0; ks < DIN / 16; ++ks) {
                    const uint32_t bh0 = *reinterpret_cast<const uint32_t*>(wh + ks * 16), bh1 = *reinterpret_cast<const uint32_t*>(wh + ks * 16 + 8);
                    const uint32_t bl0 = *reinterpret_cast<const uint32_t*>(wl + ks * 16), bl1 = *reinterpret_cast<const uint32_t*>(wl + ks * 16 + 8);
                    ga_mma16816(c, al[ks], bh0, bh1);
                    ga_mma16816(c, ah[ks], bl0, bl1);
                    ga_mma16816(c, ah[ks], bh0, bh1);
                }
                const int o0 = nt * 8 + 2 * q;
                const float bp0 = __ldg(W.bp + o0), bp1 = __ldg(W.bp + o0 + 1);
                accA = fmaf(__ldg(awA + o0), tanh_fast(c[0] + bp0), accA);
                accA = fmaf(__ldg(awA + o0 + 1), tanh_fast(c[1] + bp1), accA);
                accB = fmaf(__ldg(awB + o0), tanh_fast(c[2] + bp0), accB);
                accB = fmaf(__ldg(awB + o0 + 1), tanh_fast(c[3] + bp1), accB);
            }
            accA += __shfl_xor_sync(0xffffffffu, accA, 1); accA += __shfl_xor_sync(0xffffffffu, accA, 2);
            accB += __shfl_xor_sync(0xffffffffu, accB, 1); accB += __shfl_xor_sync(0xffffffffu, accB, 2);
            if (q == 0) {
                if (pA < npairs) { const float v = accA * inv_temp; satt[iA * lda + jA] = v; satt[jA * lda + iA] = v; }
                if (pB < npairs) { const float v = accB * inv_temp; satt[iB * lda + jB] = v; satt[jB * lda + iB] = v; }
            }
        }
    }
    __syncthreads();
    // Wp is dead: stage Wa / Wo (padded rows) for the output projection while the softmax runs
    for (int i = tid; i < DOUT * DIN; i += GA_THREADS) {
        const int o = i / DIN, d = i - o * DIN;
        sWp[o * LDW + d] = __ldg(W.Wa + i);
        sWo[o * LDW + d] = __ldg(W.Wo + i);
    }
    // ---- softmax over j (one warp per row)
    {
        const int warp = tid >> 5, lane = tid & 31;
        for (int i = warp; i < n; i += GA_THREADS / 32) {
            float* r = satt + i * lda;
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
    // ---- aggregate: sagg[i][d] = sum_j a[i][j] x[j][d]; thread = (row, 4 channels), even / odd j in two chains per channel
    // (the chains of the former thread = (row, channel) loop: same bits, half the instructions per MAC)
    for (int k = tid; k < n * (DIN / 4); k += GA_THREADS) {
        const int i = k / (DIN / 4), d = (k - i * (DIN / 4)) * 4;
        const float* r = satt + i * lda;
        float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
        int j = 0;
        for (; j + 1 < n; j += 2) {
            const float r0 = r[j], r1 = r[j + 1];
            const float4 x0 = *reinterpret_cast<const float4*>(sx + j * LDX + d);
            const float4 x1 = *reinterpret_cast<const float4*>(sx + (j + 1) * LDX + d);
            a0.x = fmaf(r0, x0.x, a0.x); a0.y = fmaf(r0, x0.y, a0.y); a0.z = fmaf(r0, x0.z, a0.z); a0.w = fmaf(r0, x0.w, a0.w);
            a1.x = fmaf(r1, x1.x, a1.x); a1.y = fmaf(r1, x1.y, a1.y); a1.z = fmaf(r1, x1.z, a1.z); a1.w = fmaf(r1, x1.w, a1.w);
        }
        if (j < n) {
            const float r0 = r[j];
            const float4 x0 = *reinterpret_cast<const float4*>(sx + j * LDX + d);
            a0.x = fmaf(r0, x0.x, a0.x); a0.y = fmaf(r0, x0.y, a0.y); a0.z = fmaf(r0, x0.z, a0.z); a0.w = fmaf(r0, x0.w, a0.w);
        }
        *reinterpret_cast<float4*>(sagg + i * DIN + d) = make_float4(a0.x + a1.x, a0.y + a1.y, a0.z + a1.z, a0.w + a1.w);
    }
    __syncthreads();
    // ---- project + BN + SELU
    for (int k = tid; k < n * DOUT; k += GA_THREADS) {
        const int i = k / DOUT, o = k - i * DOUT;
        float v0 = __ldg(W.ba + o) + __ldg(W.bo + o), v1 = 0.f;
        const float* wa = sWp + o * LDW;
        const float* wo = sWo + o * LDW;
        const float* ag = sagg + i * DIN;
        const float* xi = sx + i * LDX;
#pragma unroll 8
        for (int d = 0; d < DIN; ++d) {
            v0 = fmaf(wa[d], ag[d], v0);
            v1 = fmaf(wo[d], xi[d], v1);
        }
        const float v = selu_f((v0 + v1) * __ldg(W.bns + o) + __ldg(W.bnt + o));
        a.y[z][((long long)b * nmax + i) * DOUT + o] = v;
    }
}

template <int DIN, int DOUT>
static cudaError_t launch_gat_rows_t(const GatArgs& a, int B, int nz, cudaStream_t s);
template <int DIN, int DOUT>
static cudaError_t launch_gat_t(const GatArgs& a, int B, int nz, cudaStream_t s) {
    int nmax = a.nmax[0];
    for (int z = 1; z < nz; ++z) nmax = a.nmax[z] > nmax ? a.nmax[z] : nmax;
    const size_t smem = ((size_t)nmax * (DIN + 4) + 2 * DOUT * (DIN + 1) + (size_t)nmax * (nmax + 1) + (size_t)nmax * DIN) * sizeof(float) +
                        (size_t)nmax * (nmax + 1) / 2 * sizeof(unsigned short) + 16;
    if (smem > 200 * 1024 || nmax > 255) return launch_gat_rows_t<DIN, DOUT>(a, B, nz, s);   // pair table packs (i, j) in 16 bits
    {
        cudaError_t e = ensure_dyn_smem(reinterpret_cast<const void*>(gat_kernel<DIN, DOUT>), smem);
        if (e != cudaSuccess) return e;
    }
    dim3 grid(1 + (a.has_master ? 1 : 0), B, nz);
    (void)launch_pdl(gat_kernel<DIN, DOUT>, dim3(grid), dim3((unsigned)(GA_THREADS)), (size_t)(smem), s, a);
    return cudaGetLastError();
}
// Row-block fallback for graphs too large for the one-CTA form (n > ~150 nodes: utterances beyond ~9 s): one CTA per
// GA_ROWS output rows, every pair evaluated from both sides.
static constexpr int GA_ROWS = 8;

template <int DIN, int DOUT>
__global__ void __launch_bounds__(256)
gat_rows_kernel(const GatArgs a) {
    pdl_wait();       // launched with the PDL attribute: nothing above this line touches global memory
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
static cudaError_t launch_gat_rows_t(const GatArgs& a, int B, int nz, cudaStream_t s) {
    int nmax = a.nmax[0];
    for (int z = 1; z < nz; ++z) nmax = a.nmax[z] > nmax ? a.nmax[z] : nmax;
    const size_t smem = ((size_t)nmax * DIN + DOUT * DIN + GA_ROWS * nmax + GA_ROWS * DIN) * sizeof(float);
    {
        cudaError_t e = ensure_dyn_smem(reinterpret_cast<const void*>(gat_rows_kernel<DIN, DOUT>), smem);
        if (e != cudaSuccess) return e;
    }
    dim3 grid((nmax + GA_ROWS - 1) / GA_ROWS + (a.has_master ? 1 : 0), B, nz);
    (void)launch_pdl(gat_rows_kernel<DIN, DOUT>, dim3(grid), dim3((unsigned)(256)), (size_t)(smem), s, a);
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
    pdl_wait();       // launched with the PDL attribute: nothing above this line touches global memory
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
    if (D == 64) (void)launch_pdl(graph_pool_kernel<64>, dim3(grid), dim3((unsigned)(128)), (size_t)(smem), s, a);
    else if (D == 32) (void)launch_pdl(graph_pool_kernel<32>, dim3(grid), dim3((unsigned)(128)), (size_t)(smem), s, a);
    else return cudaErrorInvalidValue;
    return cudaGetLastError();
}

// =================================================================================================
// HS-GAL input: xc = cat( proj_type1(x1), proj_type2(x2) )   (sslassist.py:201-206), both branches per launch.
template <int D>
__global__ void __launch_bounds__(256)
hproj_kernel(const HprojArgs a) {
    pdl_wait();       // launched with the PDL attribute: nothing above this line touches global memory
    // one CTA per (utterance, branch): both projection matrices staged once in shared memory with padded rows, so that the
    // lane = output-channel reads are conflict-free (the per-thread strided __ldg walk of W made this tiny kernel
    // L1-throughput bound: 110 us at B = 64)
    __shared__ float sW[2][D][D + 1];
    __shared__ float sx[256 / D][D];
    const int z = blockIdx.z, b = blockIdx.y;
    const int n1 = a.n1[z] ? a.n1[z][b] : a.n1c[z];
    const int n2 = a.n2[z] ? a.n2[z][b] : a.n2c[z];
    const int nm = a.n1max[z] + a.n2max[z];
    const int tid = threadIdx.x;
    for (int i = tid; i < D * D; i += 256) {
        sW[0][i / D][i % D] = __ldg(a.W1[z] + i);
        sW[1][i / D][i % D] = __ldg(a.W2[z] + i);
    }
    constexpr int NPP = 256 / D;                 // nodes per pass
    const int nl = tid / D, o = tid - nl * D;
    for (int i0 = 0; i0 < n1 + n2; i0 += NPP) {
        const int i = i0 + nl;
        const bool ok = i < n1 + n2;
        const bool t1 = i < n1;
        __syncthreads();
        if (ok) {
            const float* xr = t1 ? a.x1[z] + ((long long)b * a.n1max[z] + i) * D : a.x2[z] + ((long long)b * a.n2max[z] + (i - n1)) * D;
            sx[nl][o] = xr[o];
        }
        __syncthreads();
        if (ok) {
            const float* wr = sW[t1 ? 0 : 1][o];
            const float* xs = sx[nl];
            float v0 = __ldg((t1 ? a.b1[z] : a.b2[z]) + o), v1 = 0.f;
#pragma unroll 8
            for (int d = 0; d < D; d += 2) { v0 = fmaf(wr[d], xs[d], v0); v1 = fmaf(wr[d + 1], xs[d + 1], v1); }
            a.xc[z][((long long)b * nm + i) * D + o] = v0 + v1;
        }
    }
}
cudaError_t launch_hproj(const HprojArgs& a, int B, int nz, int D, cudaStream_t s) {
    dim3 grid(1, B, nz);
    if (D == 64) (void)launch_pdl(hproj_kernel<64>, dim3(grid), dim3((unsigned)(256)), (size_t)(0), s, a);
    else if (D == 32) (void)launch_pdl(hproj_kernel<32>, dim3(grid), dim3((unsigned)(256)), (size_t)(0), s, a);
    else return cudaErrorInvalidValue;
    return cudaGetLastError();
}

// =================================================================================================
// Readout (sslassist.py:557-597).  Per branch z: T_z = T1_z + Taug_z, S_z = S1_z + Saug_z, M_z = m_z + maug_z;
// T = max(T_1, T_2) element-wise (row i of branch 1 pairs with row i of branch 2), same for S, M;
// emb = [max_i |T|, mean_i T, max_i |S|, mean_i S, M]  (5 x 32);  out = W emb + b.
__global__ void __launch_bounds__(32)
readout_kernel(const ReadoutArgs a) {
    pdl_wait();       // launched with the PDL attribute: nothing above this line touches global memory
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
    (void)launch_pdl(readout_kernel, dim3((unsigned)(B)), dim3((unsigned)(32)), (size_t)(0), s, a);
    return cudaGetLastError();
}

}  // namespace occ
