// attn_tc.cu -- fused multi-head attention of the XLS-R encoder on tcgen05 for utterances of up to 256 frames
// (64 600 samples -> T = 201).  Restates torchaudio components.py:237-310 (SelfAttention.forward; fairseq scales q,
// here folded into W_q) for one (utterance, head, 128-query tile) per CTA:
//   TMA     : Q [128 x 64], K [NK x 64], V [NK x 64] bf16 boxes of the packed qkv activations, 128-byte swizzled
//   MMA 1   : S[128 x NK] = Q K^T  (K-major A and B, fp32 accumulator in TMEM columns [0, NK))
//   softmax : thread = query row = TMEM lane; two passes over S with tcgen05.ld (row max, then exp2 / row sum);
//             un-normalised P in bf16 goes to shared memory in the K-major swizzled layout the next MMA reads
//             (over the dead Q / K tiles); keys >= the utterance length get P = 0
//   MMA 2   : O[128 x 64] = P V    (V is the MN-major B operand exactly as TMA laid it down: rows = keys)
//   epilogue: O / rowsum -> bf16 ctx [B*T, 1024]
// Two CTAs are resident per SM (256 TMEM columns, ~92 KB shared memory each), so one CTA's softmax overlaps the
// other's loads and MMAs.  The legacy mma.sync kernel (frontend.cu) remains for T > 256.
#include "common.cuh"
#include "kernels.h"
#include <cuda.h>
#include <cstdlib>

namespace occ {

bool encode_tensor_map_bf16(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                            const uint32_t* box, std::string* err);   // gemm_tc.cu

static constexpr int ATC_THREADS = 128;
static constexpr int ATC_MAXK = 256;
// shared-memory map (1024 B aligned): P blocks (4 x 16 KB, 64 keys each) alias Q (16 KB) + K (32 KB); V behind them
static constexpr int ATC_Q_OFF = 0;
static constexpr int ATC_K_OFF = 16384;
static constexpr int ATC_V_OFF = 65536;
static constexpr int ATC_BAR_OFF = ATC_V_OFF + ATC_MAXK * 128;
static constexpr int ATC_SMEM = ATC_BAR_OFF + 64 + 1024;

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// 32 lanes x 8 columns: thread `lane` writes 8 x 32 bit to its TMEM row (here: 16 bf16 probabilities of one 16-key chunk)
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&r)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// D[tmem] (+)= A[tmem] * B[smem]: A = 128 rows (lanes) x 16 bf16 (8 columns, two per 32-bit column, even k in the low half)
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// MN-major, 128-byte-swizzled B operand: rows = K index (keys), 64 N elements (128 B) per row, 8-row atoms 1024 B apart
__device__ __forceinline__ uint64_t make_sw128_mn_desc(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
    d |= (uint64_t)(1024 >> 4) << 16;                  // LBO: next 64-wide MN atom (unused, N = 64)
    d |= (uint64_t)(1024 >> 4) << 32;                  // SBO: next 8-row K atom
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;                            // SWIZZLE_128B
    return d;
}

// NCH = 16-key chunks of the key axis (NK = 16 NCH).  NCH <= 13: the whole score row (<= 208 fp32) is read from TMEM once and
// stays in registers between the row-max and the exp pass (2 CTAs x 128 threads per SM leave 255 registers per thread);
// NCH = 16 re-reads S in a second pass.
template <int NCH>
__global__ void __launch_bounds__(ATC_THREADS)
attn_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV, int T,
               const int* __restrict__ klen, bf16* __restrict__ ctx, int probe) {
    constexpr int NK = NCH * 16;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + (((base + 1023u) & ~1023u) - base);
    uint64_t* bar_load = reinterpret_cast<uint64_t*>(smem + ATC_BAR_OFF);
    uint64_t* bar_s = bar_load + 1;
    uint64_t* bar_o = bar_load + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_load + 3);

    const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int kl = klen ? min(klen[b], T) : T;
    if (qt * 128 >= kl) return;                        // whole query tile is padding (uniform over the CTA)

    if (tid == 0) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmKV);
        mbar_init(bar_load, 1);
        mbar_init(bar_s, 1);
        mbar_init(bar_o, 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(tmem_slot, 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (tid == 0) {
        mbar_arrive_expect_tx(bar_load, (uint32_t)(128 * 128 + ((probe & 1) ? 0 : NK * 128) + ((probe & 2) ? 0 : NK * 128)));
        tma_load_3d(&tmQ, bar_load, smem + ATC_Q_OFF, h * 64, qt * 128, b);
        if (!(probe & 1)) tma_load_3d(&tmKV, bar_load, smem + ATC_K_OFF, 1024 + h * 64, 0, b);
        if (!(probe & 2)) tma_load_3d(&tmKV, bar_load, smem + ATC_V_OFF, 2048 + h * 64, 0, b);
        mbar_wait(bar_load, 0, 10);
        tc_fence_after();
        // ---- S = Q K^T : M = 128, N = NK, K = 64 (4 steps of 16)
        const uint32_t idesc = make_idesc_bf16(NK);
        const uint64_t adesc = make_sw128_desc(smem_u32(smem + ATC_Q_OFF));
        const uint64_t bdesc = make_sw128_desc(smem_u32(smem + ATC_K_OFF));
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) umma_bf16(tmem_base, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc, kk > 0 ? 1u : 0u);
        umma_commit(bar_s);
    }
    __syncwarp();
    mbar_wait(bar_s, 0, 11);
    tc_fence_after();

    // ---- softmax over the keys; this thread owns query row `tid` (TMEM lane tid)
    const uint32_t trow = tmem_base + ((uint32_t)(warp * 32) << 16);
    const float LOG2E = 1.4426950408889634f;
    const int nch = (kl + 15) >> 4;                    // 16-key chunks holding a valid key
    float lsum = 0.f;
    // P row: 128 B segments of 64 keys; 16 B chunk q of segment s sits at  s*16384 + (row/8)*1024 + (row%8)*128 + ((q ^ row%8) * 16)
    uint8_t* prow = smem + (tid >> 3) * 1024 + (tid & 7) * 128;
    const int sw = tid & 7;
    auto store_chunk = [&](int c, const uint32_t (&pk)[8]) {
        uint8_t* seg = prow + (c >> 2) * 16384;
        const int q0 = (c & 3) * 2;
        *reinterpret_cast<uint4*>(seg + (((q0) ^ sw) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        *reinterpret_cast<uint4*>(seg + (((q0 + 1) ^ sw) << 4)) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
    };
    if constexpr (NCH <= 13) {
        uint32_t r[NCH][16];
#pragma unroll
        for (int c = 0; c < NCH; ++c) tmem_ld16(trow + (uint32_t)(c * 16), r[c]);     // columns >= kl hold finite junk: masked below
        tmem_ld_wait();
        if (kl >= NK - 16) {          // common case: only the last chunk holds masked keys
            const int rem = kl - (NK - 16);
#pragma unroll
            for (int j = 0; j < 16; ++j) if (j >= rem) r[NCH - 1][j] = 0xff800000u;   // -inf: P = 0
        } else {
#pragma unroll
            for (int c = 0; c < NCH; ++c)
#pragma unroll
                for (int j = 0; j < 16; ++j) if (c * 16 + j >= kl) r[c][j] = 0xff800000u;
        }
        float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
#pragma unroll
            for (int j = 0; j < 16; j += 4) {
                m0 = fmaxf(m0, fmaxf(__uint_as_float(r[c][j]), __uint_as_float(r[c][j + 1])));
                m1 = fmaxf(m1, fmaxf(__uint_as_float(r[c][j + 2]), __uint_as_float(r[c][j + 3])));
            }
        }
        const float mneg = -fmaxf(m0, m1) * LOG2E;
        float l0 = 0.f, l1 = 0.f;
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            uint32_t pk[8];
#pragma unroll
            for (int j = 0; j < 16; j += 2) {
                const float p0 = ex2_approx(fmaf(__uint_as_float(r[c][j]), LOG2E, mneg));
                const float p1 = ex2_approx(fmaf(__uint_as_float(r[c][j + 1]), LOG2E, mneg));
                l0 += p0; l1 += p1;
                pk[j >> 1] = pack_bf16(p0, p1);
            }
            if (probe & 4) tmem_st8(trow + (uint32_t)(c * 8), pk);     // P over the (already read) S columns of this lane
            else store_chunk(c, pk);
        }
        lsum = l0 + l1;
        if (probe & 4) tmem_st_wait();
    } else {
        float mx = -INFINITY;
        for (int c = 0; c < nch; ++c) {
            uint32_t r[16];
            tmem_ld16(trow + (uint32_t)(c * 16), r);
            tmem_ld_wait();
            const int rem = kl - c * 16;
#pragma unroll
            for (int j = 0; j < 16; ++j) if (j < rem) mx = fmaxf(mx, __uint_as_float(r[j]));
        }
        const float mneg = -mx * LOG2E;
        for (int c = 0; c < NCH; ++c) {
            uint32_t pk[8];
            if (c < nch) {
                uint32_t r[16];
                tmem_ld16(trow + (uint32_t)(c * 16), r);
                tmem_ld_wait();
                const int rem = kl - c * 16;
#pragma unroll
                for (int j = 0; j < 16; j += 2) {
                    float p0 = ex2_approx(fmaf(__uint_as_float(r[j]), LOG2E, mneg));
                    float p1 = ex2_approx(fmaf(__uint_as_float(r[j + 1]), LOG2E, mneg));
                    if (j >= rem) p0 = 0.f;
                    if (j + 1 >= rem) p1 = 0.f;
                    lsum += p0 + p1;
                    pk[j >> 1] = pack_bf16(p0, p1);
                }
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) pk[j] = 0u;
            }
            store_chunk(c, pk);
        }
    }
    constexpr int nchk = NCH;
    fence_proxy_async();        // P was written through the generic proxy, the MMA reads it through the async proxy
    tc_fence_before();          // all tcgen05.ld of S retired before O overwrites its columns
    __syncthreads();
    if (tid == 0) {
        tc_fence_after();
        // ---- O = P V : M = 128, N = 64, K = NK (steps of 16 keys); V is MN-major (rows = keys)
        const uint32_t idesc = make_idesc_bf16(64) | (1u << 16);
        const uint32_t pa = smem_u32(smem);
        const uint32_t va = smem_u32(smem + ATC_V_OFF);
        for (int kk = 0; kk < nchk; ++kk) {
            const uint64_t adesc = make_sw128_desc(pa + (uint32_t)((kk >> 2) * 16384 + (kk & 3) * 32));
            const uint64_t bdesc = make_sw128_mn_desc(va + (uint32_t)(kk * 2048));
            if (NCH <= 13 && (probe & 4)) umma_bf16_ts(tmem_base + 128u, tmem_base + (uint32_t)(kk * 8), bdesc, idesc, kk > 0 ? 1u : 0u);
            else umma_bf16(tmem_base, adesc, bdesc, idesc, kk > 0 ? 1u : 0u);
        }
        umma_commit(bar_o);
    }
    __syncwarp();
    mbar_wait(bar_o, 0, 12);
    tc_fence_after();

    // ---- epilogue: O / rowsum -> ctx
    {
        uint32_t r0[32], r1[32];
        const uint32_t ocol = (NCH <= 13 && (probe & 4)) ? 128u : 0u;
        tmem_ld32(trow + ocol, r0);
        tmem_ld32(trow + ocol + 32u, r1);
        tmem_ld_wait();
        const int q = qt * 128 + tid;
        if (q < kl) {
            const float inv = 1.0f / lsum;
            uint4* op = reinterpret_cast<uint4*>(ctx + ((long long)b * T + q) * 1024 + h * 64);
#pragma unroll
            for (int j = 0; j < 4; ++j)
                op[j] = make_uint4(pack_bf16(__uint_as_float(r0[8 * j]) * inv, __uint_as_float(r0[8 * j + 1]) * inv),
                                   pack_bf16(__uint_as_float(r0[8 * j + 2]) * inv, __uint_as_float(r0[8 * j + 3]) * inv),
                                   pack_bf16(__uint_as_float(r0[8 * j + 4]) * inv, __uint_as_float(r0[8 * j + 5]) * inv),
                                   pack_bf16(__uint_as_float(r0[8 * j + 6]) * inv, __uint_as_float(r0[8 * j + 7]) * inv));
#pragma unroll
            for (int j = 0; j < 4; ++j)
                op[4 + j] = make_uint4(pack_bf16(__uint_as_float(r1[8 * j]) * inv, __uint_as_float(r1[8 * j + 1]) * inv),
                                       pack_bf16(__uint_as_float(r1[8 * j + 2]) * inv, __uint_as_float(r1[8 * j + 3]) * inv),
                                       pack_bf16(__uint_as_float(r1[8 * j + 4]) * inv, __uint_as_float(r1[8 * j + 5]) * inv),
                                       pack_bf16(__uint_as_float(r1[8 * j + 6]) * inv, __uint_as_float(r1[8 * j + 7]) * inv));
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, 256);
}

template <int NCH>
static cudaError_t launch_attn_tc_t(const bf16* qkv, int B, int T, const int* klen, bf16* ctx, cudaStream_t s, std::string* err) {
    constexpr int NK = NCH * 16;
    CUtensorMap tq, tkv;
    const uint64_t dims[3] = {3072, (uint64_t)T, (uint64_t)B};
    const uint64_t st[2] = {3072ull * 2, (uint64_t)T * 3072ull * 2};
    const uint32_t boxq[3] = {64, 128, 1};
    const uint32_t boxk[3] = {64, (uint32_t)NK, 1};
    if (!encode_tensor_map_bf16(&tq, qkv, 3, dims, st, boxq, err)) return cudaErrorInvalidValue;
    if (!encode_tensor_map_bf16(&tkv, qkv, 3, dims, st, boxk, err)) return cudaErrorInvalidValue;
    static bool attr = false;
    if (!attr) {
        cudaError_t e = cudaFuncSetAttribute(attn_tc_kernel<NCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, ATC_SMEM);
        if (e != cudaSuccess) return e;
        attr = true;
    }
    dim3 grid((T + 127) / 128, 16, B);
    static const int probe = [] { const char* e = getenv("OCC_ATTN_PROBE"); return e ? atoi(e) : 0; }();   // dev: 1 = no K load, 2 = no V load (timing only)
    attn_tc_kernel<NCH><<<grid, ATC_THREADS, ATC_SMEM, s>>>(tq, tkv, T, klen, ctx, probe);
    return cudaGetLastError();
}

// qkv bf16 [B*T, 3072] -> ctx bf16 [B*T, 1024]; T <= 256.  Returns cudaErrorNotSupported when the shape is outside the fast path.
cudaError_t launch_attn_tc(const bf16* qkv, int B, int T, const int* klen, bf16* ctx, cudaStream_t s, std::string* err) {
    if (T <= 0 || T > ATC_MAXK || B <= 0) return cudaErrorNotSupported;
    if (T <= 64) return launch_attn_tc_t<4>(qkv, B, T, klen, ctx, s, err);
    if (T <= 128) return launch_attn_tc_t<8>(qkv, B, T, klen, ctx, s, err);
    if (T <= 208) return launch_attn_tc_t<13>(qkv, B, T, klen, ctx, s, err);
    return launch_attn_tc_t<16>(qkv, B, T, klen, ctx, s, err);
}

}  // namespace occ
