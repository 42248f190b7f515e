// attn_tc3.cu -- tcgen05 attention for utterances LONGER than 256 frames (BASELINE.json configs[3]: variable length up to 10 s =
// 499 frames; configs[4]: 30 s = 1 499 frames).  Same arithmetic as attn_tc2.cu (torchaudio components.py:237-310: softmax(q k^T) v
// per head, q pre-scaled by 1/8 through the packed weights, keys beyond the utterance masked), but the key axis is walked in chunks
// of 128 with an online softmax, so the score row never has to fit TMEM / registers.
//
// One persistent CTA per SM, 12 warps:
//   warps 0-3 / 4-7 : softmax group 0 / 1 (thread = query row = TMEM lane).  A group owns TMEM slot g (256 columns: S in
//                     [0,128), the chunk's O in [128,192), bf16 P in [192,256)) and every second work item of the CTA.
//   warp 8 / 10     : TMA producer of slot 0 / 1 -- the item's Q tile once, then its K and V chunks through a 3-deep ring
//   warp 9          : MMA issuer for both slots (polling two small state machines):
//                       S(k)  = Q K_k^T              (4 x K=16 steps, N = 128, SS form)
//                       O_k   = P_k V_k  (fresh)     (<= 8 x K=16 steps, N = 64, P read from TENSOR MEMORY, V MN-major in smem)
//                     S(k+1) is issued right behind PV(k): it touches neither P_k nor O_k, and its completion barrier
//                     (tcgen05.commit covers every earlier MMA) is also what lets the group overwrite P.
// Per chunk a softmax thread: loads its 128 scores, takes the running maximum m, p = 2^(s log2e - m), stores bf16 P,
// and -- one chunk late, under the next chunk's MMAs -- folds the previous chunk's O_k into its 64 fp32 output registers:
// o <- o * 2^(m_old - m_new) + O_k.  No rescaling pass over TMEM, no second pass over the keys.
// Work item = (utterance b, head h, 128-query tile qt), statically strided over the CTAs; tiles beyond an utterance are skipped.
#include "common.cuh"
#include "kernels.h"
#include <cuda.h>

namespace occ {

bool encode_tensor_map_bf16(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                            const uint32_t* box, std::string* err);   // gemm_tc.cu

static constexpr int A3_THREADS = 384;
static constexpr int A3_KC = 128;                          // keys per chunk
static constexpr int A3_NS = 3;                            // K/V ring depth per slot
static constexpr int A3_QB = 16384;                        // Q tile: 128 rows x 128 B
static constexpr int A3_KVB = 2 * A3_KC * 128;             // one ring stage: K chunk + V chunk
static constexpr int A3_SLOT = A3_QB + A3_NS * A3_KVB;     // 112 KB per slot
static constexpr int A3_BAR_OFF = 2 * A3_SLOT;
static constexpr int A3_SMEM = A3_BAR_OFF + 512 + 1024;    // barriers + alignment slack (<= 227 KB)

__device__ __forceinline__ void a3_tmem_st8(uint32_t taddr, const uint32_t* r) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
}
__device__ __forceinline__ void a3_tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float a3_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// D[tmem] (+)= A[tmem] * B[smem]: A = 128 rows (lanes) x 16 bf16 in 8 columns (two per 32-bit column, even k in the low half)
__device__ __forceinline__ void a3_umma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// MN-major, 128-byte-swizzled B operand (V: rows = keys, 64 head dims = 128 B per row, 8-row atoms 1024 B apart)
__device__ __forceinline__ uint64_t a3_sw128_mn_desc(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
    d |= (uint64_t)(1024 >> 4) << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

struct A3Item { int b, h, qt, kl; };

// The CTA's work items in order (static stride over the grid, tiles beyond the utterance skipped); slot s takes every second
// one.  Every role of the CTA walks the same sequence, so no item list has to be communicated.
__device__ __forceinline__ bool a3_next(int& it, int& j, int slot, int total, int QT, int T, const int* __restrict__ klen, A3Item& x) {
    while (it < total) {
        const int cur = it;
        it += (int)gridDim.x;
        const int qt = cur % QT, h = (cur / QT) & 15, b = cur / (QT * 16);
        const int kl = klen ? min(__ldg(klen + b), T) : T;
        if (qt * 128 >= kl) continue;
        const int jj = j++;
        if ((jj & 1) != slot) continue;
        x.b = b; x.h = h; x.qt = qt; x.kl = kl;
        return true;
    }
    return false;
}

__global__ void __launch_bounds__(A3_THREADS, 1)
attn_tc3_kernel(const __grid_constant__ CUtensorMap tm, int T, int B, const int* __restrict__ klen, bf16* __restrict__ ctx) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + (((base + 1023u) & ~1023u) - base);
    // barriers, per slot s (index s * 12 + ...): q_full, q_free, kv_full[3], kv_free[3], s_full, p_full, o_full, o_free
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + A3_BAR_OFF);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 24);
    auto q_full = [&](int s) { return bars + s * 12 + 0; };
    auto q_free = [&](int s) { return bars + s * 12 + 1; };
    auto kv_full = [&](int s, int st) { return bars + s * 12 + 2 + st; };
    auto kv_free = [&](int s, int st) { return bars + s * 12 + 5 + st; };
    auto s_full = [&](int s) { return bars + s * 12 + 8; };
    auto p_full = [&](int s) { return bars + s * 12 + 9; };
    auto o_full = [&](int s) { return bars + s * 12 + 10; };
    auto o_free = [&](int s) { return bars + s * 12 + 11; };

    const int tid = threadIdx.x, warp = tid >> 5;
    const int QT = (T + 127) >> 7;
    const int total = B * 16 * QT;

    if (tid == 0) {
        tma_prefetch_desc(&tm);
        for (int s = 0; s < 2; ++s) {
            mbar_init(q_full(s), 1); mbar_init(q_free(s), 128);
            for (int i = 0; i < A3_NS; ++i) { mbar_init(kv_full(s, i), 1); mbar_init(kv_free(s, i), 1); }
            mbar_init(s_full(s), 1); mbar_init(p_full(s), 128);
            mbar_init(o_full(s), 1); mbar_init(o_free(s), 128);
        }
        fence_barrier_init();
    }
    if (warp == 9) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_trigger();
    pdl_wait();

    if (warp >= 8) {
        // register budget: 8 softmax warps x 224 + 4 x 56 = 2016 of the SM's 2048 register-warps.  An exact fit (224 / 64) left some
        // softmax warps blocked in setmaxnreg.inc forever (measured: one group per CTA never started) -- keep a margin.
        asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
        if (warp == 8 || warp == 10) {
            // ------------------------------------------------------------------ TMA producer of slot s
            const int s = warp == 8 ? 0 : 1;
            uint8_t* sl = smem + s * A3_SLOT;
            int it = blockIdx.x, j = 0, ni = 0, u = 0;
            A3Item x;
            while (a3_next(it, j, s, total, QT, T, klen, x)) {
                mbar_wait(q_free(s), ((uint32_t)ni & 1u) ^ 1u, 30);      // the group has finished with the previous item's Q tile (ctx staging)
                if (elect_one()) {
                    mbar_arrive_expect_tx(q_full(s), (uint32_t)A3_QB);
                    tma_load_3d(&tm, q_full(s), sl, x.h * 64, x.qt * 128, x.b);          // rows >= T: zero fill
                }
                __syncwarp();
                ++ni;
                const int nch = (x.kl + A3_KC - 1) / A3_KC;
                for (int k = 0; k < nch; ++k, ++u) {
                    const int st = u % A3_NS;
                    mbar_wait(kv_free(s, st), ((uint32_t)(u / A3_NS) & 1u) ^ 1u, 31);
                    if (elect_one()) {
                        uint8_t* kv = sl + A3_QB + st * A3_KVB;
                        mbar_arrive_expect_tx(kv_full(s, st), (uint32_t)A3_KVB);
                        tma_load_3d(&tm, kv_full(s, st), kv, 1024 + x.h * 64, k * A3_KC, x.b);
                        tma_load_3d(&tm, kv_full(s, st), kv + A3_KC * 128, 2048 + x.h * 64, k * A3_KC, x.b);
                    }
                    __syncwarp();
                }
            }
        } else if (warp == 9) {
            // ------------------------------------------------------------------ MMA issuer (both slots)
            const uint32_t idesc_s = make_idesc_bf16(A3_KC);
            const uint32_t idesc_o = make_idesc_bf16(64) | (1u << 16);          // B (= V) is MN-major
            // one state machine per slot: chunk k of the slot's current item, unit counter u (barrier parities / ring stage), item
            // counter ni, step 0 = S(k) pending / 1 = PV(k) pending
            struct Slot { int it, j, k, u, ni, step, kl; bool act; };
            auto advance = [&](Slot& z, const int s) -> bool {
                if (!z.act) return false;
                const int st = z.u % A3_NS;
                const uint32_t tslot = tmem_base + (uint32_t)(s * 256);
                uint8_t* sl = smem + s * A3_SLOT;
                if (z.step == 0) {
                    bool ok = mbar_test_wait(kv_full(s, st), (uint32_t)(z.u / A3_NS) & 1u);
                    if (ok && z.k == 0) ok = mbar_test_wait(q_full(s), (uint32_t)z.ni & 1u);
                    if (!ok) return false;
                    tc_fence_after();
                    const uint64_t adesc = make_sw128_desc(smem_u32(sl));
                    const uint64_t bdesc = make_sw128_desc(smem_u32(sl + A3_QB + st * A3_KVB));
                    if (elect_one()) {
#pragma unroll
                        for (int kk = 0; kk < 4; ++kk)
                            umma_bf16(tslot, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc_s, kk > 0 ? 1u : 0u);
                        umma_commit(s_full(s));
                    }
                    __syncwarp();
                    z.step = 1;
                    return true;
                }
                // P(u) stored by the group, and the group has read the previous chunk's O out of the slot
                if (!(mbar_test_wait(p_full(s), (uint32_t)z.u & 1u) && mbar_test_wait(o_free(s), ((uint32_t)z.u & 1u) ^ 1u))) return false;
                tc_fence_after();
                const uint32_t va = smem_u32(sl + A3_QB + st * A3_KVB + A3_KC * 128);
                const int nvalid = min(A3_KC, z.kl - z.k * A3_KC);
                const int nkk = (nvalid + 15) >> 4;            // 16-key steps holding a valid key (P = 0 beyond)
                if (elect_one()) {
                    for (int kk = 0; kk < nkk; ++kk)
                        a3_umma_ts(tslot + 128u, tslot + 192u + (uint32_t)(kk * 8), a3_sw128_mn_desc(va + (uint32_t)(kk * 2048)), idesc_o, kk > 0 ? 1u : 0u);
                    umma_commit(o_full(s));
                    umma_commit(kv_free(s, st));
                }
                __syncwarp();
                z.step = 0;
                ++z.u; ++z.k;
                if (z.k * A3_KC >= z.kl) {
                    z.k = 0; ++z.ni;
                    A3Item x;
                    z.act = a3_next(z.it, z.j, s, total, QT, T, klen, x);
                    z.kl = x.kl;
                }
                return true;
            };
            Slot z0, z1;
            {
                A3Item x;
                z0.it = blockIdx.x; z0.j = 0; z0.k = 0; z0.u = 0; z0.ni = 0; z0.step = 0;
                z0.act = a3_next(z0.it, z0.j, 0, total, QT, T, klen, x); z0.kl = x.kl;
                z1.it = blockIdx.x; z1.j = 0; z1.k = 0; z1.u = 0; z1.ni = 0; z1.step = 0;
                z1.act = a3_next(z1.it, z1.j, 1, total, QT, T, klen, x); z1.kl = x.kl;
            }
            long long t0 = clock64();
            while (z0.act || z1.act) {
                bool did = advance(z0, 0);
                did |= advance(z1, 1);
                if (did) t0 = clock64();
                else if (clock64() - t0 > 4000000000LL) {
                    if ((tid & 31) == 0) printf("occ: attn_tc3 MMA issuer timeout block=%d u=%d/%d step=%d/%d\n", (int)blockIdx.x, z0.u, z1.u, z0.step, z1.step);
                    __trap();
                }
            }
        }
    } else {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 224;");
        // ---------------------------------------------------------------------- softmax groups (thread = query row)
        const int g = warp >> 2;
        const int lane = tid & 31;
        const uint32_t trow = tmem_base + (uint32_t)(g * 256) + ((uint32_t)((warp & 3) * 32) << 16);
        const float LOG2E = 1.4426950408889634f;
        int it = blockIdx.x, j = 0, u = 0;
        A3Item x;
        while (a3_next(it, j, g, total, QT, T, klen, x)) {
            const int nch = (x.kl + A3_KC - 1) / A3_KC;
            float m = -INFINITY, l = 0.f, sc_prev = 0.f;       // m: running maximum, already multiplied by log2(e)
            float o[64];
#pragma unroll
            for (int i = 0; i < 64; ++i) o[i] = 0.f;
            for (int k = 0; k < nch; ++k, ++u) {
                mbar_wait(s_full(g), (uint32_t)u & 1u, 32);
                tc_fence_after();
                uint32_t r[4][32];
#pragma unroll
                for (int c = 0; c < 4; ++c) tmem_ld32(trow + (uint32_t)(c * 32), r[c]);
                tmem_ld_wait();
                const int nvalid = x.kl - k * A3_KC;           // >= 1; keys at or beyond it are masked
                if (nvalid < A3_KC) {
#pragma unroll
                    for (int c = 0; c < 4; ++c)
#pragma unroll
                        for (int i = 0; i < 32; ++i) if (c * 32 + i >= nvalid) r[c][i] = 0xff800000u;      // -inf: P = 0
                }
                float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int i = 0; i < 32; i += 4) {
                        mx0 = fmaxf(mx0, fmaxf(__uint_as_float(r[c][i]), __uint_as_float(r[c][i + 1])));
                        mx1 = fmaxf(mx1, fmaxf(__uint_as_float(r[c][i + 2]), __uint_as_float(r[c][i + 3])));
                    }
                const float mn = fmaxf(m, fmaxf(mx0, mx1) * LOG2E);       // finite: every chunk holds a valid key
                const float sc = a3_ex2(m - mn);                           // first chunk: 2^(-inf) = 0
                m = mn;
                float l0 = 0.f, l1 = 0.f;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    uint32_t pk[16];
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {
                        const float p0 = a3_ex2(fmaf(__uint_as_float(r[c][i]), LOG2E, -mn));
                        const float p1 = a3_ex2(fmaf(__uint_as_float(r[c][i + 1]), LOG2E, -mn));
                        l0 += p0; l1 += p1;
                        pk[i >> 1] = pack_bf16(p0, p1);
                    }
                    a3_tmem_st8(trow + 192u + (uint32_t)(c * 16), pk);    // P: columns 192..255 of the slot
                    a3_tmem_st8(trow + 192u + (uint32_t)(c * 16 + 8), pk + 8);
                }
                l = fmaf(l, sc, l0 + l1);
                a3_tmem_st_wait();
                tc_fence_before();
                mbar_arrive(p_full(g));
                if (k > 0) {
                    // previous chunk's O_k (columns 128..191), under this chunk's PV MMAs
                    mbar_wait(o_full(g), (uint32_t)(u - 1) & 1u, 33);
                    tc_fence_after();
                    tmem_ld32(trow + 128u, r[0]);
                    tmem_ld32(trow + 160u, r[1]);
                    tmem_ld_wait();
                    tc_fence_before();
                    mbar_arrive(o_free(g));
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        o[i] = fmaf(o[i], sc_prev, __uint_as_float(r[0][i]));
                        o[32 + i] = fmaf(o[32 + i], sc_prev, __uint_as_float(r[1][i]));
                    }
                }
                sc_prev = sc;
            }
            {
                uint32_t r0[32], r1[32];
                mbar_wait(o_full(g), (uint32_t)(u - 1) & 1u, 34);
                tc_fence_after();
                tmem_ld32(trow + 128u, r0);
                tmem_ld32(trow + 160u, r1);
                tmem_ld_wait();
                tc_fence_before();
                mbar_arrive(o_free(g));
                const float inv = 1.0f / l;
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    o[i] = fmaf(o[i], sc_prev, __uint_as_float(r0[i])) * inv;
                    o[32 + i] = fmaf(o[32 + i], sc_prev, __uint_as_float(r1[i])) * inv;
                }
            }
            // ---- ctx: bf16, transposed through the (now dead) Q tile of the slot so that every global store instruction writes
            // four full 128 B rows; the tile goes back to the producer only after these reads (q_free counts the 128 threads).
            {
                uint8_t* stg = smem + g * A3_SLOT + (warp & 3) * 4096;
                uint8_t* wr = stg + lane * 128;
                const int sw = lane & 7;
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    *reinterpret_cast<uint4*>(wr + ((i ^ sw) << 4)) =
                        make_uint4(pack_bf16(o[8 * i], o[8 * i + 1]), pack_bf16(o[8 * i + 2], o[8 * i + 3]),
                                   pack_bf16(o[8 * i + 4], o[8 * i + 5]), pack_bf16(o[8 * i + 6], o[8 * i + 7]));
                __syncwarp();
                const int cc = lane & 7;
                const int q0 = x.qt * 128 + (warp & 3) * 32;
                bf16* ob = ctx + ((long long)x.b * T + q0) * 1024 + x.h * 64 + cc * 8;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int rr = i * 4 + (lane >> 3);
                    const uint4 v = *reinterpret_cast<const uint4*>(stg + rr * 128 + ((cc ^ (rr & 7)) << 4));
                    if (q0 + rr < x.kl) *reinterpret_cast<uint4*>(ob + (long long)rr * 1024) = v;
                }
                fence_proxy_async();          // generic accesses to the tile are ordered before the producer's next TMA write
                mbar_arrive(q_free(g));
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 9) tmem_dealloc(tmem_base, 512);
}

// qkv bf16 [B*T, 3072] -> ctx bf16 [B*T, 1024]; any T (used for T > 256)
cudaError_t launch_attn_tc3(const bf16* qkv, int B, int T, const int* klen, bf16* ctx, int num_sms, cudaStream_t s, std::string* err) {
    if (T <= 0 || B <= 0) return cudaErrorInvalidValue;
    CUtensorMap tm;
    const uint64_t dims[3] = {3072, (uint64_t)T, (uint64_t)B};
    const uint64_t st[2] = {3072ull * 2, (uint64_t)T * 3072ull * 2};
    const uint32_t box[3] = {64, 128, 1};
    if (!encode_tensor_map_bf16(&tm, qkv, 3, dims, st, box, err)) return cudaErrorInvalidValue;
    cudaError_t e = ensure_dyn_smem(reinterpret_cast<const void*>(attn_tc3_kernel), A3_SMEM);
    if (e != cudaSuccess) return e;
    const int total = B * 16 * ((T + 127) / 128);
    const int want = (total + 1) / 2;                 // two slots per CTA
    const int grid = want < num_sms ? want : num_sms;
    return launch_pdl(attn_tc3_kernel, dim3((unsigned)grid), dim3(A3_THREADS), (size_t)A3_SMEM, s, tm, T, B, klen, ctx);
}

}  // namespace occ
