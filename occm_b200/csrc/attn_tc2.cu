// attn_tc2.cu -- persistent, software-pipelined form of the tcgen05 attention (attn_tc.cu holds the one-tile-per-CTA form
// and the description of the arithmetic: torchaudio components.py:237-310).  One CTA per SM, 3 warpgroups:
//   warps 0-3 / 4-7 : softmax groups 0 / 1 (thread = query row = TMEM lane); a group owns TMEM slot g (256 columns) and takes
//                     every second work item of the CTA (ping-pong)
//   warp 8          : TMA producer -- Q / K / V boxes of the next items into a 3-deep (2-deep for T > 208) shared-memory ring
//   warp 9          : MMA issuer  -- S = Q K^T into the slot, later O = P V with P read from TENSOR MEMORY (tcgen05.mma A
//                     operand in TMEM: the softmax group stores bf16 P over the S columns it has consumed), so no P tile in
//                     shared memory and the ring can hold loads that are two items ahead
// Per slot the chain  S-MMA -> softmax -> O-MMA -> epilogue  is serial, but the two slots overlap each other and every
// load; the per-CTA set-up (TMEM alloc, barrier init, descriptor prefetch) is paid once per SM instead of once per tile.
// Work item = (utterance b, head h, 128-query tile qt), statically strided over the CTAs; items beyond the utterance skipped.
#include "common.cuh"
#include "kernels.h"
#include <cuda.h>
#include <cstdlib>

namespace occ {

bool encode_tensor_map_bf16(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                            const uint32_t* box, std::string* err);   // gemm_tc.cu

static constexpr int A2_THREADS = 384;

// Development trace (OCC_ATTN_FLAGS bit 2): clock64 stamps of the first A2_TRACE_ITEMS work items of every CTA,
// [cta][item][slot * 8 + event]; events: 0 S issued, 1 first-half O issued, 2 O issued + committed (MMA warp);
// 3 s_full seen, 4 S in registers, 5 P stored, 6 o_full seen, 7 epilogue done (thread 0 of the softmax group).
static constexpr int A2_TRACE_ITEMS = 8;
__device__ long long g_attn_trace[160 * A2_TRACE_ITEMS * 16];
#define A2_STAMP(cond, item, slot, ev)                                                                              \
    do {                                                                                                            \
        if ((flags & 4) && (cond) && (item) < A2_TRACE_ITEMS && blockIdx.x < 160)                                   \
            g_attn_trace[((int)blockIdx.x * A2_TRACE_ITEMS + (item)) * 16 + (slot) * 8 + (ev)] = clock64();         \
    } while (0)

__device__ __forceinline__ void a2_tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void a2_tmem_st8(uint32_t taddr, const uint32_t (&r)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
}
__device__ __forceinline__ void a2_tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void a2_bar_sync(int id) { asm volatile("bar.sync %0, 256;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void a2_bar_arrive(int id) { asm volatile("bar.arrive %0, 256;" ::"r"(id) : "memory"); }
__device__ __forceinline__ float a2_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// D[tmem] (+)= A[tmem] * B[smem]: A = 128 rows (lanes) x 16 bf16 in 8 columns (two per 32-bit column, even k in the low half)
__device__ __forceinline__ void a2_umma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// MN-major, 128-byte-swizzled B operand (V: rows = keys, 64 head dims = 128 B per row, 8-row atoms 1024 B apart)
__device__ __forceinline__ uint64_t a2_sw128_mn_desc(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
    d |= (uint64_t)(1024 >> 4) << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

// PAIRED (129 <= T <= 256, fixed length): a ring stage holds BOTH 128-query tiles of one (utterance, head) plus its K and V, and
// the two softmax groups work on the two tiles of that pair at the same time -- K and V are staged once instead of twice
// (68 -> 42 KB of L2->SM traffic per query tile at T = 201) and two stages keep a whole pair in flight ahead of the math.
template <int NCH, bool PAIRED = false>
struct A2Cfg {
    static constexpr int NK = NCH * 16;
    static constexpr int NS = PAIRED ? 2 : (NCH <= 13 ? 3 : 2);     // shared-memory ring depth
    static constexpr int QBYTES = PAIRED ? 32768 : 16384;
    static constexpr int STAGE = QBYTES + 2 * NK * 128;             // Q (+ Q') + K + V
    static constexpr int BAR_OFF = NS * STAGE;
    static constexpr int SMEM = BAR_OFF + 256 + 1024;
};

template <int NCH, bool PAIRED>
__global__ void __launch_bounds__(A2_THREADS, 1)
attn_tc2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV, int T, int B,
                const int* __restrict__ klen, bf16* __restrict__ ctx, int flags) {
    using C = A2Cfg<NCH, PAIRED>;
    constexpr int NK = C::NK, NS = C::NS, QB = C::QBYTES;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + (((base + 1023u) & ~1023u) - base);
    uint64_t* ld_full = reinterpret_cast<uint64_t*>(smem + C::BAR_OFF);      // [NS]
    uint64_t* ld_free = ld_full + 3;                                         // [NS]
    uint64_t* s_full = ld_free + 3;                                          // [2]
    uint64_t* p_full = s_full + 2;                                           // [2]  128 arrivals
    uint64_t* o_full = p_full + 2;                                           // [2]
    uint64_t* t_free = o_full + 2;                                           // [2]  128 arrivals
    uint64_t* p_half = t_free + 2;                                           // [2]  128 arrivals (split O-MMA: first HALF P chunks stored)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(p_half + 2);
    // flags bit 0: the two softmax groups take TURNS at the exponential section (named barriers 1 / 2, strict alternation
    //              g0, g1, g0, ...): left alone both groups run in lock step, fight over the MUFU pipe during the exponentials
    //              and then idle together through the MMA / barrier / epilogue latencies; in turns one group's exponentials
    //              run under the other's latencies
    //       bit 1: split O-MMA (the first HALF chunks of P V are issued while the group still computes the remaining P chunks)
    constexpr int HALF = (NCH + 1) / 2;
    const bool split = NCH <= 13 && (flags & 2) != 0;
    const bool turns = (flags & 1) != 0;
    const int turn_limit = (flags >> 4) > 0 ? (flags >> 4) : 0x7fffffff;      // turns only for a group's first N units (0 = always)

    const int tid = threadIdx.x, warp = tid >> 5;
    const int QT = (T + 127) >> 7;
    const int total = PAIRED ? B * 16 : B * 16 * QT;       // PAIRED: work item = (b, h), its two query tiles go to the two groups

    if (tid == 0) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmKV);
        for (int i = 0; i < NS; ++i) { mbar_init(&ld_full[i], 1); mbar_init(&ld_free[i], PAIRED ? 256 : 128); }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&s_full[i], 1); mbar_init(&p_full[i], 128);
            mbar_init(&o_full[i], 1); mbar_init(&t_free[i], 128);
            mbar_init(&p_half[i], 128);
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
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        if (warp == 8) {
            // ------------------------------------------------------------------ TMA producer
            int j = 0;
            for (int it = blockIdx.x; it < total; it += gridDim.x) {
                const int qt = PAIRED ? 0 : it % QT, h = PAIRED ? (it & 15) : ((it / QT) & 15), b = PAIRED ? (it >> 4) : it / (QT * 16);
                const int kl = klen ? min(__ldg(klen + b), T) : T;
                if (qt * 128 >= kl) continue;
                const int st = j % NS;
                mbar_wait(&ld_free[st], ((uint32_t)(j / NS) & 1u) ^ 1u, 20);
                if (elect_one()) {
                    uint8_t* sq = smem + st * C::STAGE;
                    mbar_arrive_expect_tx(&ld_full[st], (uint32_t)C::STAGE);
                    tma_load_3d(&tmQ, &ld_full[st], sq, h * 64, qt * 128, b);
                    if (PAIRED) tma_load_3d(&tmQ, &ld_full[st], sq + 16384, h * 64, 128, b);      // rows >= T: zero fill
                    tma_load_3d(&tmKV, &ld_full[st], sq + QB, 1024 + h * 64, 0, b);
                    tma_load_3d(&tmKV, &ld_full[st], sq + QB + NK * 128, 2048 + h * 64, 0, b);
                }
                __syncwarp();
                ++j;
            }
        } else if (warp == 9 || warp == 10) {
            // ------------------------------------------------------------------ MMA issuers: one warp per TMEM slot
            // Each slot's chain  S-MMA -> (softmax) -> O-MMA  is issued by its own warp with plain (suspending) barrier waits.
            // The earlier single issuer polled both slots with mbarrier.try_wait; try_wait parks the thread for a
            // system-dependent time, which in practice serialised the two slots (tools/attn_trace.py: S of slot 1 went out only
            // after O of slot 0), and a spinning test_wait poll took issue slots from the two softmax warps of its SM
            // sub-partition, whose late P arrivals then held back the whole group.
            const int slot = warp - 9;
            const uint32_t idesc_s = make_idesc_bf16(NK);
            const uint32_t idesc_o = make_idesc_bf16(64) | (1u << 16);          // B (= V) is MN-major
            const uint32_t tb = tmem_base + (uint32_t)(slot * 256);
            int j = 0;             // PAIRED: pair counter (both slots take every pair); else unit counter (slots alternate)
            for (int it = blockIdx.x; it < total; it += gridDim.x) {
                if (!PAIRED) {
                    const int qt = it % QT, b = it / (QT * 16);
                    const int kl = klen ? min(__ldg(klen + b), T) : T;
                    if (qt * 128 >= kl) continue;
                    if ((j & 1) != slot) { ++j; continue; }
                }
                const int js = j++;
                const int st = js % NS;
                const uint32_t ph = PAIRED ? ((uint32_t)js & 1u) : ((uint32_t)(js >> 1) & 1u);     // units this slot has done, mod 2
                const int titem = js;
                mbar_wait(&ld_full[st], (uint32_t)(js / NS) & 1u, 30);
                mbar_wait(&t_free[slot], ph ^ 1u, 31);                       // epilogue of the slot's previous unit has drained it
                tc_fence_after();
                const uint32_t qa = smem_u32(smem + st * C::STAGE);
                const uint32_t va = qa + (uint32_t)(QB + NK * 128);
                if (elect_one()) {
                    const uint64_t adesc = make_sw128_desc(qa + (PAIRED ? (uint32_t)(slot * 16384) : 0u)), bdesc = make_sw128_desc(qa + (uint32_t)QB);
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk)
                        umma_bf16(tb, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc_s, kk > 0 ? 1u : 0u);
                    umma_commit(&s_full[slot]);
                }
                __syncwarp();
                A2_STAMP((tid & 31) == 0, titem, PAIRED ? slot : 0, 0);
                if (split) {
                    mbar_wait(&p_half[slot], ph, 32);
                    tc_fence_after();
                    if (elect_one()) {
#pragma unroll
                        for (int kk = 0; kk < HALF; ++kk)
                            a2_umma_ts(tb + 128u, tb + (uint32_t)(kk * 8), a2_sw128_mn_desc(va + (uint32_t)(kk * 2048)), idesc_o, kk > 0 ? 1u : 0u);
                    }
                    __syncwarp();
                    A2_STAMP((tid & 31) == 0, titem, PAIRED ? slot : 0, 1);
                }
                mbar_wait(&p_full[slot], ph, 33);
                tc_fence_after();
                if (elect_one()) {
                    if (split) {
#pragma unroll
                        for (int kk = HALF; kk < NCH; ++kk)
                            a2_umma_ts(tb + 128u, tb + (uint32_t)(kk * 8), a2_sw128_mn_desc(va + (uint32_t)(kk * 2048)), idesc_o, 1u);
                    } else {
#pragma unroll
                        for (int kk = 0; kk < NCH; ++kk)
                            a2_umma_ts(tb + 128u, tb + (uint32_t)(kk * 8), a2_sw128_mn_desc(va + (uint32_t)(kk * 2048)), idesc_o, kk > 0 ? 1u : 0u);
                    }
                    umma_commit(&o_full[slot]);
                }
                __syncwarp();
                A2_STAMP((tid & 31) == 0, titem, PAIRED ? slot : 0, 2);
            }
        }
    } else {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
        // ---------------------------------------------------------------------- softmax groups (thread = query row)
        const int g = warp >> 2;                            // group = TMEM slot
        const uint32_t trow = tmem_base + (uint32_t)(g * 256) + ((uint32_t)((warp & 3) * 32) << 16);
        const float LOG2E = 1.4426950408889634f;
        if (turns && g == 1) a2_bar_arrive(1);              // group 0 goes first
        int j = 0, nu = 0;                                  // nu: units this group has taken
        for (int it = blockIdx.x; it < total; it += gridDim.x) {
            if (!PAIRED && klen == nullptr && (j & 1) != g) { ++j; continue; }       // fixed length: no item is skipped, no decode needed
            const int qt = PAIRED ? g : it % QT, h = PAIRED ? (it & 15) : ((it / QT) & 15), b = PAIRED ? (it >> 4) : it / (QT * 16);
            const int kl = klen ? min(__ldg(klen + b), T) : T;
            if (!PAIRED && qt * 128 >= kl) continue;
            if (!PAIRED && (j & 1) != g) { ++j; continue; }
            // PAIRED: j counts pairs and this group takes tile g of every pair; else j counts units and the groups alternate
            const uint32_t ph = PAIRED ? ((uint32_t)j & 1u) : ((uint32_t)(j >> 1) & 1u);
            const int st = j % NS;
            const int tj = j;                                   // trace index: pair (PAIRED) or unit
            ++j;
            mbar_wait(&s_full[g], ph, 24);
            tc_fence_after();
            A2_STAMP((tid & 127) == 0, tj, PAIRED ? g : 0, 3);
            float lsum;
            if constexpr (NCH <= 13) {
                uint32_t r[NCH][16];
#pragma unroll
                for (int c = 0; c < NCH; ++c) a2_tmem_ld16(trow + (uint32_t)(c * 16), r[c]);    // columns >= kl: finite junk, masked below
                tmem_ld_wait();
                A2_STAMP((tid & 127) == 0, tj, PAIRED ? g : 0, 4);
                if (kl >= NK - 16) {          // common case: only the last chunk holds masked keys
                    const int rem = kl - (NK - 16);
#pragma unroll
                    for (int i = 0; i < 16; ++i) if (i >= rem) r[NCH - 1][i] = 0xff800000u;       // -inf: P = 0
                } else {
#pragma unroll
                    for (int c = 0; c < NCH; ++c)
#pragma unroll
                        for (int i = 0; i < 16; ++i) if (c * 16 + i >= kl) r[c][i] = 0xff800000u;
                }
                float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
                for (int c = 0; c < NCH; ++c) {
#pragma unroll
                    for (int i = 0; i < 16; i += 4) {
                        m0 = fmaxf(m0, fmaxf(__uint_as_float(r[c][i]), __uint_as_float(r[c][i + 1])));
                        m1 = fmaxf(m1, fmaxf(__uint_as_float(r[c][i + 2]), __uint_as_float(r[c][i + 3])));
                    }
                }
                const float mneg = -fmaxf(m0, m1) * LOG2E;
                float l0 = 0.f, l1 = 0.f;
                if (turns && nu < turn_limit) a2_bar_sync(1 + g);
#pragma unroll
                for (int c = 0; c < NCH; ++c) {
                    uint32_t pk[8];
#pragma unroll
                    for (int i = 0; i < 16; i += 2) {
                        const float p0 = a2_ex2(fmaf(__uint_as_float(r[c][i]), LOG2E, mneg));
                        const float p1 = a2_ex2(fmaf(__uint_as_float(r[c][i + 1]), LOG2E, mneg));
                        l0 += p0; l1 += p1;
                        pk[i >> 1] = pack_bf16(p0, p1);
                    }
                    a2_tmem_st8(trow + (uint32_t)(c * 8), pk);      // P over S columns this lane has already read
                    if (c == HALF - 1 && split) {
                        a2_tmem_st_wait();
                        tc_fence_before();
                        mbar_arrive(&p_half[g]);
                    }
                }
                lsum = l0 + l1;
            } else {
                // NCH = 16: two passes over S; P chunk c lands on columns [8c, 8c+8), inside S chunk c/2 which pass 2 has consumed
                float mx = -INFINITY;
#pragma unroll 1
                for (int c0 = 0; c0 < NCH; c0 += 4) {
                    uint32_t r[4][16];
#pragma unroll
                    for (int u = 0; u < 4; ++u) a2_tmem_ld16(trow + (uint32_t)((c0 + u) * 16), r[u]);
                    tmem_ld_wait();
#pragma unroll
                    for (int u = 0; u < 4; ++u)
#pragma unroll
                        for (int i = 0; i < 16; ++i) if ((c0 + u) * 16 + i < kl) mx = fmaxf(mx, __uint_as_float(r[u][i]));
                }
                const float mneg = -mx * LOG2E;
                lsum = 0.f;
                if (turns && nu < turn_limit) a2_bar_sync(1 + g);
#pragma unroll 1
                for (int c0 = 0; c0 < NCH; c0 += 4) {
                    uint32_t r[4][16];
#pragma unroll
                    for (int u = 0; u < 4; ++u) a2_tmem_ld16(trow + (uint32_t)((c0 + u) * 16), r[u]);
                    tmem_ld_wait();
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        uint32_t pk[8];
#pragma unroll
                        for (int i = 0; i < 16; i += 2) {
                            float p0 = a2_ex2(fmaf(__uint_as_float(r[u][i]), LOG2E, mneg));
                            float p1 = a2_ex2(fmaf(__uint_as_float(r[u][i + 1]), LOG2E, mneg));
                            if ((c0 + u) * 16 + i >= kl) p0 = 0.f;
                            if ((c0 + u) * 16 + i + 1 >= kl) p1 = 0.f;
                            lsum += p0 + p1;
                            pk[i >> 1] = pack_bf16(p0, p1);
                        }
                        a2_tmem_st8(trow + (uint32_t)((c0 + u) * 8), pk);
                    }
                }
            }
            if (turns && nu < turn_limit) a2_bar_arrive(2 - g);                // the other group's turn
            ++nu;
            a2_tmem_st_wait();
            tc_fence_before();
            mbar_arrive(&p_full[g]);
            A2_STAMP((tid & 127) == 0, tj, PAIRED ? g : 0, 5);
            // ---- epilogue: O (columns 128..191 of the slot) / rowsum -> ctx
            mbar_wait(&o_full[g], ph, 25);
            tc_fence_after();
            A2_STAMP((tid & 127) == 0, tj, PAIRED ? g : 0, 6);
            uint32_t r0[32], r1[32];
            tmem_ld32(trow + 128u, r0);
            tmem_ld32(trow + 160u, r1);
            tmem_ld_wait();
            tc_fence_before();
            mbar_arrive(&t_free[g]);
            // O / rowsum -> bf16, transposed through the (dead) Q tile of this item's ring stage so that every global store
            // instruction writes four full 128 B rows (thread-per-row stores touched 32 lines per instruction: ~1200 cycles/item).
            // The stage goes back to the producer only after these reads: ld_free counts the 128 threads of the group.
            {
                const float inv = 1.0f / lsum;
                const int lane = tid & 31;
                uint8_t* stg = smem + st * C::STAGE + (PAIRED ? g * 16384 : 0) + (warp & 3) * 4096;
                uint8_t* wr = stg + lane * 128;
                const int sw = lane & 7;
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    *reinterpret_cast<uint4*>(wr + ((i ^ sw) << 4)) =
                        make_uint4(pack_bf16(__uint_as_float(r0[8 * i]) * inv, __uint_as_float(r0[8 * i + 1]) * inv),
                                   pack_bf16(__uint_as_float(r0[8 * i + 2]) * inv, __uint_as_float(r0[8 * i + 3]) * inv),
                                   pack_bf16(__uint_as_float(r0[8 * i + 4]) * inv, __uint_as_float(r0[8 * i + 5]) * inv),
                                   pack_bf16(__uint_as_float(r0[8 * i + 6]) * inv, __uint_as_float(r0[8 * i + 7]) * inv));
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    *reinterpret_cast<uint4*>(wr + (((4 + i) ^ sw) << 4)) =
                        make_uint4(pack_bf16(__uint_as_float(r1[8 * i]) * inv, __uint_as_float(r1[8 * i + 1]) * inv),
                                   pack_bf16(__uint_as_float(r1[8 * i + 2]) * inv, __uint_as_float(r1[8 * i + 3]) * inv),
                                   pack_bf16(__uint_as_float(r1[8 * i + 4]) * inv, __uint_as_float(r1[8 * i + 5]) * inv),
                                   pack_bf16(__uint_as_float(r1[8 * i + 6]) * inv, __uint_as_float(r1[8 * i + 7]) * inv));
                __syncwarp();
                const int cc = lane & 7;
                const int q0 = qt * 128 + (warp & 3) * 32;
                bf16* ob = ctx + ((long long)b * T + q0) * 1024 + h * 64 + cc * 8;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int rr = i * 4 + (lane >> 3);
                    const uint4 v = *reinterpret_cast<const uint4*>(stg + rr * 128 + ((cc ^ (rr & 7)) << 4));
                    if (q0 + rr < kl) *reinterpret_cast<uint4*>(ob + (long long)rr * 1024) = v;
                }
                fence_proxy_async();          // generic accesses to the stage are ordered before the producer's next TMA write
                mbar_arrive(&ld_free[st]);
                A2_STAMP((tid & 127) == 0, tj, PAIRED ? g : 0, 7);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 9) tmem_dealloc(tmem_base, 512);
}

template <int NCH, bool PAIRED>
static cudaError_t launch_attn_tc2_t(const bf16* qkv, int B, int T, const int* klen, bf16* ctx, int num_sms, cudaStream_t s, std::string* err) {
    using C = A2Cfg<NCH, PAIRED>;
    CUtensorMap tq, tkv;
    const uint64_t dims[3] = {3072, (uint64_t)T, (uint64_t)B};
    const uint64_t st[2] = {3072ull * 2, (uint64_t)T * 3072ull * 2};
    const uint32_t boxq[3] = {64, 128, 1};
    const uint32_t boxk[3] = {64, (uint32_t)C::NK, 1};
    if (!encode_tensor_map_bf16(&tq, qkv, 3, dims, st, boxq, err)) return cudaErrorInvalidValue;
    if (!encode_tensor_map_bf16(&tkv, qkv, 3, dims, st, boxk, err)) return cudaErrorInvalidValue;
    {
        cudaError_t e = ensure_dyn_smem(reinterpret_cast<const void*>(attn_tc2_kernel<NCH, PAIRED>), C::SMEM);
        if (e != cudaSuccess) return e;
    }
    const int total = PAIRED ? B * 16 : B * 16 * ((T + 127) / 128);
    const int grid = total < num_sms ? total : num_sms;
    static const int flags = [] { const char* e = getenv("OCC_ATTN_FLAGS"); return e ? atoi(e) : 3; }();   // A/B switch; 3 = product path
    return launch_pdl(attn_tc2_kernel<NCH, PAIRED>, dim3((unsigned)grid), dim3(A2_THREADS), (size_t)C::SMEM, s, tq, tkv, T, B, klen, ctx, flags);
}

cudaError_t attn_tc2_trace_read(long long* host_out, int n) {
    const size_t bytes = sizeof(long long) * (size_t)(n < 160 * A2_TRACE_ITEMS * 16 ? n : 160 * A2_TRACE_ITEMS * 16);
    return cudaMemcpyFromSymbol(host_out, g_attn_trace, bytes);
}

// qkv bf16 [B*T, 3072] -> ctx bf16 [B*T, 1024]; T <= 256
cudaError_t launch_attn_tc2(const bf16* qkv, int B, int T, const int* klen, bf16* ctx, int num_sms, cudaStream_t s, std::string* err) {
    if (T <= 0 || T > 256 || B <= 0) return cudaErrorNotSupported;
    if (T <= 64) return launch_attn_tc2_t<4, false>(qkv, B, T, klen, ctx, num_sms, s, err);
    if (T <= 128) return launch_attn_tc2_t<8, false>(qkv, B, T, klen, ctx, num_sms, s, err);
    static const bool no_pair = [] { const char* e = getenv("OCC_ATTN_NOPAIR"); return e && e[0] == '1'; }();
    const bool paired = klen == nullptr && !no_pair;       // ragged batches keep the per-tile items (tiles beyond an utterance are skipped)
    if (T <= 208) return paired ? launch_attn_tc2_t<13, true>(qkv, B, T, klen, ctx, num_sms, s, err)
                                : launch_attn_tc2_t<13, false>(qkv, B, T, klen, ctx, num_sms, s, err);
    return paired ? launch_attn_tc2_t<16, true>(qkv, B, T, klen, ctx, num_sms, s, err)
                  : launch_attn_tc2_t<16, false>(qkv, B, T, klen, ctx, num_sms, s, err);
}

}  // namespace occ
