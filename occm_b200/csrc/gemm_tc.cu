// gemm_tc.cu -- the tcgen05/TMEM/TMA "shift-GEMM" that carries every dense contraction of the
// XLS-R frontend (conv layers 1-6, feature projection, grouped positional conv, QKV / out-proj / FC1 / FC2,
// LL).  One persistent CTA per SM; warp-specialised:
//   warp 0   : TMA producer (one lane) -- 5-D A boxes {64 cols, 1 sub, 128 rows}, 3-D W boxes {64, BN}
//   warp 1   : tcgen05.mma issuer (one lane) -- 128 x BN x 16 bf16 MMAs, fp32 accumulators in TMEM
//   warp 2   : TMEM allocator
//   warps 4-15: epilogue -- tcgen05.ld 32x32b, bias / erf-GELU / residual, vectorised global stores
// Two TMEM accumulator stages (2 x BN columns) overlap the epilogue of tile i with the MMAs of tile i+1.
#include "common.cuh"
#include "occ_internal.h"
#include <cuda.h>
#include <mutex>
#include <map>
#include <utility>
#include <cstdlib>

namespace occ {

static constexpr int BM = 128;            // rows per CTA; a CTA pair (cta_group::2) owns a 256-row tile
static constexpr int BK = 64;
static constexpr int NUM_EPI_WARPS = 12;
static constexpr int NUM_THREADS = 128 + NUM_EPI_WARPS * 32;
static constexpr int EPI_PITCH = 144;     // bytes per staged row: 32 fp32 + 16 B pad -> conflict-free 16 B stores and loads

struct GemmParams {
    int tiles_m_per_batch, tiles_n, num_tiles;
    int nkb, kb_per_tap, ntaps_w, shift0, shift_h, shift_w, a_sub, a_col_group_stride, nseg;
    int rows_per_batch, N;
    const float* bias;
    const float* resid;
    float* out32;
    float* out32_copy;
    bf16* out16;
    long long ldo;            // leading dimension (elements) shared by resid / out32 / out16
    long long out16_plane;
    int gelu;
    const int* row_valid;
    int conv_mode, conv_wp, conv_h0, conv_h1, conv_wm;
    const float *cs1, *ct1, *cs2, *ct2;
    const bf16* cresid;
    long long cresid_plane;
    const int* conv_wlen;
    int a_rank2;    // A is addressed through a plain 2-D map {cols, rows} (linear layers, single plane)
    int flags;      // EPI_* combination (selects the compiled epilogue variant)
    int halo;       // BN = 64 shift-GEMMs with one 64-channel k-block per tap: the A rows every tap of a tile reads are loaded
    int halo_min;   // ONCE (256 rows from t0 + halo_min, per plane) and each tap's MMAs address them by a row-shifted descriptor
    int cat;        // conv mode, split, halo: pass 1 = A_hi x [W_hi ; W_lo] as ONE N = 128 MMA per K step (CTA 0 stages the 64 W_hi rows, CTA 1 the
                    // 64 W_lo rows -> accumulator columns 0..63 / 64..127), pass 2 = A_lo x W_hi (N = 64) on top of columns 0..63; the epilogue
                    // adds the two column blocks.  2 MMAs per K step instead of 3 (every tcgen05.mma here sits on its ~110-cycle floor).
    int halo_rows2; // rows of the second halo box (tmA2): the taps reach halo_max - halo_min <= 128 rows past a CTA's own 128, rounded up to 8
    int ln_mode;    // fused LayerNorm(512) + GELU epilogue: a CTA pair walks M-tiles and computes both 256-column halves of each
    int num_m_tiles;// back to back (accumulator stage 0 = columns 0..255, stage 1 = columns 256..511 of the SAME rows)
    const float *ln_g, *ln_b;
    bf16* x16_out;          // LayerNorm fold, producer side: bf16 copy of the output + per-row partial sums
    float* stat_out;
    const float* stat_in;   // LayerNorm fold, consumer side: per-row partial sums [rows][12][2], column sums of W'
    const float* colsum;
    int N_stat;             // width of the normalised rows (= K of the consumer GEMM)
    int toep_T, toep_Tp, toep_B;   // Toeplitz pos-conv output mapping (GemmDesc::toep_T)
    int kk_steps;   // K = 16 MMA steps issued per 64-wide k-block (4; 2 when only the first 32 channels of every k-block are non-zero)
    int probe;      // development probes (occ_debug_gemm_bench): 1 = no TMA loads, 2 = no MMAs, 4 = no epilogue work
};

// BN = N extent of the pair's tile.  Each CTA stages its own 128 rows of A and its own BN/2 rows of W per k-block:
// (128 + BN/2) * 128 B of L2->SM traffic for 128 x BN x 64 MACs, i.e. 256 FLOP/B at BN = 256 against the
// 170 FLOP/B of a single-CTA 128 x 256 tile, which profiles/r1a showed pinned at the L2 throughput cap.
// Tensor maps of the TMA epilogue (MODE 5): the fp32 residual stream (loaded as the residual AND stored to, in place), the
// optional second fp32 destination (hidden-state export) and the optional bf16 copy (LayerNorm-fold operand).
struct EpiMaps { CUtensorMap x32, copy32, x16; };
static constexpr int EPI5_WARP_BYTES = 2 * 2048 + 1024;       // per epilogue warp: two 32 x 16 fp32 tiles (residual in / result out) + one bf16 tile
static constexpr int LN_SLOTS = 12;                            // LayerNorm-fold partial sums per row: 4 column tiles x 3 epilogue column groups
static constexpr int LN_STAT_BYTES = 2 * 4 * 3 * 32 * 4;      // fused-LN epilogue: per quadrant, per column group, 32 row partials x {sum, sumsq}
template <int BN, int ST = 0>     // ST != 0: stage-count override of the development probes (no epilogue staging above 5)
struct Cfg {
    static constexpr int STAGES = ST != 0 ? ST : (BN == 256 ? 5 : (BN == 128 ? 7 : 8));
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = (BN / 2) * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int TMEM_COLS = BN == 64 ? 256 : 2 * BN;      // BN = 64: room for two 128-column accumulators (cat mode)
    static constexpr int BAR_BYTES = (2 * STAGES + 8) * 8 + 16;
    static constexpr int EPI_OFF = STAGES * STAGE_BYTES + ((BAR_BYTES + 15) / 16) * 16;
    static constexpr int EPI_BYTES = ST > 5 ? 0 : NUM_EPI_WARPS * 32 * EPI_PITCH;   // one 32 x 32 fp32 chunk per epilogue warp
    static constexpr int LN_OFF = EPI_OFF + EPI_BYTES;
    static constexpr int SMEM = EPI_OFF + EPI_BYTES + 1024;               // + alignment slack
    static constexpr int SMEM_LN = SMEM + LN_STAT_BYTES;
    static constexpr int EPI5_OFF = (EPI_OFF + 1023) & ~1023;             // swizzled TMA tiles: 512 B atoms
    static constexpr int EPI5_BAR_OFF = EPI5_OFF + NUM_EPI_WARPS * EPI5_WARP_BYTES;
    static constexpr int SMEM_EPI5 = EPI5_BAR_OFF + NUM_EPI_WARPS * 2 * 8 + 1024;
};

__device__ __forceinline__ int floordiv(int a, int b) { int q = a / b; return (a % b != 0 && ((a < 0) != (b < 0))) ? q - 1 : q; }

// Epilogue variants (compile-time): what is applied to / written from the accumulator.  EPI_ANY keeps every test at run
// time (used by the less frequent launches: feature projection, fp32-mode planes, positional conv).
enum : int { EPI_GELU = 1, EPI_RESID = 2, EPI_OUT32 = 4, EPI_OUT16 = 8, EPI_PLANE = 16, EPI_ZROW = 32, EPI_ANY = 64 };

// The it-th tile of CTA pair `pair`.  Default: tiles (m-tile major, n-tile minor) strided over the pairs.  ln_mode: the pair walks
// M-tiles and does both 256-column halves of each back to back, so that accumulator stages 0 / 1 hold the two halves of one
// 512-wide row block when the fused LayerNorm epilogue needs the whole row.
__device__ __forceinline__ bool tile_at(const GemmParams& p, int it, int pair, int npairs, int& n_tile, int& mt) {
    if (p.ln_mode) {
        mt = pair + (it >> 1) * npairs;
        n_tile = it & 1;
        return mt < p.num_m_tiles;
    }
    const int tile = pair + it * npairs;
    if (tile >= p.num_tiles) return false;
    n_tile = tile % p.tiles_n;
    mt = tile / p.tiles_n;
    return true;
}

// One epilogue warp: TMEM lane quadrant `warp & 3` (32 accumulator rows), every second 32-column chunk of each tile.
// thread = accumulator row after tcgen05.ld: bias (+GELU) there, then the 32 x 32 chunk goes through shared memory so
// that global accesses are row-contiguous (the row-per-lane form touched 32 cache lines per instruction).
// MODE: 0 plain, 1 residual prefetch (out-proj / FC2), 2 = 1 + LayerNorm-fold producer (bf16 copy + per-row partial sums),
//       4 LayerNorm-fold consumer (QKV / FC1: v = rstd * (acc - mean * s[n]) + c[n]; `stat` = [4 quadrants][32 rows] float2 in smem)
template <int BN, int F, int MODE = 0>
__device__ __forceinline__ void epilogue_tiles(const GemmParams& p, uint64_t* tfull, uint64_t* tempty, uint32_t tmem_base,
                                               uint8_t* epi_smem, uint32_t rank, int pair, int npairs, float* stat = nullptr) {
    constexpr bool PF = MODE == 1 || MODE == 2;
    constexpr bool FOLD_OUT = MODE == 2;
    constexpr bool FOLD_IN = MODE == 4;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int quad = warp & 3;
    const int cg = (warp - 4) >> 2;
    const bool gelu = (F & EPI_ANY) ? p.gelu != 0 : (F & EPI_GELU) != 0;
    const bool has_resid = (F & EPI_ANY) ? p.resid != nullptr : (F & EPI_RESID) != 0;
    const bool has32 = (F & EPI_ANY) ? p.out32 != nullptr : (F & EPI_OUT32) != 0;
    const bool has16 = (F & EPI_ANY) ? p.out16 != nullptr : (F & EPI_OUT16) != 0;
    const bool plane = (F & EPI_ANY) ? p.out16_plane != 0 : false;
    constexpr bool WIDE16 = F == EPI_OUT16 || F == (EPI_OUT16 | EPI_GELU);   // bf16-only output: 8 columns (16 B) per lane
    // read-back mapping of a staged chunk: WIDE16: lane -> row (lane>>2) + 8j, columns 8*(lane&3)..+7
    //                                      else  : lane -> row (lane>>3) + 4j, columns 4*(lane&7)..+3
    const int sub_r = WIDE16 ? (lane >> 2) : (lane >> 3);
    const int c0 = WIDE16 ? (lane & 3) * 8 : (lane & 7) * 4;
    const uint8_t* epi_rd = epi_smem + sub_r * EPI_PITCH + c0 * 4;
    float4* epi_wr = reinterpret_cast<float4*>(epi_smem + lane * EPI_PITCH);
    const long long ldo = p.ldo;
    const int r_warp0 = (int)rank * BM + quad * 32;
    const uint32_t tempty_leader0 = mapa_u32(smem_u32(&tempty[0]), 0), tempty_leader1 = mapa_u32(smem_u32(&tempty[1]), 0);   // (no array: a dynamic index puts it in local memory)
    int it = 0;
    for (int tile = pair; tile < p.num_tiles; tile += npairs, ++it) {
        const int as = it & 1;
        const uint32_t aphase = (uint32_t)(it >> 1) & 1u;
        const int n_tile = tile % p.tiles_n;
        const int mt = tile / p.tiles_n;
        const int b = mt / p.tiles_m_per_batch;
        const int t_base = (mt - b * p.tiles_m_per_batch) * (2 * BM) + r_warp0;   // first row of this warp's 32
        const int lim = p.rows_per_batch - t_base - sub_r;       // staged rows rr < lim are inside the batch item
        int zlim = 0x3fffffff;                                   // staged rows rr >= zlim are beyond the utterance: zeros
        if ((F & EPI_ANY) && p.row_valid != nullptr) zlim = __ldg(p.row_valid + b) - t_base - sub_r;
        const long long roff = ((long long)b * p.rows_per_batch + t_base + sub_r) * ldo + n_tile * BN + c0;
        const float* bias_t = p.bias + n_tile * BN;
        // Residual variant (out-proj / FC2: x <- x + y in the fp32 residual stream, in place): the residual of a chunk is
        // requested one chunk ahead -- the first one before the wait for the accumulator -- so its L2 / HBM latency runs under
        // the wait, the TMEM load and the staging instead of inside the store loop (profiles/r1b: the in-loop read made these
        // epilogues latency-bound).  A 128 B line of x is read and written by the same 8 lanes only.
        constexpr bool RESID_PF = PF && F == (EPI_OUT32 | EPI_RESID);
        float ps[FOLD_OUT ? 8 : 1], pq[FOLD_OUT ? 8 : 1];           // row sums of this warp's chunks: rows sub_r + 4j
        if constexpr (FOLD_OUT) {
#pragma unroll
            for (int j = 0; j < 8; ++j) { ps[j] = 0.f; pq[j] = 0.f; }
        }
        float f_mean = 0.f, f_rstd = 0.f;
        if constexpr (FOLD_IN) {
            // the row statistics depend only on the previous launch: fetched and shared among the quadrant's three warps BEFORE the
            // wait for the accumulator, i.e. under the mainloop.  Column group 0 reads the row's 12 partial {sum, sumsq} pairs.
            float2* st2 = reinterpret_cast<float2*>(stat) + quad * 32;
            asm volatile("bar.sync %0, 96;" ::"r"(1 + quad) : "memory");      // the previous tile's readers are done with the slot
            if (cg == 0) {
                const int row = t_base + lane;
                float sum = 0.f, sq = 0.f;
                if (row < p.rows_per_batch) {
                    const float4* sp = reinterpret_cast<const float4*>(p.stat_in + ((long long)b * p.rows_per_batch + row) * (2 * LN_SLOTS));
#pragma unroll
                    for (int q = 0; q < LN_SLOTS / 2; ++q) { const float4 v = __ldg(sp + q); sum += v.x; sq += v.y; sum += v.z; sq += v.w; }
                }
                const float mean = sum * (1.0f / (float)p.N_stat);
                const float var = fmaxf(sq * (1.0f / (float)p.N_stat) - mean * mean, 0.f);
                st2[lane] = make_float2(mean, rsqrtf(var + 1e-5f));
            }
            asm volatile("bar.sync %0, 96;" ::"r"(1 + quad) : "memory");
            const float2 mr = st2[lane];
            f_mean = mr.x; f_rstd = mr.y;
        }
        float4 qn[RESID_PF ? 8 : 1];
        if constexpr (RESID_PF) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (4 * j < lim) qn[j] = *reinterpret_cast<const float4*>(p.resid + roff + cg * 32 + (long long)(4 * j) * ldo);
        }
        mbar_wait_cl(&tfull[as], aphase, 4);
        tc_fence_after();
        const uint32_t tcol = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(as * BN);
        for (int c = cg; c < BN / 32 && !(p.probe & 4); c += NUM_EPI_WARPS / 4) {
            uint32_t r[32];
            tmem_ld32(tcol + (uint32_t)(c * 32), r);
            float4 qc[RESID_PF ? 8 : 1];
            if constexpr (RESID_PF) {
#pragma unroll
                for (int j = 0; j < 8; ++j) qc[j] = qn[j];
                const int cn = c + NUM_EPI_WARPS / 4;
                if (cn < BN / 32) {
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        if (4 * j < lim) qn[j] = *reinterpret_cast<const float4*>(p.resid + roff + cn * 32 + (long long)(4 * j) * ldo);
                }
            }
            tmem_ld_wait();
            {
                const float4* bp = reinterpret_cast<const float4*>(bias_t + c * 32);
                const float4* cp = reinterpret_cast<const float4*>(p.colsum + (FOLD_IN ? n_tile * BN + c * 32 : 0));
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float4 bb = __ldg(bp + j);
                    float4 v;
                    if constexpr (FOLD_IN) {
                        const float4 ss = __ldg(cp + j);
                        v = make_float4(fmaf(fmaf(-f_mean, ss.x, __uint_as_float(r[4 * j])), f_rstd, bb.x),
                                        fmaf(fmaf(-f_mean, ss.y, __uint_as_float(r[4 * j + 1])), f_rstd, bb.y),
                                        fmaf(fmaf(-f_mean, ss.z, __uint_as_float(r[4 * j + 2])), f_rstd, bb.z),
                                        fmaf(fmaf(-f_mean, ss.w, __uint_as_float(r[4 * j + 3])), f_rstd, bb.w));
                    } else {
                        v = make_float4(__uint_as_float(r[4 * j]) + bb.x, __uint_as_float(r[4 * j + 1]) + bb.y,
                                        __uint_as_float(r[4 * j + 2]) + bb.z, __uint_as_float(r[4 * j + 3]) + bb.w);
                    }
                    if (gelu) { gelu_erf_fast2(v.x, v.y); gelu_erf_fast2(v.z, v.w); }
                    epi_wr[j] = v;
                }
            }
            __syncwarp();
            const long long coff = roff + c * 32;
            if constexpr (WIDE16) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float4 x0 = *reinterpret_cast<const float4*>(epi_rd + j * (8 * EPI_PITCH));
                    const float4 x1 = *reinterpret_cast<const float4*>(epi_rd + j * (8 * EPI_PITCH) + 16);
                    if (8 * j < lim)
                        *reinterpret_cast<uint4*>(p.out16 + coff + (long long)(8 * j) * ldo) =
                            make_uint4(pack_bf16(x0.x, x0.y), pack_bf16(x0.z, x0.w), pack_bf16(x1.x, x1.y), pack_bf16(x1.z, x1.w));
                }
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float4 x = *reinterpret_cast<const float4*>(epi_rd + j * (4 * EPI_PITCH));
                    if (4 * j < lim) {
                        const long long off = coff + (long long)(4 * j) * ldo;
                        if constexpr (RESID_PF) {
                            x.x += qc[j].x; x.y += qc[j].y; x.z += qc[j].z; x.w += qc[j].w;
                        } else if (has_resid) {
                            const float4 q = *reinterpret_cast<const float4*>(p.resid + off);
                            x.x += q.x; x.y += q.y; x.z += q.z; x.w += q.w;
                        }
                        if ((F & EPI_ANY) && 4 * j >= zlim) x = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (has32) {
                            *reinterpret_cast<float4*>(p.out32 + off) = x;
                            if (p.out32_copy != nullptr) *reinterpret_cast<float4*>(p.out32_copy + off) = x;
                        }
                        if constexpr (FOLD_OUT) {
                            ps[j] += (x.x + x.y) + (x.z + x.w);
                            pq[j] = fmaf(x.x, x.x, fmaf(x.y, x.y, fmaf(x.z, x.z, fmaf(x.w, x.w, pq[j]))));
                            *reinterpret_cast<uint2*>(p.x16_out + off) = make_uint2(pack_bf16(x.x, x.y), pack_bf16(x.z, x.w));
                        }
                        if (has16) {
                            const uint32_t h0 = pack_bf16(x.x, x.y), h1 = pack_bf16(x.z, x.w);
                            bf16* op = p.out16 + off;
                            *reinterpret_cast<uint2*>(op) = make_uint2(h0, h1);
                            if (plane) {
                                const __nv_bfloat162 a0 = *reinterpret_cast<const __nv_bfloat162*>(&h0);
                                const __nv_bfloat162 a1 = *reinterpret_cast<const __nv_bfloat162*>(&h1);
                                *reinterpret_cast<uint2*>(op + p.out16_plane) =
                                    make_uint2(pack_bf16(x.x - __low2float(a0), x.y - __high2float(a0)),
                                               pack_bf16(x.z - __low2float(a1), x.w - __high2float(a1)));
                            }
                        }
                    }
                }
            }
            __syncwarp();      // chunk buffer is rewritten by the next column chunk
        }
        tc_fence_before();
        __syncwarp();
        // this accumulator stage may be overwritten once all 16 epilogue warps of the pair have drained it
        if (lane == 0) mbar_arrive_cluster(as ? tempty_leader1 : tempty_leader0);
        if constexpr (FOLD_OUT) {
            // row sums over this warp's columns: the 8 lanes that share a row (same lane >> 3) combine, lane & 7 == 0 writes the
            // row's slot (column tile, column group) -- after the accumulator stage has been handed back
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float a = ps[j], q = pq[j];
                a += __shfl_xor_sync(0xffffffffu, a, 1); q += __shfl_xor_sync(0xffffffffu, q, 1);
                a += __shfl_xor_sync(0xffffffffu, a, 2); q += __shfl_xor_sync(0xffffffffu, q, 2);
                a += __shfl_xor_sync(0xffffffffu, a, 4); q += __shfl_xor_sync(0xffffffffu, q, 4);
                if ((lane & 7) == 0 && 4 * j < lim)
                    *reinterpret_cast<float2*>(p.stat_out + ((long long)b * p.rows_per_batch + t_base + sub_r + 4 * j) * (2 * LN_SLOTS) +
                                               (n_tile * 3 + cg) * 2) = make_float2(a, q);
            }
        }
    }
}

// conv2d epilogue of the AASIST residual blocks (GemmDesc::conv_mode): N = 64 channel-last, so thread = accumulator row
// already reads / writes 64 contiguous bytes per plane; no staging.  See occ_internal.h for the arithmetic.
template <int BN>
__device__ __forceinline__ void epilogue_conv_tiles(const GemmParams& p, uint64_t* tfull, uint64_t* tempty, uint32_t tmem_base,
                                                    uint32_t rank, int pair, int npairs) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int quad = warp & 3;
    const int cg = (warp - 4) >> 2;
    const int r_warp0 = (int)rank * BM + quad * 32;
    const uint32_t tempty_leader0 = mapa_u32(smem_u32(&tempty[0]), 0), tempty_leader1 = mapa_u32(smem_u32(&tempty[1]), 0);   // (no array: a dynamic index puts it in local memory)
    int it = 0;
    for (int tile = pair; tile < p.num_tiles; tile += npairs, ++it) {
        const int as = it & 1;
        const uint32_t aphase = (uint32_t)(it >> 1) & 1u;
        const int n_tile = tile % p.tiles_n;
        const int mt = tile / p.tiles_n;
        const int b = mt / p.tiles_m_per_batch;
        const int t = (mt - b * p.tiles_m_per_batch) * (2 * BM) + r_warp0 + lane;      // padded-grid position of this row
        const bool in_batch = t < p.rows_per_batch;
        const int hp = t / p.conv_wp, wp = t - hp * p.conv_wp;
        const int wl = p.conv_wlen != nullptr ? __ldg(p.conv_wlen + b) : p.conv_wm;
        const bool valid = in_batch && hp >= p.conv_h0 && hp < p.conv_h1 && wp >= 1 && wp <= wl;
        const long long orow = ((long long)b * p.rows_per_batch + t) * p.ldo + n_tile * BN;
        mbar_wait_cl(&tfull[as], aphase, 4);
        tc_fence_after();
        const uint32_t tcol = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(as * (p.cat ? 2 * BN : BN));
        for (int c = cg; c < BN / 32; c += NUM_EPI_WARPS / 4) {
            uint32_t r[32];
            tmem_ld32(tcol + (uint32_t)(c * 32), r);
            tmem_ld_wait();
            if (p.probe & 0x100) continue;
            if (p.cat) {       // + the A_hi x W_lo block of the accumulator
                uint32_t r2[32];
                tmem_ld32(tcol + (uint32_t)(BN + c * 32), r2);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; ++j) r[j] = __float_as_uint(__uint_as_float(r[j]) + __uint_as_float(r2[j]));
            }
            if (!in_batch) continue;
            const int n = n_tile * BN + c * 32;
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = 0.f;
            if (valid) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) + __ldg(p.bias + n + j);
                if (p.cs1 != nullptr) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = selu_f(fmaf(v[j], __ldg(p.cs1 + n + j), __ldg(p.ct1 + n + j)));
                }
                if (p.cresid != nullptr) {
                    const uint4* rh = reinterpret_cast<const uint4*>(p.cresid + orow + c * 32);
                    const uint4* rl = reinterpret_cast<const uint4*>(p.cresid + p.cresid_plane + orow + c * 32);
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const uint4 a = rh[q], l = rl[q];
                        const uint32_t ah[4] = {a.x, a.y, a.z, a.w}, al[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const __nv_bfloat162 x = *reinterpret_cast<const __nv_bfloat162*>(&ah[e]);
                            const __nv_bfloat162 y = *reinterpret_cast<const __nv_bfloat162*>(&al[e]);
                            v[8 * q + 2 * e] += __low2float(x) + __low2float(y);
                            v[8 * q + 2 * e + 1] += __high2float(x) + __high2float(y);
                        }
                    }
                }
                if (p.cs2 != nullptr) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = selu_f(fmaf(v[j], __ldg(p.cs2 + n + j), __ldg(p.ct2 + n + j)));
                }
            }
            if (p.out32 != nullptr) {
                float4* op = reinterpret_cast<float4*>(p.out32 + orow + c * 32);
#pragma unroll
                for (int q = 0; q < 8; ++q) op[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
            }
            if (p.out16 != nullptr) {
                uint32_t h[16], l[16];
#pragma unroll
                for (int q = 0; q < 16; ++q) {
                    h[q] = pack_bf16(v[2 * q], v[2 * q + 1]);
                    const __nv_bfloat162 hh = *reinterpret_cast<const __nv_bfloat162*>(&h[q]);
                    l[q] = pack_bf16(v[2 * q] - __low2float(hh), v[2 * q + 1] - __high2float(hh));
                }
                uint4* oh = reinterpret_cast<uint4*>(p.out16 + orow + c * 32);
                uint4* ol = reinterpret_cast<uint4*>(p.out16 + p.out16_plane + orow + c * 32);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    oh[q] = make_uint4(h[4 * q], h[4 * q + 1], h[4 * q + 2], h[4 * q + 3]);
                    ol[q] = make_uint4(l[4 * q], l[4 * q + 1], l[4 * q + 2], l[4 * q + 3]);
                }
            }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(as ? tempty_leader1 : tempty_leader0);
    }
}

// Fused LayerNorm(512) + erf-GELU epilogue of the conv feature encoder (GemmDesc::ln_gamma; torchaudio components.py:90-93).
// Accumulator stage 0 / 1 hold columns 0..255 / 256..511 of the SAME 256 rows (tile_at, ln_mode).  thread = accumulator row;
// the three warps that share a TMEM lane quadrant split the sixteen 32-column chunks of the row (6 / 6 / 4):
//   A  one pass over the warp's chunks of v = acc + bias: shifted sums  S1 = sum (v - p), S2 = sum (v - p)^2  around the pivot
//      p = the warp's first value of the row (no cancellation whatever the row mean), giving the partial mean and centred second
//      moment (n_i, m_i, M2_i); the three partials are combined through shared memory with the pairwise update of Chan et al.
//      (mean = sum n_i m_i / 512, M2 = sum M2_i + sum n_i (m_i - mean)^2; biased variance, eps 1e-5 as F.layer_norm).
//      Stage 0's chunks are summed as soon as stage 0 is complete, i.e. under stage 1's MMAs.
//   B  y = gelu((v - mean) * rstd * g + b) -> bf16, staged through shared memory, 16 B row-contiguous stores; stage 0 goes back
//      to the MMA warp after its half, so the next tile's MMAs run under stage 1's half of pass B.
// Tensor-memory loads are software-pipelined (the next chunk is requested before the current one is processed).
__device__ __forceinline__ void epilogue_ln_tiles(const GemmParams& p, uint64_t* tfull, uint64_t* tempty, uint32_t tmem_base,
                                                  uint8_t* epi_smem, float* stat, uint32_t rank, int pair, int npairs) {
    constexpr int BN = 256;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int quad = warp & 3;
    const int cg = (warp - 4) >> 2;                       // 0..2: this warp's chunks of a half are cg, cg + 3, cg + 6 (< 8)
    const int nc = cg < 2 ? 3 : 2;
    float* st_m = stat + quad * 96;                       // [3][32] partial means, per quadrant
    float* st_q = stat + 384 + quad * 96;                 // [3][32] partial centred second moments
    const int sub_r = lane >> 2, c0 = (lane & 3) * 8;     // read-back mapping of a staged chunk: row (lane>>2) + 8j, 8 columns
    const uint8_t* epi_rd = epi_smem + sub_r * EPI_PITCH + c0 * 4;
    float4* epi_wr = reinterpret_cast<float4*>(epi_smem + lane * EPI_PITCH);
    const long long ldo = p.ldo;
    const int r_warp0 = (int)rank * BM + quad * 32;
    const uint32_t tempty_leader0 = mapa_u32(smem_u32(&tempty[0]), 0), tempty_leader1 = mapa_u32(smem_u32(&tempty[1]), 0);
    const uint32_t trow = tmem_base + ((uint32_t)(quad * 32) << 16);
    const int bar_id = 1 + quad;                          // named barrier of the quadrant's 3 warps (0 is __syncthreads)
    const float n_own = (float)(64 * nc);                 // values of the row this warp sums: 2 halves x nc chunks x 32
    for (int it2 = 0;; ++it2) {
        const int mt = pair + it2 * npairs;
        if (mt >= p.num_m_tiles) break;
        const uint32_t aphase = (uint32_t)it2 & 1u;
        const int b = mt / p.tiles_m_per_batch;
        const int t_base = (mt - b * p.tiles_m_per_batch) * (2 * BM) + r_warp0;
        const int lim = p.rows_per_batch - t_base - sub_r;
        const long long roff = ((long long)b * p.rows_per_batch + t_base + sub_r) * ldo + c0;
        uint32_t ra[32], rb[32];
        // ---- A: shifted sums over this warp's chunks (stage 0 first: its MMAs finish while stage 1's are still running)
        float pivot = 0.f;
        uint64_t s1 = f2_splat(0.f), s2 = f2_splat(0.f);
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            mbar_wait_cl(&tfull[half], aphase, 4);
            tc_fence_after();
            tmem_ld32(trow + (uint32_t)(half * BN + cg * 32), ra);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                uint32_t (&cur)[32] = (i & 1) ? rb : ra;
                uint32_t (&nxt)[32] = (i & 1) ? ra : rb;
                if (i + 1 < nc) tmem_ld32(trow + (uint32_t)(half * BN + (cg + 3 * (i + 1)) * 32), nxt);
                if (i < nc) {
                    const float4* bp = reinterpret_cast<const float4*>(p.bias + half * BN + (cg + 3 * i) * 32);
                    if (half == 0 && i == 0) pivot = __uint_as_float(cur[0]) + __ldg(p.bias + cg * 32);
                    const uint64_t np2 = f2_splat(-pivot);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float4 bb = __ldg(bp + j);
                        const uint64_t d0 = f2_add(f2_add(f2_pack(__uint_as_float(cur[4 * j]), __uint_as_float(cur[4 * j + 1])), f2_pack(bb.x, bb.y)), np2);
                        const uint64_t d1 = f2_add(f2_add(f2_pack(__uint_as_float(cur[4 * j + 2]), __uint_as_float(cur[4 * j + 3])), f2_pack(bb.z, bb.w)), np2);
                        s1 = f2_add(s1, f2_add(d0, d1));
                        s2 = f2_fma(d0, d0, s2);
                        s2 = f2_fma(d1, d1, s2);
                    }
                }
                tmem_ld_wait();
            }
        }
        {
            float a, c, q0, q1;
            f2_unpack(s1, a, c);
            f2_unpack(s2, q0, q1);
            const float S1 = a + c, S2 = q0 + q1;
            st_m[cg * 32 + lane] = pivot + S1 / n_own;                    // partial mean
            st_q[cg * 32 + lane] = S2 - S1 * S1 / n_own;                  // partial centred second moment
        }
        asm volatile("bar.sync %0, 96;" ::"r"(bar_id) : "memory");
        float mean, rstd;
        {
            const float m0 = st_m[lane], m1 = st_m[32 + lane], m2 = st_m[64 + lane];
            mean = (192.f * m0 + 192.f * m1 + 128.f * m2) * (1.0f / 512.0f);
            const float e0 = m0 - mean, e1 = m1 - mean, e2 = m2 - mean;
            const float M2 = st_q[lane] + st_q[32 + lane] + st_q[64 + lane] + 192.f * e0 * e0 + 192.f * e1 * e1 + 128.f * e2 * e2;
            rstd = rsqrtf(M2 * (1.0f / 512.0f) + 1e-5f);
        }
        const uint64_t nm2 = f2_splat(-mean), rs2 = f2_splat(rstd);
        // ---- B: normalise, affine, GELU, bf16 store
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            tmem_ld32(trow + (uint32_t)(half * BN + cg * 32), ra);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                uint32_t (&cur)[32] = (i & 1) ? rb : ra;
                uint32_t (&nxt)[32] = (i & 1) ? ra : rb;
                if (i + 1 < nc) tmem_ld32(trow + (uint32_t)(half * BN + (cg + 3 * (i + 1)) * 32), nxt);
                if (i < nc) {
                    const int col = half * BN + (cg + 3 * i) * 32;
                    const float4* bp = reinterpret_cast<const float4*>(p.bias + col);
                    const float4* gp = reinterpret_cast<const float4*>(p.ln_g + col);
                    const float4* ep = reinterpret_cast<const float4*>(p.ln_b + col);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float4 bb = __ldg(bp + j), gg = __ldg(gp + j), ee = __ldg(ep + j);
                        float4 v;
                        f2_unpack(f2_fma(f2_mul(f2_add(f2_add(f2_pack(__uint_as_float(cur[4 * j]), __uint_as_float(cur[4 * j + 1])), f2_pack(bb.x, bb.y)), nm2), rs2),
                                         f2_pack(gg.x, gg.y), f2_pack(ee.x, ee.y)), v.x, v.y);
                        f2_unpack(f2_fma(f2_mul(f2_add(f2_add(f2_pack(__uint_as_float(cur[4 * j + 2]), __uint_as_float(cur[4 * j + 3])), f2_pack(bb.z, bb.w)), nm2), rs2),
                                         f2_pack(gg.z, gg.w), f2_pack(ee.z, ee.w)), v.z, v.w);
                        gelu_erf_fast2(v.x, v.y);
                        gelu_erf_fast2(v.z, v.w);
                        epi_wr[j] = v;
                    }
                    __syncwarp();
                    const long long coff = roff + col;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float4 x0 = *reinterpret_cast<const float4*>(epi_rd + j * (8 * EPI_PITCH));
                        const float4 x1 = *reinterpret_cast<const float4*>(epi_rd + j * (8 * EPI_PITCH) + 16);
                        if (8 * j < lim)
                            *reinterpret_cast<uint4*>(p.out16 + coff + (long long)(8 * j) * ldo) =
                                make_uint4(pack_bf16(x0.x, x0.y), pack_bf16(x0.z, x0.w), pack_bf16(x1.x, x1.y), pack_bf16(x1.z, x1.w));
                    }
                    __syncwarp();      // chunk buffer is rewritten by the next column chunk
                }
                tmem_ld_wait();
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(half ? tempty_leader1 : tempty_leader0);
        }
        // the quadrant's partial slots are rewritten by the next tile only after all three warps have read them
        asm volatile("bar.sync %0, 96;" ::"r"(bar_id) : "memory");
    }
}

// Residual epilogue on the TMA engine (MODE 5: out-proj / FC2, x <- x + (A W^T + b) in place on the fp32 residual stream).
// Everything stays in the accumulator's own layout (thread = row): per 32-row x 16-column chunk the warp
//   * has the residual tile brought into shared memory by a TMA load it issued one chunk ahead (64 B-swizzled: conflict-free
//     16 B accesses at one row per lane), double-buffered, the first one requested before the wait for the accumulator;
//   * adds accumulator + bias + residual in registers, keeps the row's running sum / sum of squares (the LayerNorm-fold
//     statistics cost two FMAs per element and no shuffles here), writes the fp32 result over the residual tile and, for the
//     LayerNorm fold, its bf16 copy into a third tile;
//   * hands the tiles to TMA stores (UTMASTG): the residual stream itself, the optional hidden-state copy, the optional bf16
//     operand of the next GEMM.  Rows beyond the matrix are clipped by the tensor maps -- no per-lane address arithmetic, no
//     transposing read-back through shared memory, no global load / store instructions at all in this epilogue.
struct ChunkPos { int tile, c; };
template <int BN>
__device__ __forceinline__ void epilogue_resid_tma(const GemmParams& p, const EpiMaps& em, uint64_t* tfull, uint64_t* tempty, uint32_t tmem_base,
                                                   uint8_t* wbuf, uint64_t* lbar, uint32_t rank, int pair, int npairs) {
    constexpr int CW = 16, NCH = BN / CW;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int quad = warp & 3;
    const int cg = (warp - 4) >> 2;
    const uint32_t tempty_leader0 = mapa_u32(smem_u32(&tempty[0]), 0), tempty_leader1 = mapa_u32(smem_u32(&tempty[1]), 0);
    const int swz = (lane >> 1) & 3;          // SWIZZLE_64B: 16-byte piece index ^= (row >> 1) & 3   (rows of 64 B)
    const int swh = (lane >> 2) & 1;          // SWIZZLE_32B: 16-byte piece index ^= (row >> 2) & 1   (rows of 32 B)
    uint8_t* hb = wbuf + 4096;
    const bool has16 = p.x16_out != nullptr, has_copy = p.out32_copy != nullptr, has_stat = p.stat_out != nullptr;
    auto row_of = [&](int tile) { return (tile / p.tiles_n) * (2 * BM) + (int)rank * BM + quad * 32; };     // nbatch == 1
    auto load = [&](const ChunkPos& q, int buf) {
        if (lane == 0) {
            mbar_arrive_expect_tx(&lbar[buf], 2048u);
            tma_load_2d(&em.x32, &lbar[buf], wbuf + buf * 2048, (q.tile % p.tiles_n) * BN + q.c * CW, row_of(q.tile));
        }
    };
    ChunkPos cur{pair, cg};
    bool valid = cur.tile < p.num_tiles;
    if (valid) load(cur, 0);
    int q = 0, it = 0;
    float s1 = 0.f, s2 = 0.f;
    while (valid) {
        const int as = it & 1;
        if (cur.c == cg) {           // first chunk of a tile
            mbar_wait_cl(&tfull[as], (uint32_t)(it >> 1) & 1u, 4);
            tc_fence_after();
            s1 = 0.f; s2 = 0.f;
        }
        ChunkPos nxt{cur.tile, cur.c + 3};
        if (nxt.c >= NCH) { nxt.tile += npairs; nxt.c = cg; }
        const bool nvalid = nxt.tile < p.num_tiles;
        // every earlier store of this warp has finished reading shared memory: the other fp32 tile and the bf16 tile are free
        if (lane == 0) tma_store_wait_read0();
        __syncwarp();
        if (nvalid) load(nxt, (q + 1) & 1);
        const int n_tile = cur.tile % p.tiles_n;
        const int col = n_tile * BN + cur.c * CW;
        uint32_t r[16];
        tmem_ld16(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(as * BN + cur.c * CW), r);
        uint8_t* ra = wbuf + (q & 1) * 2048 + lane * 64;
        const float4* bp = reinterpret_cast<const float4*>(p.bias + col);
        float4 bb[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) bb[j] = __ldg(bp + j);
        tmem_ld_wait();
        mbar_wait(&lbar[q & 1], (uint32_t)(q >> 1) & 1u, 9);
        float4 x[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float4 rs = *reinterpret_cast<const float4*>(ra + ((j ^ swz) << 4));
            x[j] = make_float4(__uint_as_float(r[4 * j]) + bb[j].x + rs.x, __uint_as_float(r[4 * j + 1]) + bb[j].y + rs.y,
                               __uint_as_float(r[4 * j + 2]) + bb[j].z + rs.z, __uint_as_float(r[4 * j + 3]) + bb[j].w + rs.w);
            s1 += (x[j].x + x[j].y) + (x[j].z + x[j].w);
            s2 = fmaf(x[j].x, x[j].x, fmaf(x[j].y, x[j].y, fmaf(x[j].z, x[j].z, fmaf(x[j].w, x[j].w, s2))));
            *reinterpret_cast<float4*>(ra + ((j ^ swz) << 4)) = x[j];
        }
        if (has16) {
            uint8_t* hr = hb + lane * 32;
#pragma unroll
            for (int j = 0; j < 2; ++j)
                *reinterpret_cast<uint4*>(hr + ((j ^ swh) << 4)) =
                    make_uint4(pack_bf16(x[2 * j].x, x[2 * j].y), pack_bf16(x[2 * j].z, x[2 * j].w),
                               pack_bf16(x[2 * j + 1].x, x[2 * j + 1].y), pack_bf16(x[2 * j + 1].z, x[2 * j + 1].w));
        }
        fence_proxy_async();          // generic-proxy writes to the tiles are visible to the TMA engine
        __syncwarp();
        const int row0 = row_of(cur.tile);
        if (lane == 0) {
            tma_store_2d(&em.x32, wbuf + (q & 1) * 2048, col, row0);
            if (has_copy) tma_store_2d(&em.copy32, wbuf + (q & 1) * 2048, col, row0);
            if (has16) tma_store_2d(&em.x16, hb, col, row0);
            tma_store_commit();
        }
        if (nxt.tile != cur.tile || !nvalid) {          // last chunk of this tile
            if (has_stat && row0 + lane < p.rows_per_batch)
                *reinterpret_cast<float2*>(p.stat_out + (long long)(row0 + lane) * (2 * LN_SLOTS) + (n_tile * 3 + cg) * 2) = make_float2(s1, s2);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(as ? tempty_leader1 : tempty_leader0);
            ++it;
        }
        cur = nxt; valid = nvalid; ++q;
    }
    if (lane == 0) tma_store_wait0();       // the tiles (and the global writes) outlive the loop
    __syncwarp();
}

// bf16-output epilogue on the TMA engine (MODE 6: QKV / FC1): bias (or the LayerNorm fold: v = rstd * (acc - mean * s[n]) + c[n]),
// optional erf-GELU, all in the accumulator's own layout (thread = row); the 32 x 32 bf16 chunk goes into a 64 B-swizzled
// shared-memory tile (4 conflict-free 16 B stores per lane) and leaves through one TMA store -- two tiles per warp, so the
// store of chunk i drains while chunk i+1 is computed.  `stat`: [4 quadrants][32 rows] float2 (fold only).
template <int BN, bool GELU>
__device__ __forceinline__ void epilogue_bf16_tma(const GemmParams& p, const EpiMaps& em, uint64_t* tfull, uint64_t* tempty, uint32_t tmem_base,
                                                  uint8_t* wbuf, float* stat, uint32_t rank, int pair, int npairs) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int quad = warp & 3;
    const int cg = (warp - 4) >> 2;
    const bool fold = p.stat_in != nullptr;
    const uint32_t tempty_leader0 = mapa_u32(smem_u32(&tempty[0]), 0), tempty_leader1 = mapa_u32(smem_u32(&tempty[1]), 0);
    const int swz = (lane >> 1) & 3;          // SWIZZLE_64B
    int it = 0, q = 0;
    for (int tile = pair; tile < p.num_tiles; tile += npairs, ++it) {
        const int as = it & 1;
        const int n_tile = tile % p.tiles_n;
        const int mt = tile / p.tiles_n;
        const int row0 = mt * (2 * BM) + (int)rank * BM + quad * 32;            // nbatch == 1
        float f_mean = 0.f, f_rstd = 0.f;
        if (fold) {
            // row statistics of the previous launch, shared among the quadrant's three warps, fetched under the mainloop
            float2* st2 = reinterpret_cast<float2*>(stat) + quad * 32;
            asm volatile("bar.sync %0, 96;" ::"r"(1 + quad) : "memory");
            if (cg == 0) {
                const int row = row0 + lane;
                float sum = 0.f, sq = 0.f;
                if (row < p.rows_per_batch) {
                    const float4* sp = reinterpret_cast<const float4*>(p.stat_in + (long long)row * (2 * LN_SLOTS));
#pragma unroll
                    for (int k = 0; k < LN_SLOTS / 2; ++k) { const float4 v = __ldg(sp + k); sum += v.x; sq += v.y; sum += v.z; sq += v.w; }
                }
                const float mean = sum * (1.0f / (float)p.N_stat);
                const float var = fmaxf(sq * (1.0f / (float)p.N_stat) - mean * mean, 0.f);
                st2[lane] = make_float2(mean, rsqrtf(var + 1e-5f));
            }
            asm volatile("bar.sync %0, 96;" ::"r"(1 + quad) : "memory");
            const float2 mr = st2[lane];
            f_mean = mr.x; f_rstd = mr.y;
        }
        const uint64_t f_nmean2 = f2_splat(-f_mean), f_rstd2 = f2_splat(f_rstd);
        mbar_wait_cl(&tfull[as], (uint32_t)(it >> 1) & 1u, 4);
        tc_fence_after();
        const uint32_t tcol = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(as * BN);
        // this warp's chunks cg, cg + 3, cg + 6 (< BN / 32): the next chunk's accumulator is requested from tensor memory before the
        // current one is processed (two register buffers)
        constexpr int NC_MAX = (BN / 32 + 2) / 3;
        const int nc = (BN / 32 - cg + 2) / 3;
        uint32_t ra[32], rb[32];
        tmem_ld32(tcol + (uint32_t)(cg * 32), ra);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < NC_MAX; ++i) {
            uint32_t (&r)[32] = (i & 1) ? rb : ra;
            uint32_t (&rn)[32] = (i & 1) ? ra : rb;
            if (i + 1 < nc) tmem_ld32(tcol + (uint32_t)((cg + 3 * (i + 1)) * 32), rn);
            if (i < nc) {
            const int c = cg + 3 * i;
            // the tile written two chunks ago has been read by its store
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            __syncwarp();
            const int col = n_tile * BN + c * 32;
            const float4* bp = reinterpret_cast<const float4*>(p.bias + col);
            const float4* cp = reinterpret_cast<const float4*>(p.colsum + (fold ? col : 0));
            uint8_t* tb = wbuf + (q & 1) * 2048 + lane * 64;
#pragma unroll
            for (int j = 0; j < 4; ++j) {           // 8 columns = one 16 B piece of the bf16 row
                // adjacent column pairs on the packed fp32x2 pipe (same operations, per lane IEEE-identical to the scalar form)
                uint32_t o[4];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const float4 bb = __ldg(bp + 2 * j + h);
                    uint64_t v01 = f2_pack(__uint_as_float(r[8 * j + 4 * h]), __uint_as_float(r[8 * j + 4 * h + 1]));
                    uint64_t v23 = f2_pack(__uint_as_float(r[8 * j + 4 * h + 2]), __uint_as_float(r[8 * j + 4 * h + 3]));
                    if (fold) {
                        const float4 ss = __ldg(cp + 2 * j + h);
                        v01 = f2_fma(f2_fma(f_nmean2, f2_pack(ss.x, ss.y), v01), f_rstd2, f2_pack(bb.x, bb.y));
                        v23 = f2_fma(f2_fma(f_nmean2, f2_pack(ss.z, ss.w), v23), f_rstd2, f2_pack(bb.z, bb.w));
                    } else {
                        v01 = f2_add(v01, f2_pack(bb.x, bb.y));
                        v23 = f2_add(v23, f2_pack(bb.z, bb.w));
                    }
                    float y0, y1, y2, y3;
                    f2_unpack(v01, y0, y1);
                    f2_unpack(v23, y2, y3);
                    if (GELU) { gelu_erf_fast2(y0, y1); gelu_erf_fast2(y2, y3); }
                    o[2 * h] = pack_bf16(y0, y1);
                    o[2 * h + 1] = pack_bf16(y2, y3);
                }
                *reinterpret_cast<uint4*>(tb + ((j ^ swz) << 4)) = make_uint4(o[0], o[1], o[2], o[3]);
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
                tma_store_2d(&em.x16, wbuf + (q & 1) * 2048, col, row0);
                tma_store_commit();
            }
            ++q;
            }
            tmem_ld_wait();
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(as ? tempty_leader1 : tempty_leader0);
    }
    if (lane == 0) tma_store_wait0();
    __syncwarp();
}

// Toeplitz pos-conv epilogue (MODE 7, GemmDesc::toep_T): bias + erf-GELU in the accumulator's layout, chunk staged through shared
// memory, then each staged row (= 4 frames of the flat padded sequence) scatters its 32 channels of frame 4*row + jj to the un-padded
// [B, T, 1024] output with 16 B stores (four rows x 64 B per store instruction); gap frames and rows past the batch are dropped.
template <int BN>
__device__ __forceinline__ void epilogue_toeplitz(const GemmParams& p, uint64_t* tfull, uint64_t* tempty, uint32_t tmem_base,
                                                  uint8_t* epi_smem, uint32_t rank, int pair, int npairs) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int quad = warp & 3;
    const int cg = (warp - 4) >> 2;
    const int sub_r = lane >> 2, c0 = (lane & 3) * 8;
    const uint8_t* epi_rd = epi_smem + sub_r * EPI_PITCH + c0 * 4;
    float4* epi_wr = reinterpret_cast<float4*>(epi_smem + lane * EPI_PITCH);
    const int r_warp0 = (int)rank * BM + quad * 32;
    const uint32_t tempty_leader0 = mapa_u32(smem_u32(&tempty[0]), 0), tempty_leader1 = mapa_u32(smem_u32(&tempty[1]), 0);
    int it = 0;
    for (int tile = pair; tile < p.num_tiles; tile += npairs, ++it) {
        const int as = it & 1;
        const int grp = tile % p.tiles_n;
        const int mt = tile / p.tiles_n;
        const int t_base = mt * (2 * BM) + r_warp0;            // first flat row (4-frame group) of this warp's 32
        // destination of the four staged rows this lane stores: row -> (utterance, first frame)
        int dst_b[4], dst_t[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int f0 = 4 * (t_base + sub_r + 8 * j);
            dst_b[j] = f0 / p.toep_Tp;
            dst_t[j] = f0 - dst_b[j] * p.toep_Tp;
        }
        mbar_wait_cl(&tfull[as], (uint32_t)(it >> 1) & 1u, 4);
        tc_fence_after();
        const uint32_t tcol = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(as * BN);
        for (int c = cg; c < BN / 32; c += NUM_EPI_WARPS / 4) {
            uint32_t r[32];
            tmem_ld32(tcol + (uint32_t)(c * 32), r);
            tmem_ld_wait();
            const int jj = c >> 1;                               // frame within the 4-frame group
            const int ch0 = grp * 64 + (c & 1) * 32;             // first of this chunk's 32 channels
            const float4* bp = reinterpret_cast<const float4*>(p.bias + ch0);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float4 bb = __ldg(bp + j);
                float4 v = make_float4(__uint_as_float(r[4 * j]) + bb.x, __uint_as_float(r[4 * j + 1]) + bb.y,
                                       __uint_as_float(r[4 * j + 2]) + bb.z, __uint_as_float(r[4 * j + 3]) + bb.w);
                gelu_erf_fast2(v.x, v.y);
                gelu_erf_fast2(v.z, v.w);
                epi_wr[j] = v;
            }
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float4 x0 = *reinterpret_cast<const float4*>(epi_rd + j * (8 * EPI_PITCH));
                const float4 x1 = *reinterpret_cast<const float4*>(epi_rd + j * (8 * EPI_PITCH) + 16);
                const int t = dst_t[j] + jj;
                if (dst_b[j] < p.toep_B && t < p.toep_T)
                    *reinterpret_cast<uint4*>(p.out16 + ((long long)dst_b[j] * p.toep_T + t) * p.ldo + ch0 + c0) =
                        make_uint4(pack_bf16(x0.x, x0.y), pack_bf16(x0.z, x0.w), pack_bf16(x1.x, x1.y), pack_bf16(x1.z, x1.w));
            }
            __syncwarp();
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(as ? tempty_leader1 : tempty_leader0);
    }
}

// MODE (own kernel instantiations, so that one variant's register needs do not touch the code generated for the others):
//   0 default epilogues, 1 residual prefetch (fp32 out + residual), 2 = 1 + LayerNorm-fold producer, 3 fused LayerNorm(512) + GELU
//   (conv encoder), 4 LayerNorm-fold consumer, 5 residual epilogue on the TMA engine (+ fold producer; replaces 1 / 2),
//   6 bf16-output epilogue on the TMA engine (+ fold consumer; replaces 4 and the bf16 cases of 0), 7 Toeplitz pos-conv scatter.
// RP: the residual-prefetch epilogue (fp32 out + residual) as its own instantiation, so that its register needs do not
// touch the code generated for the other variants (as one kernel every variant slowed down by 1.6 %).
template <int BN, int ST = 0, int MODE = 0>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmA2,
               const __grid_constant__ EpiMaps em, const GemmParams p) {
    using C = Cfg<BN, ST>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + (((base + 1023u) & ~1023u) - base);   // SWIZZLE_128B tiles need 1024 B alignment
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + C::STAGES * C::A_BYTES;
    uint8_t* smem_w1 = smem + ((C::EPI_OFF + 1023) & ~1023);      // cat mode: resident pass-1 W (<= 6 x 8 KB) over the unused epilogue staging
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + C::STAGES * C::STAGE_BYTES);   // used in the leader CTA
    uint64_t* empty = full + C::STAGES;                                                // used in both CTAs
    uint64_t* tfull = empty + C::STAGES;                                               // both CTAs
    uint64_t* tempty = tfull + 2;                                                      // leader CTA
    uint64_t* hfull = tempty + 2;                                                      // halo tiles (BN = 64): leader CTA
    uint64_t* hempty = hfull + 2;                                                      // both CTAs
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(hempty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();          // 0 = leader (issues the MMAs), 1 = peer
    const int pair = blockIdx.x >> 1;
    const int npairs = gridDim.x >> 1;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        if constexpr (MODE == 5) { tma_prefetch_desc(&em.x32); tma_prefetch_desc(&em.x16); }
        if constexpr (MODE == 6) tma_prefetch_desc(&em.x16);
    }
    if constexpr (MODE == 5) {
        if (warp == 3 && lane == 0) {
            uint64_t* lb = reinterpret_cast<uint64_t*>(smem + C::EPI5_BAR_OFF);
            for (int i = 0; i < 2 * NUM_EPI_WARPS; ++i) mbar_init(&lb[i], 1);
            fence_barrier_init();
        }
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < C::STAGES; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tfull[i], 1);
            mbar_init(&tempty[i], 2 * NUM_EPI_WARPS);
            mbar_init(&hfull[i], 1);
            mbar_init(&hempty[i], 1);
        }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc_2cta(tmem_slot, C::TMEM_COLS);
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_trigger();       // persistent single-wave grid: the next kernel may take an SM as soon as this kernel's CTA leaves it
    pdl_wait();          // everything above touched only shared / tensor memory and the kernel parameters

    // Roles are warp-uniform: every lane of the producer / MMA warp runs the loop (so tile, stage and descriptor values live
    // in uniform registers) and one elected lane issues the TMA / tcgen05 instructions.  With the loop under `lane == 0`
    // ptxas wrapped every UTMALDG / UTCHMMA in an ELECT + R2UR waterfall (~200 cycles per MMA regardless of N: profiles/r1b).
    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer (both CTAs)
        int stage = 0;
        uint32_t phase = 0;
        const bool skip_sync = (p.probe & 3) == 3;
        int hit = 0;
        for (int n_tile, mt; tile_at(p, hit, pair, npairs, n_tile, mt); ++hit) {
            const int b = mt / p.tiles_m_per_batch;
            const int t0 = (mt - b * p.tiles_m_per_batch) * (2 * BM) + (int)rank * BM;
            const int n0 = n_tile * BN + (int)rank * (BN / 2);
            const int a_col0 = n_tile * p.a_col_group_stride;
            if (BN == 64 && p.halo) {
                // ---- halo mode: this tile's A rows once (2 boxes of 128 rows per plane), then only W k-blocks stream
                const int hs = hit & 1;
                const int nplanes = p.nseg == 3 ? 2 : 1;
                if (p.cat && hit == 0) {
                    // W-resident (cat) mode: every W k-block of both passes is staged ONCE per CTA (the N = 64 tile is the whole N):
                    // pass 1 = this CTA's plane, 64 rows per k-block (8 KB, in the unused epilogue staging region), pass 2 = its 32
                    // rows of W_hi (4 KB, in the W ring region).  full[0] of the leader counts both CTAs' bytes.
                    if (elect_one()) {
                        if (rank == 0) mbar_arrive_expect_tx(&full[0], (uint32_t)(2 * p.nkb * 3 * C::B_BYTES));
                        const uint32_t fb = mapa_u32(smem_u32(&full[0]), 0);
                        for (int kb = 0; kb < p.nkb; ++kb) {
                            tma_load_3d_2cta(&tmB, fb, smem_w1 + kb * (2 * C::B_BYTES), kb * BK, n_tile * BN, (int)rank);
                            tma_load_3d_2cta(&tmB, fb, smem_w1 + kb * (2 * C::B_BYTES) + C::B_BYTES, kb * BK, n_tile * BN + BN / 2, (int)rank);
                            tma_load_3d_2cta(&tmB, fb, smem_b + kb * C::B_BYTES, kb * BK, n0, 0);
                        }
                    }
                    __syncwarp();
                }
                mbar_wait_cl(&hempty[hs], ((uint32_t)(hit >> 1) & 1u) ^ 1u, 6);
                if (elect_one()) {
                    // second box: only the halo_rows2 rows the taps can reach (tmA2; the rows behind it keep stale, never-read data)
                    if (rank == 0) mbar_arrive_expect_tx(&hfull[hs], (uint32_t)(2 * nplanes * (C::A_BYTES + p.halo_rows2 * 128)));
                    const uint32_t fb = mapa_u32(smem_u32(&hfull[hs]), 0);
                    for (int pl = 0; pl < nplanes; ++pl) {
                        tma_load_5d_2cta(&tmA, fb, smem_a + ((hs * 2 + pl) * 2) * C::A_BYTES, a_col0, 0, t0 + p.halo_min, b, pl);
                        tma_load_5d_2cta(&tmA2, fb, smem_a + ((hs * 2 + pl) * 2 + 1) * C::A_BYTES, a_col0, 0, t0 + p.halo_min + BM, b, pl);
                    }
                }
                __syncwarp();
                if (p.cat) continue;          // nothing else streams: the next tile's halo goes out as soon as its slot is free
                for (int seg = 0; seg < p.nseg; ++seg) {
                    const int b_plane = seg == 1 ? 1 : 0;
                    for (int kb = 0; kb < p.nkb; ++kb) {
                        mbar_wait_cl(&empty[stage], phase ^ 1u, 1);
                        if (elect_one()) {
                            if (rank == 0) mbar_arrive_expect_tx(&full[stage], 2 * C::B_BYTES);
                            const uint32_t fb = mapa_u32(smem_u32(&full[stage]), 0);
                            tma_load_3d_2cta(&tmB, fb, smem_b + stage * C::B_BYTES, kb * BK, n0, b_plane);
                        }
                        __syncwarp();
                        if (++stage == C::STAGES) { stage = 0; phase ^= 1u; }
                    }
                }
                continue;
            }
            for (int seg = 0; seg < p.nseg; ++seg) {
                const int a_plane = seg == 2 ? 1 : 0;
                const int b_plane = seg == 1 ? 1 : 0;
                for (int kb = 0; kb < p.nkb; ++kb) {
                    if (!skip_sync) mbar_wait_cl(&empty[stage], phase ^ 1u, 1);
                    const int tap = kb / p.kb_per_tap;
                    const int cib = kb - tap * p.kb_per_tap;
                    const int ti = tap / p.ntaps_w;
                    const int tj = tap - ti * p.ntaps_w;
                    const int lin = p.shift0 + ti * p.shift_h + tj * p.shift_w;
                    const int rs = floordiv(lin, p.a_sub);
                    const int sub = lin - rs * p.a_sub;
                    if (elect_one()) {
                        if (p.probe & 1) {
                            if (rank == 0 && !(p.probe & 2)) mbar_arrive(&full[stage]);
                        } else {
                            // the leader's barrier counts the bytes of both CTAs' boxes
                            if (rank == 0) mbar_arrive_expect_tx(&full[stage], 2 * C::STAGE_BYTES);
                            const uint32_t fb = mapa_u32(smem_u32(&full[stage]), 0);
                            tma_load_5d_2cta(&tmA, fb, smem_a + stage * C::A_BYTES, a_col0 + cib * BK, sub, t0 + rs, b, a_plane);
                            tma_load_3d_2cta(&tmB, fb, smem_b + stage * C::B_BYTES, kb * BK, n0, b_plane);
                        }
                    }
                    __syncwarp();
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1u; }
                }
            }
        }
        // producer tail: every stage this CTA filled has been consumed (the leader's commits target this CTA's
        // barriers, so it must not exit before they landed)
        for (int i = 0; i < C::STAGES && !skip_sync && !(BN == 64 && p.cat); ++i) {      // (cat mode does not use the ring)
            mbar_wait_cl(&empty[stage], phase ^ 1u, 5);
            if (++stage == C::STAGES) { stage = 0; phase ^= 1u; }
        }
        if (BN == 64 && p.halo) {
            for (int i = 0; i < 2 && i < hit; ++i) {
                const int j = hit - 1 - i;          // the last (two) halo fills of this CTA have been consumed
                mbar_wait_cl(&hempty[j & 1], (uint32_t)(j >> 1) & 1u, 7);
            }
        }
    } else if (warp == 1) {
        if (rank == 0) {
            // ------------------------------------------------------------------ MMA issuer (leader CTA only)
            const uint32_t idesc = make_idesc_bf16_m256(BN);
            const int nk_total = p.nseg * p.nkb;
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int n_tile_, mt_; tile_at(p, it, pair, npairs, n_tile_, mt_); ++it) {
                const int as = it & 1;
                const uint32_t aphase = (uint32_t)(it >> 1) & 1u;
                mbar_wait_cl(&tempty[as], aphase ^ 1u, 2);
                tc_fence_after();
                if ((p.probe & 3) == 3) {      // pure-epilogue probe: hand the (stale) accumulator over immediately
                    if (elect_one()) {
                        mbar_arrive(&tfull[as]);
                        mbar_arrive_cluster(mapa_u32(smem_u32(&tfull[as]), 1));
                    }
                    __syncwarp();
                    continue;
                }
                const uint32_t d_tmem = tmem_base + (uint32_t)(as * ((BN == 64 && p.cat) ? 2 * BN : BN));
                if (BN == 64 && p.halo && p.cat) {
                    const int hs = it & 1;
                    if (it == 0) { mbar_wait_cl(&full[0], 0u, 3); tc_fence_after(); }       // resident W (both CTAs) has landed
                    mbar_wait_cl(&hfull[hs], aphase, 8);
                    tc_fence_after();
                    const uint32_t idesc2 = make_idesc_bf16_m256(2 * BN);
                    for (int pass = 0; pass < 2; ++pass) {
                        const uint32_t a_base = smem_u32(smem_a + (hs * 2 + pass) * 2 * C::A_BYTES);      // pass 0: hi plane, pass 1: lo plane
                        for (int kb = 0; kb < p.nkb; ++kb) {
                            const int ti = kb / p.ntaps_w;
                            const int tj = kb - ti * p.ntaps_w;
                            const int off = p.shift0 + ti * p.shift_h + tj * p.shift_w - p.halo_min;
                            const uint64_t adesc = make_sw128_desc(a_base + (uint32_t)off * 128u);
                            const uint64_t bdesc = make_sw128_desc(pass == 0 ? smem_u32(smem_w1 + kb * (2 * C::B_BYTES)) : smem_u32(smem_b + kb * C::B_BYTES));
                            if (elect_one()) {
#pragma unroll
                                for (int kk = 0; kk < BK / 16; ++kk)
                                    if (kk < p.kk_steps && !(p.probe & 0x200))
                                        umma_bf16_2cta(d_tmem, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), pass == 0 ? idesc2 : idesc,
                                                       (pass > 0 || kb > 0 || kk > 0) ? 1u : 0u);
                                if (pass == 1 && kb == p.nkb - 1) { umma_commit_2cta(&tfull[as]); umma_commit_2cta(&hempty[hs]); }
                            }
                            __syncwarp();
                        }
                    }
                    continue;
                }
                if (BN == 64 && p.halo) {
                    const int hs = it & 1;
                    mbar_wait_cl(&hfull[hs], aphase, 8);
                    tc_fence_after();
                    int k = 0;
                    for (int seg = 0; seg < p.nseg; ++seg) {
                        const int a_plane = seg == 2 ? 1 : 0;
                        const uint32_t a_base = smem_u32(smem_a + (hs * 2 + a_plane) * 2 * C::A_BYTES);
                        for (int kb = 0; kb < p.nkb; ++kb, ++k) {
                            const int ti = kb / p.ntaps_w;
                            const int tj = kb - ti * p.ntaps_w;
                            const int off = p.shift0 + ti * p.shift_h + tj * p.shift_w - p.halo_min;     // 0 .. 128 rows into the halo
                            mbar_wait_cl(&full[stage], phase, 3);
                            tc_fence_after();
                            const uint32_t a_addr = a_base + (uint32_t)off * 128u;
                            // rows are 128 B: a row-shifted start address keeps the 8-row / 1024 B swizzle period, and the 128 B swizzle is a
                            // function of the absolute shared-memory address (as TMA wrote it), so base_offset stays 0 (measured on B200:
                            // bit-identical to the per-tap loads; base_offset = row phase gives wrong operands)
                            const uint64_t adesc = make_sw128_desc(a_addr);
                            const uint64_t bdesc = make_sw128_desc(smem_u32(smem_b + stage * C::B_BYTES));
                            if (elect_one()) {
#pragma unroll
                                for (int kk = 0; kk < BK / 16; ++kk)
                                    if (kk < p.kk_steps)
                                        umma_bf16_2cta(d_tmem, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc, (k > 0 || kk > 0) ? 1u : 0u);
                                umma_commit_2cta(&empty[stage]);
                                if (k == nk_total - 1) { umma_commit_2cta(&tfull[as]); umma_commit_2cta(&hempty[hs]); }
                            }
                            __syncwarp();
                            if (++stage == C::STAGES) { stage = 0; phase ^= 1u; }
                        }
                    }
                    continue;
                }
                for (int k = 0; k < nk_total; ++k) {
                    mbar_wait_cl(&full[stage], phase, 3);
                    tc_fence_after();
                    const uint64_t adesc = make_sw128_desc(smem_u32(smem_a + stage * C::A_BYTES));
                    const uint64_t bdesc = make_sw128_desc(smem_u32(smem_b + stage * C::B_BYTES));
                    if (elect_one()) {
                        if (p.probe & 2) {
                            mbar_arrive(&empty[stage]);
                            mbar_arrive_cluster(mapa_u32(smem_u32(&empty[stage]), 1));
                            if (k == nk_total - 1) { mbar_arrive(&tfull[as]); mbar_arrive_cluster(mapa_u32(smem_u32(&tfull[as]), 1)); }
                        } else {
#pragma unroll
                            for (int kk = 0; kk < BK / 16; ++kk) {
                                // advance 16 bf16 = 32 B inside the 128 B swizzle atom: +2 in the (addr >> 4) field
                                umma_bf16_2cta(d_tmem, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc, (k > 0 || kk > 0) ? 1u : 0u);
                            }
                            umma_commit_2cta(&empty[stage]);         // frees this smem stage in both CTAs
                            if (k == nk_total - 1) umma_commit_2cta(&tfull[as]);   // accumulator complete (both CTAs' epilogues)
                        }
                    }
                    __syncwarp();
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1u; }
                }
            }
        }
    } else if (warp >= 4) {
        // ---------------------------------------------------------------------- epilogue (both CTAs, own 128 rows)
        uint8_t* epi_smem = smem + C::EPI_OFF + (warp - 4) * (32 * EPI_PITCH);
        if constexpr (MODE == 7) epilogue_toeplitz<BN>(p, tfull, tempty, tmem_base, epi_smem, rank, pair, npairs);
        else if constexpr (MODE == 6) {
            uint8_t* wb = smem + C::EPI5_OFF + (warp - 4) * 4096;
            float* st = reinterpret_cast<float*>(smem + C::LN_OFF);
            if (p.flags & EPI_GELU) epilogue_bf16_tma<BN, true>(p, em, tfull, tempty, tmem_base, wb, st, rank, pair, npairs);
            else epilogue_bf16_tma<BN, false>(p, em, tfull, tempty, tmem_base, wb, st, rank, pair, npairs);
        } else if constexpr (MODE == 5)
            epilogue_resid_tma<BN>(p, em, tfull, tempty, tmem_base, smem + C::EPI5_OFF + (warp - 4) * EPI5_WARP_BYTES,
                                   reinterpret_cast<uint64_t*>(smem + C::EPI5_BAR_OFF) + (warp - 4) * 2, rank, pair, npairs);
        else if constexpr (MODE == 3) epilogue_ln_tiles(p, tfull, tempty, tmem_base, epi_smem, reinterpret_cast<float*>(smem + C::LN_OFF), rank, pair, npairs);
        else if constexpr (MODE == 1) epilogue_tiles<BN, EPI_OUT32 | EPI_RESID, 1>(p, tfull, tempty, tmem_base, epi_smem, rank, pair, npairs);
        else if constexpr (MODE == 2) epilogue_tiles<BN, EPI_OUT32 | EPI_RESID, 2>(p, tfull, tempty, tmem_base, epi_smem, rank, pair, npairs);
        else if constexpr (MODE == 4) {
            float* st = reinterpret_cast<float*>(smem + C::LN_OFF);
            if (p.flags & EPI_GELU) epilogue_tiles<BN, EPI_OUT16 | EPI_GELU, 4>(p, tfull, tempty, tmem_base, epi_smem, rank, pair, npairs, st);
            else epilogue_tiles<BN, EPI_OUT16, 4>(p, tfull, tempty, tmem_base, epi_smem, rank, pair, npairs, st);
        }
        else if (BN == 64 && p.conv_mode) epilogue_conv_tiles<BN>(p, tfull, tempty, tmem_base, rank, pair, npairs);
        else switch (p.flags) {      // uniform over the grid: each variant is a branch-free chunk loop
            case EPI_OUT16: epilogue_tiles<BN, EPI_OUT16>(p, tfull, tempty, tmem_base, epi_smem, rank, pair, npairs); break;
            case EPI_OUT16 | EPI_GELU: epilogue_tiles<BN, EPI_OUT16 | EPI_GELU>(p, tfull, tempty, tmem_base, epi_smem, rank, pair, npairs); break;
            case EPI_OUT32: epilogue_tiles<BN, EPI_OUT32>(p, tfull, tempty, tmem_base, epi_smem, rank, pair, npairs); break;
            case EPI_OUT32 | EPI_RESID: epilogue_tiles<BN, EPI_OUT32 | EPI_RESID>(p, tfull, tempty, tmem_base, epi_smem, rank, pair, npairs); break;
            default: epilogue_tiles<BN, EPI_ANY>(p, tfull, tempty, tmem_base, epi_smem, rank, pair, npairs); break;
        }
    }

    tc_fence_before();
    cluster_sync_all();      // no CTA of the pair exits (or frees TMEM) while the other may still signal it
    if (warp == 2) tmem_dealloc_2cta(tmem_base, C::TMEM_COLS);
}

// ------------------------------------------------------------------------------------------------
// Host side: tensor-map encoding (driver entry point fetched through the runtime, so the library does
// not link libcuda and still loads on a machine without a driver).
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encoder(std::string* err) {
    static PFN_encodeTiled fn = nullptr;
    static std::mutex mu;
    std::lock_guard<std::mutex> lk(mu);
    if (fn) return fn;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || p == nullptr) {
        if (err) *err = std::string("cuTensorMapEncodeTiled entry point unavailable: ") + cudaGetErrorString(e);
        return nullptr;
    }
    fn = reinterpret_cast<PFN_encodeTiled>(p);
    return fn;
}

static bool encode_map(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                       const uint32_t* box, std::string* err, CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16,
                       CUtensorMapSwizzle sw = CU_TENSOR_MAP_SWIZZLE_128B) {
    PFN_encodeTiled enc = get_encoder(err);
    if (!enc) return false;
    cuuint64_t gd[5];
    cuuint64_t gs[4];
    cuuint32_t bx[5], es[5];
    for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
    for (int i = 0; i < rank - 1; ++i) gs[i] = strides_bytes[i];
    CUresult r = enc(m, dt, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        if (err) {
            char buf[512];
            snprintf(buf, sizeof buf,
                     "cuTensorMapEncodeTiled failed (%d): rank %d dims [%llu %llu %llu %llu %llu] strides [%llu %llu %llu %llu] box [%u %u %u %u %u] base %p",
                     (int)r, rank, (unsigned long long)gd[0], (unsigned long long)(rank > 1 ? gd[1] : 0),
                     (unsigned long long)(rank > 2 ? gd[2] : 0), (unsigned long long)(rank > 3 ? gd[3] : 0),
                     (unsigned long long)(rank > 4 ? gd[4] : 0), (unsigned long long)gs[0],
                     (unsigned long long)(rank > 2 ? gs[1] : 0), (unsigned long long)(rank > 3 ? gs[2] : 0),
                     (unsigned long long)(rank > 4 ? gs[3] : 0), bx[0], bx[1], rank > 2 ? bx[2] : 0, rank > 3 ? bx[3] : 0,
                     rank > 4 ? bx[4] : 0, base);
            *err = buf;
        }
        return false;
    }
    return true;
}

bool encode_tensor_map_bf16(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                            const uint32_t* box, std::string* err) {
    return encode_map(m, base, rank, dims, strides_bytes, box, err);
}

cudaError_t ensure_dyn_smem(const void* func, size_t bytes) {
    static std::mutex mu;
    static std::map<std::pair<int, const void*>, size_t> cur;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lk(mu);
    size_t& c = cur[std::make_pair(dev, func)];
    if (bytes > c) {
        e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        if (e != cudaSuccess) return e;
        c = bytes;
    }
    return cudaSuccess;
}

static EpiMaps g_no_maps;       // zero-initialised placeholder for the variants that do not use the TMA epilogue
template <int BN, int ST = 0, int MODE = 0>
static cudaError_t launch_t(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& ta2, const GemmParams& p, int grid, cudaStream_t s,
                            const EpiMaps& em = g_no_maps) {
    constexpr int SMEM = MODE == 5 ? Cfg<BN, ST>::SMEM_EPI5 : ((MODE == 3 || MODE == 4 || MODE == 6) ? Cfg<BN, ST>::SMEM_LN : Cfg<BN, ST>::SMEM);
    {
        cudaError_t e = ensure_dyn_smem(reinterpret_cast<const void*>(gemm_tc_kernel<BN, ST, MODE>), SMEM);
        if (e != cudaSuccess) return e;
    }
    // __cluster_dims__(2,1,1): grid is even
    return launch_pdl(gemm_tc_kernel<BN, ST, MODE>, dim3((unsigned)grid), dim3(NUM_THREADS), (size_t)SMEM, s, ta, tb, ta2, em, p);
}

cudaError_t launch_gemm_tc(const GemmDesc& d_in, cudaStream_t s, int num_sms, std::string* err, int probe) {
    GemmDesc d = d_in;
    if (((probe >> 12) & 3) == 1) d.BN = 128;       // probes: N-tile override
    if (((probe >> 12) & 3) == 2) d.BN = 64;
    if (d.BN != 64 && d.BN != 128 && d.BN != 256) { if (err) *err = "BN must be 64/128/256"; return cudaErrorInvalidValue; }
    if (d.N % d.BN != 0 || d.nkb <= 0 || d.kb_per_tap <= 0 || d.nkb % d.kb_per_tap != 0) {
        if (err) *err = "gemm: N must be a multiple of BN and nkb of kb_per_tap";
        return cudaErrorInvalidValue;
    }
    const int nplanes = d.split ? 2 : 1;
    CUtensorMap ta, tb, ta2;
    bool a_rank2 = false;
    {
        // A: {cols, sub, rows, batch, plane}
        uint64_t dims[5] = {(uint64_t)d.a_cols, (uint64_t)d.a_sub, (uint64_t)d.a_rows, (uint64_t)d.nbatch, (uint64_t)nplanes};
        uint64_t st[4] = {(uint64_t)(d.a_sub > 1 ? d.a_sub_stride : d.a_row_stride) * 2, (uint64_t)d.a_row_stride * 2,
                          (uint64_t)(d.nbatch > 1 ? d.a_batch_stride : d.a_row_stride * (int64_t)d.a_rows) * 2,
                          (uint64_t)(nplanes > 1 ? d.a_plane_stride : d.a_row_stride * (int64_t)d.a_rows * d.nbatch) * 2};
        if (d.nbatch > 1 && d.a_batch_stride == 0) st[2] = (uint64_t)d.a_row_stride * 2;
        uint32_t box[5] = {(uint32_t)BK, 1, (uint32_t)BM, 1, 1};
        a_rank2 = (probe & 32) && d.a_sub == 1 && d.nbatch == 1 && nplanes == 1;
        if (a_rank2) {
            uint64_t d2[2] = {(uint64_t)d.a_cols, (uint64_t)d.a_rows};
            uint64_t s2[1] = {(uint64_t)d.a_row_stride * 2};
            uint32_t b2[2] = {(uint32_t)BK, (uint32_t)BM};
            if (!encode_map(&ta, d.A, 2, d2, s2, b2, err)) return cudaErrorInvalidValue;
        } else if (!encode_map(&ta, d.A, 5, dims, st, box, err)) return cudaErrorInvalidValue;
    }
    {
        // W: {ktot, N, plane}
        const uint64_t ktot = (uint64_t)d.nkb * BK;
        uint64_t dims[3] = {ktot, (uint64_t)d.N, (uint64_t)nplanes};
        uint64_t st[2] = {ktot * 2, (uint64_t)(nplanes > 1 ? d.w_plane_stride : (int64_t)ktot * d.N) * 2};
        uint32_t box[3] = {(uint32_t)BK, (uint32_t)(d.BN / 2), 1};      // each CTA of the pair loads half of the N tile
        if (!encode_map(&tb, d.W, 3, dims, st, box, err)) return cudaErrorInvalidValue;
    }
    GemmParams p;
    p.tiles_m_per_batch = (d.rows_per_batch + 2 * BM - 1) / (2 * BM);   // 256-row tiles, one per CTA pair
    p.tiles_n = d.N / d.BN;
    p.num_tiles = p.tiles_m_per_batch * d.nbatch * p.tiles_n;
    p.nkb = d.nkb; p.kb_per_tap = d.kb_per_tap; p.ntaps_w = d.ntaps_w;
    p.shift0 = d.shift0; p.shift_h = d.shift_h; p.shift_w = d.shift_w;
    p.a_sub = d.a_sub; p.a_col_group_stride = d.a_col_group_stride; p.nseg = d.split ? 3 : 1;
    p.rows_per_batch = d.rows_per_batch; p.N = d.N;
    p.bias = d.bias; p.resid = d.resid; p.out32 = d.out32; p.out32_copy = d.out32 ? d.out32_copy : nullptr; p.out16 = d.out16;
    p.ldo = d.out32 ? d.ld32 : d.ld16;
    if ((d.resid && d.ld_resid != p.ldo) || (d.out32 && d.ld32 != p.ldo) || (d.out16 && d.ld16 != p.ldo) || p.ldo % 4 != 0) {
        if (err) *err = "gemm: resid / out32 / out16 must share one leading dimension (multiple of 4)";
        return cudaErrorInvalidValue;
    }
    static const int conv_probe = [] { const char* e = getenv("OCC_CONV_PROBE"); return e ? atoi(e) : 0; }();    // development: 1 no conv epilogue, 2 no cat MMAs
    p.out16_plane = d.out16_plane_stride; p.gelu = d.gelu; p.row_valid = d.row_valid; p.probe = (probe & 255) | ((conv_probe & 3) << 8); p.a_rank2 = a_rank2 ? 1 : 0;
    p.flags = (d.gelu ? EPI_GELU : 0) | (d.resid ? EPI_RESID : 0) | (d.out32 ? EPI_OUT32 : 0) | (d.out16 ? EPI_OUT16 : 0) |
              (d.out16_plane_stride ? EPI_PLANE : 0) | (d.row_valid ? EPI_ZROW : 0);
    p.conv_mode = d.conv_mode; p.conv_wp = d.conv_wp; p.conv_h0 = d.conv_h0; p.conv_h1 = d.conv_h1; p.conv_wm = d.conv_wm;
    p.cs1 = d.cs1; p.ct1 = d.ct1; p.cs2 = d.cs2; p.ct2 = d.ct2; p.cresid = d.cresid; p.cresid_plane = d.cresid_plane;
    p.conv_wlen = d.conv_wlen;
    p.halo = 0; p.halo_min = 0; p.kk_steps = BK / 16; p.cat = 0; p.halo_rows2 = BM;
    p.ln_mode = 0; p.num_m_tiles = p.tiles_m_per_batch * d.nbatch; p.ln_g = d.ln_gamma; p.ln_b = d.ln_beta;
    p.toep_T = d.toep_T; p.toep_Tp = d.toep_Tp; p.toep_B = d.toep_B;
    p.x16_out = d.ln_x16_out; p.stat_out = d.ln_stat_out; p.stat_in = d.ln_stat_in; p.colsum = d.ln_colsum; p.N_stat = d.nkb * BK;
    if (d.ln_stat_out != nullptr || d.ln_x16_out != nullptr) {
        if (d.BN != 256 || d.N != 256 * (LN_SLOTS / 3) || !d.ln_stat_out || !d.ln_x16_out || !d.out32 || !d.resid || d.out16 || d.split || d.gelu || d.row_valid ||
            d.ld32 != d.N || (probe & 255)) {
            if (err) *err = "gemm: the LayerNorm-fold producer epilogue needs N = 1024, BN = 256, fp32 output + residual, no split";
            return cudaErrorInvalidValue;
        }
    }
    if (d.ln_stat_in != nullptr) {
        if (d.BN != 256 || !d.ln_colsum || !d.out16 || d.out32 || d.resid || d.split || d.out16_plane_stride != 0 || d.row_valid || d.nbatch != 1 || (probe & 255)) {
            if (err) *err = "gemm: the LayerNorm-fold consumer epilogue needs BN = 256, a single bf16 output, no split / residual";
            return cudaErrorInvalidValue;
        }
    }
    if (d.ln_gamma != nullptr) {
        if (d.BN != 256 || d.N != 512 || d.split || d.out32 || !d.out16 || d.out16_plane_stride != 0 || d.resid || d.conv_mode || !d.ln_beta || (probe & 255)) {
            if (err) *err = "gemm: the fused LayerNorm epilogue needs N = 512, BN = 256, a single bf16 output and no split / residual";
            return cudaErrorInvalidValue;
        }
        p.ln_mode = 1;
    }
    ta2 = ta;
    {
        // halo mode (BN = 64, one k-block per tap, unit sub-stride): all taps of a tile lie within 128 rows of each other
        static const bool no_halo = [] { const char* e = getenv("OCC_GEMM_NO_HALO"); return e && e[0] == '1'; }();
        const int ntaps = d.nkb;
        if (d.BN == 64 && d.kb_per_tap == 1 && d.a_sub == 1 && ntaps > 1 && d.ntaps_w > 0 && !no_halo && !(probe & 255)) {
            int lo = 1 << 30, hi = -(1 << 30);
            for (int t = 0; t < ntaps; ++t) {
                const int lin = d.shift0 + (t / d.ntaps_w) * d.shift_h + (t % d.ntaps_w) * d.shift_w;
                lo = lin < lo ? lin : lo; hi = lin > hi ? lin : hi;
            }
            if (hi - lo <= BM) {
                p.halo = 1; p.halo_min = lo;
                static const bool no_trim = [] { const char* e = getenv("OCC_GEMM_NO_TRIM"); return e && e[0] == '1'; }();
                const int rows2 = ((hi - lo + 7) / 8) * 8;        // whole 8-row swizzle atoms
                if (rows2 < BM && rows2 > 0 && !no_trim) {
                    uint64_t dims[5] = {(uint64_t)d.a_cols, (uint64_t)d.a_sub, (uint64_t)d.a_rows, (uint64_t)d.nbatch, (uint64_t)nplanes};
                    uint64_t st[4] = {(uint64_t)d.a_row_stride * 2, (uint64_t)d.a_row_stride * 2,
                                      (uint64_t)(d.nbatch > 1 ? d.a_batch_stride : d.a_row_stride * (int64_t)d.a_rows) * 2,
                                      (uint64_t)(nplanes > 1 ? d.a_plane_stride : d.a_row_stride * (int64_t)d.a_rows * d.nbatch) * 2};
                    if (d.nbatch > 1 && d.a_batch_stride == 0) st[2] = (uint64_t)d.a_row_stride * 2;
                    uint32_t box[5] = {(uint32_t)BK, 1, (uint32_t)rows2, 1, 1};
                    if (!encode_map(&ta2, d.A, 5, dims, st, box, err)) return cudaErrorInvalidValue;
                    p.halo_rows2 = rows2;
                }
                if (d.k_real > 0 && d.k_real <= 32) p.kk_steps = (d.k_real + 15) / 16;    // channels >= k_real are zero in A and W
                static const bool no_cat = [] { const char* e = getenv("OCC_GEMM_NO_CAT"); return e && e[0] == '1'; }();
                if (d.conv_mode && d.split && d.nkb <= 6 && d.N == d.BN && !no_cat) p.cat = 1;      // 6 taps of resident W fit beside the halo
            }
        }
    }
    if (d.conv_mode && (d.BN != 64 || d.conv_wp <= 0 || (d.out16 && d.out16_plane_stride == 0))) {
        if (err) *err = "gemm: conv mode needs BN = 64, conv_wp > 0 and hi/lo output planes";
        return cudaErrorInvalidValue;
    }
    if (d.bias == nullptr) { if (err) *err = "gemm: bias is required"; return cudaErrorInvalidValue; }
    if (d.out16 && (p.ldo % 8 != 0)) { if (err) *err = "gemm: bf16 output needs a leading dimension multiple of 8"; return cudaErrorInvalidValue; }
    if (p.num_tiles <= 0) return cudaSuccess;
    const int max_pairs = num_sms / 2;
    if (d.toep_T > 0) {
        if (d.BN != 256 || d.N % 256 != 0 || d.a_sub != 4 || d.nbatch != 1 || !d.out16 || d.out32 || d.split || !d.gelu || d.toep_Tp % 4 != 0 || d.toep_Tp < d.toep_T ||
            d.toep_B < 1 || (probe & 255)) {
            if (err) *err = "gemm: the Toeplitz pos-conv form needs BN = 256, a_sub = 4, one flat batch, bf16 output with GELU";
            return cudaErrorInvalidValue;
        }
        const int g7 = 2 * (p.num_tiles < max_pairs ? p.num_tiles : max_pairs);
        return launch_t<256, 0, 7>(ta, tb, ta2, p, g7, s);
    }
    if (p.ln_mode) {
        const int g2 = 2 * (p.num_m_tiles < max_pairs ? p.num_m_tiles : max_pairs);
        return launch_t<256, 0, 3>(ta, tb, ta2, p, g2, s);
    }
    const int grid = 2 * (p.num_tiles < max_pairs ? p.num_tiles : max_pairs);
    switch (d.BN + (d.BN == 256 ? ((probe >> 8) & 15) : 0)) {
        case 64: return launch_t<64>(ta, tb, ta2, p, grid, s);
        case 256 + 3: return launch_t<256, 3>(ta, tb, ta2, p, grid, s);      // probes only (probe >> 8 = stage override)
        case 256 + 4: return launch_t<256, 4>(ta, tb, ta2, p, grid, s);
        case 256 + 6: return launch_t<256, 6>(ta, tb, ta2, p, grid, s);
        case 128: return launch_t<128>(ta, tb, ta2, p, grid, s);
        default:
            if (p.flags == (EPI_OUT32 | EPI_RESID) && !p.probe) {
                static const bool epi_tma = [] { const char* e = getenv("OCC_EPI_TMA"); return !(e && e[0] == '0'); }();     // 0: previous form, for A/B runs
                if (epi_tma && d.nbatch == 1 && d.resid == d.out32 && d.ld32 == d.N && d.N % 16 == 0) {
                    // in-place residual epilogue on the TMA engine: maps over the [rows, N] fp32 stream / copy and the bf16 operand
                    EpiMaps em;
                    const uint64_t dims[2] = {(uint64_t)d.N, (uint64_t)d.rows_per_batch};
                    const uint64_t st32[1] = {(uint64_t)d.N * 4}, st16[1] = {(uint64_t)d.N * 2};
                    const uint32_t box[2] = {16, 32};
                    if (!encode_map(&em.x32, d.out32, 2, dims, st32, box, err, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, CU_TENSOR_MAP_SWIZZLE_64B)) return cudaErrorInvalidValue;
                    em.copy32 = em.x32; em.x16 = em.x32;
                    if (d.out32_copy && !encode_map(&em.copy32, d.out32_copy, 2, dims, st32, box, err, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, CU_TENSOR_MAP_SWIZZLE_64B)) return cudaErrorInvalidValue;
                    if (d.ln_x16_out && !encode_map(&em.x16, d.ln_x16_out, 2, dims, st16, box, err, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, CU_TENSOR_MAP_SWIZZLE_32B)) return cudaErrorInvalidValue;
                    return launch_t<256, 0, 5>(ta, tb, ta2, p, grid, s, em);
                }
                return p.stat_out ? launch_t<256, 0, 2>(ta, tb, ta2, p, grid, s) : launch_t<256, 0, 1>(ta, tb, ta2, p, grid, s);
            }
            {
                static const bool epi_tma16 = [] { const char* e = getenv("OCC_EPI_TMA16"); return !(e && e[0] == '0'); }();     // 0: previous form, for A/B runs
                const bool bf16_only = (p.flags == EPI_OUT16 || p.flags == (EPI_OUT16 | EPI_GELU)) && !p.probe;
                if (epi_tma16 && bf16_only && d.nbatch == 1 && d.ld16 == d.N && d.N % 32 == 0) {
                    EpiMaps em;
                    const uint64_t dims[2] = {(uint64_t)d.N, (uint64_t)d.rows_per_batch};
                    const uint64_t st16[1] = {(uint64_t)d.N * 2};
                    const uint32_t box[2] = {32, 32};
                    if (!encode_map(&em.x16, d.out16, 2, dims, st16, box, err, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, CU_TENSOR_MAP_SWIZZLE_64B)) return cudaErrorInvalidValue;
                    em.x32 = em.x16; em.copy32 = em.x16;
                    return launch_t<256, 0, 6>(ta, tb, ta2, p, grid, s, em);
                }
            }
            if (p.stat_in) return launch_t<256, 0, 4>(ta, tb, ta2, p, grid, s);
            return launch_t<256>(ta, tb, ta2, p, grid, s);
    }
}

// ------------------------------------------------------------------------------------------------
// SIMT check kernel: same GemmDesc semantics, one thread per output element, fp32 accumulate.
// Used only by occ_debug_gemm / OCC_GEMM_SIMT=1 to cross-check the tensor-core path on the GPU.
__global__ void gemm_simt_kernel(const GemmDesc d, long long total) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const int n = (int)(idx % d.N);
    const long long orow = idx / d.N;
    const int b = (int)(orow / d.rows_per_batch);
    const int t = (int)(orow - (long long)b * d.rows_per_batch);
    const int a_col0 = (n / d.BN) * d.a_col_group_stride;
    const long long ktot = (long long)d.nkb * BK;
    float acc = 0.f;
    const int nseg = d.split ? 3 : 1;
    for (int seg = 0; seg < nseg; ++seg) {
        const bf16* Ap = d.A + (seg == 2 ? d.a_plane_stride : 0);
        const bf16* Wp = d.W + (seg == 1 ? d.w_plane_stride : 0) + (long long)n * ktot;
        for (int kb = 0; kb < d.nkb; ++kb) {
            const int tap = kb / d.kb_per_tap;
            const int cib = kb - tap * d.kb_per_tap;
            const int ti = tap / d.ntaps_w, tj = tap - ti * d.ntaps_w;
            const int lin = d.shift0 + ti * d.shift_h + tj * d.shift_w;
            int rs = lin / d.a_sub;
            if (lin % d.a_sub != 0 && lin < 0) rs -= 1;
            const int sub = lin - rs * d.a_sub;
            const int row = t + rs;
            if (row < 0 || row >= d.a_rows) continue;
            const bf16* ar = Ap + (long long)b * d.a_batch_stride + (long long)row * d.a_row_stride + (long long)sub * d.a_sub_stride + a_col0 + cib * BK;
            const bf16* wr = Wp + (long long)kb * BK;
            for (int c = 0; c < BK; ++c) {
                if (a_col0 + cib * BK + c < d.a_cols) acc += __bfloat162float(ar[c]) * __bfloat162float(wr[c]);
            }
        }
    }
    float v = acc;
    if (d.bias) v += d.bias[n];
    if (d.gelu) v = gelu_erf(v);
    if (d.resid) v += d.resid[orow * d.ld_resid + n];
    if (d.row_valid && t >= d.row_valid[b]) v = 0.f;
    if (d.out32) d.out32[orow * d.ld32 + n] = v;
    if (d.out32 && d.out32_copy) d.out32_copy[orow * d.ld32 + n] = v;
    if (d.out16) {
        const bf16 hi = __float2bfloat16_rn(v);
        d.out16[orow * d.ld16 + n] = hi;
        if (d.out16_plane_stride != 0) d.out16[d.out16_plane_stride + orow * d.ld16 + n] = __float2bfloat16_rn(v - __bfloat162float(hi));
    }
}

cudaError_t launch_gemm_simt(const GemmDesc& d, cudaStream_t s) {
    const long long total = (long long)d.nbatch * d.rows_per_batch * d.N;
    if (total <= 0) return cudaSuccess;
    const int threads = 256;
    const long long blocks = (total + threads - 1) / threads;
    gemm_simt_kernel<<<(unsigned)blocks, threads, 0, s>>>(d, total);
    return cudaGetLastError();
}

}  // namespace occ
