// gemm_tc.cu -- the tcgen05/TMEM/TMA "shift-GEMM" that carries every dense contraction of the
// XLS-R frontend (conv layers 1-6, feature projection, grouped positional conv, QKV / out-proj / FC1 / FC2,
// LL).  One persistent CTA per SM; warp-specialised:
//   warp 0   : TMA producer (one lane) -- 5-D A boxes {64 cols, 1 sub, 128 rows}, 3-D W boxes {64, BN}
//   warp 1   : tcgen05.mma issuer (one lane) -- 128 x BN x 16 bf16 MMAs, fp32 accumulators in TMEM
//   warp 2   : TMEM allocator
//   warps 4-11: epilogue -- tcgen05.ld 32x32b, bias / erf-GELU / residual, vectorised global stores
// Two TMEM accumulator stages (2 x BN columns) overlap the epilogue of tile i with the MMAs of tile i+1.
#include "common.cuh"
#include "occ_internal.h"
#include <cuda.h>
#include <mutex>

namespace occ {

static constexpr int BM = 128;
static constexpr int BK = 64;
static constexpr int NUM_EPI_WARPS = 8;
static constexpr int NUM_THREADS = 128 + NUM_EPI_WARPS * 32;

struct GemmParams {
    int tiles_m_per_batch, tiles_n, num_tiles;
    int nkb, kb_per_tap, ntaps_w, shift0, shift_h, shift_w, a_sub, a_col_group_stride, nseg;
    int rows_per_batch, N;
    const float* bias;
    const float* resid;
    long long ld_resid;
    float* out32;
    long long ld32;
    bf16* out16;
    long long ld16;
    long long out16_plane;
    int gelu;
    const int* row_valid;
};

template <int BN>
struct Cfg {
    static constexpr int STAGES = BN == 256 ? 4 : (BN == 128 ? 6 : 8);
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int TMEM_COLS = 2 * BN;
    static constexpr int BAR_BYTES = (2 * STAGES + 4) * 8 + 16;
    static constexpr int SMEM = STAGES * STAGE_BYTES + BAR_BYTES + 1024;  // + alignment slack
};

__device__ __forceinline__ int floordiv(int a, int b) { int q = a / b; return (a % b != 0 && ((a < 0) != (b < 0))) ? q - 1 : q; }

template <int BN>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmParams p) {
    using C = Cfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + (((base + 1023u) & ~1023u) - base);   // SWIZZLE_128B tiles need 1024 B alignment
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + C::STAGES * C::A_BYTES;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + C::STAGES * C::STAGE_BYTES);
    uint64_t* empty = full + C::STAGES;
    uint64_t* tfull = empty + C::STAGES;
    uint64_t* tempty = tfull + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < C::STAGES; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tfull[i], 1);
            mbar_init(&tempty[i], NUM_EPI_WARPS);
        }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            // ------------------------------------------------------------------ TMA producer
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
                const int n_tile = tile % p.tiles_n;
                const int mt = tile / p.tiles_n;
                const int b = mt / p.tiles_m_per_batch;
                const int t0 = (mt - b * p.tiles_m_per_batch) * BM;
                const int n0 = n_tile * BN;
                const int a_col0 = n_tile * p.a_col_group_stride;
                for (int seg = 0; seg < p.nseg; ++seg) {
                    const int a_plane = seg == 2 ? 1 : 0;
                    const int b_plane = seg == 1 ? 1 : 0;
                    for (int kb = 0; kb < p.nkb; ++kb) {
                        mbar_wait(&empty[stage], phase ^ 1u, 1);
                        const int tap = kb / p.kb_per_tap;
                        const int cib = kb - tap * p.kb_per_tap;
                        const int ti = tap / p.ntaps_w;
                        const int tj = tap - ti * p.ntaps_w;
                        const int lin = p.shift0 + ti * p.shift_h + tj * p.shift_w;
                        const int rs = floordiv(lin, p.a_sub);
                        const int sub = lin - rs * p.a_sub;
                        mbar_arrive_expect_tx(&full[stage], C::STAGE_BYTES);
                        tma_load_5d(&tmA, &full[stage], smem_a + stage * C::A_BYTES, a_col0 + cib * BK, sub, t0 + rs, b, a_plane);
                        tma_load_3d(&tmB, &full[stage], smem_b + stage * C::B_BYTES, kb * BK, n0, b_plane);
                        if (++stage == C::STAGES) { stage = 0; phase ^= 1u; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ------------------------------------------------------------------ MMA issuer
            const uint32_t idesc = make_idesc_bf16(BN);
            const int nk_total = p.nseg * p.nkb;
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
                const int as = it & 1;
                const uint32_t aphase = (uint32_t)(it >> 1) & 1u;
                mbar_wait(&tempty[as], aphase ^ 1u, 2);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(as * BN);
                for (int k = 0; k < nk_total; ++k) {
                    mbar_wait(&full[stage], phase, 3);
                    tc_fence_after();
                    const uint64_t adesc = make_sw128_desc(smem_u32(smem_a + stage * C::A_BYTES));
                    const uint64_t bdesc = make_sw128_desc(smem_u32(smem_b + stage * C::B_BYTES));
#pragma unroll
                    for (int kk = 0; kk < BK / 16; ++kk) {
                        // advance 16 bf16 = 32 B inside the 128 B swizzle atom: +2 in the (addr >> 4) field
                        umma_bf16(d_tmem, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc, (k > 0 || kk > 0) ? 1u : 0u);
                    }
                    umma_commit(&empty[stage]);          // frees this smem stage when the MMAs have read it
                    if (k == nk_total - 1) umma_commit(&tfull[as]);   // accumulator complete
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1u; }
                }
            }
        }
    } else if (warp >= 4) {
        // ---------------------------------------------------------------------- epilogue
        const int ew = warp - 4;
        const int quad = warp & 3;            // TMEM lane quadrant this warp may access
        const int cg = ew >> 2;               // column group (2 groups)
        const int r_in_tile = quad * 32 + lane;
        int it = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
            const int as = it & 1;
            const uint32_t aphase = (uint32_t)(it >> 1) & 1u;
            const int n_tile = tile % p.tiles_n;
            const int mt = tile / p.tiles_n;
            const int b = mt / p.tiles_m_per_batch;
            const int t = (mt - b * p.tiles_m_per_batch) * BM + r_in_tile;
            const int n0 = n_tile * BN;
            const bool valid = t < p.rows_per_batch;
            const long long orow = (long long)b * p.rows_per_batch + t;
            const bool zero_row = (p.row_valid != nullptr) && valid && (t >= p.row_valid[b]);
            mbar_wait(&tfull[as], aphase, 4);
            tc_fence_after();
            for (int c = cg; c < BN / 32; c += NUM_EPI_WARPS / 4) {
                uint32_t r[32];
                tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(as * BN + c * 32), r);
                tmem_ld_wait();
                if (valid) {
                    const int n = n0 + c * 32;
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
                    if (p.bias != nullptr) {
                        const float4* bp = reinterpret_cast<const float4*>(p.bias + n);
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 bb = __ldg(bp + j);
                            v[4 * j + 0] += bb.x; v[4 * j + 1] += bb.y; v[4 * j + 2] += bb.z; v[4 * j + 3] += bb.w;
                        }
                    }
                    if (p.gelu) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
                    }
                    if (p.resid != nullptr) {
                        const float4* rp = reinterpret_cast<const float4*>(p.resid + orow * p.ld_resid + n);
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 rr = rp[j];
                            v[4 * j + 0] += rr.x; v[4 * j + 1] += rr.y; v[4 * j + 2] += rr.z; v[4 * j + 3] += rr.w;
                        }
                    }
                    if (zero_row) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = 0.f;
                    }
                    if (p.out32 != nullptr) {
                        float4* op = reinterpret_cast<float4*>(p.out32 + orow * p.ld32 + n);
#pragma unroll
                        for (int j = 0; j < 8; ++j) op[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                    }
                    if (p.out16 != nullptr) {
                        uint4* op = reinterpret_cast<uint4*>(p.out16 + orow * p.ld16 + n);
                        uint32_t h[16];
#pragma unroll
                        for (int j = 0; j < 16; ++j) h[j] = pack_bf16(v[2 * j], v[2 * j + 1]);
#pragma unroll
                        for (int j = 0; j < 4; ++j) op[j] = make_uint4(h[4 * j], h[4 * j + 1], h[4 * j + 2], h[4 * j + 3]);
                        if (p.out16_plane != 0) {
                            uint4* lp = reinterpret_cast<uint4*>(p.out16 + p.out16_plane + orow * p.ld16 + n);
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                const __nv_bfloat162 hh = *reinterpret_cast<__nv_bfloat162*>(&h[j]);
                                h[j] = pack_bf16(v[2 * j] - __low2float(hh), v[2 * j + 1] - __high2float(hh));
                            }
#pragma unroll
                            for (int j = 0; j < 4; ++j) lp[j] = make_uint4(h[4 * j], h[4 * j + 1], h[4 * j + 2], h[4 * j + 3]);
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[as]);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, C::TMEM_COLS);
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
                       const uint32_t* box, std::string* err) {
    PFN_encodeTiled enc = get_encoder(err);
    if (!enc) return false;
    cuuint64_t gd[5];
    cuuint64_t gs[4];
    cuuint32_t bx[5], es[5];
    for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
    for (int i = 0; i < rank - 1; ++i) gs[i] = strides_bytes[i];
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
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

template <int BN>
static cudaError_t launch_t(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid, cudaStream_t s) {
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::SMEM);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    gemm_tc_kernel<BN><<<grid, NUM_THREADS, Cfg<BN>::SMEM, s>>>(ta, tb, p);
    return cudaGetLastError();
}

cudaError_t launch_gemm_tc(const GemmDesc& d, cudaStream_t s, int num_sms, std::string* err) {
    if (d.BN != 64 && d.BN != 128 && d.BN != 256) { if (err) *err = "BN must be 64/128/256"; return cudaErrorInvalidValue; }
    if (d.N % d.BN != 0 || d.nkb <= 0 || d.kb_per_tap <= 0 || d.nkb % d.kb_per_tap != 0) {
        if (err) *err = "gemm: N must be a multiple of BN and nkb of kb_per_tap";
        return cudaErrorInvalidValue;
    }
    const int nplanes = d.split ? 2 : 1;
    CUtensorMap ta, tb;
    {
        // A: {cols, sub, rows, batch, plane}
        uint64_t dims[5] = {(uint64_t)d.a_cols, (uint64_t)d.a_sub, (uint64_t)d.a_rows, (uint64_t)d.nbatch, (uint64_t)nplanes};
        uint64_t st[4] = {(uint64_t)(d.a_sub > 1 ? d.a_sub_stride : d.a_row_stride) * 2, (uint64_t)d.a_row_stride * 2,
                          (uint64_t)(d.nbatch > 1 ? d.a_batch_stride : d.a_row_stride * (int64_t)d.a_rows) * 2,
                          (uint64_t)(nplanes > 1 ? d.a_plane_stride : d.a_row_stride * (int64_t)d.a_rows * d.nbatch) * 2};
        if (d.nbatch > 1 && d.a_batch_stride == 0) st[2] = (uint64_t)d.a_row_stride * 2;
        uint32_t box[5] = {(uint32_t)BK, 1, (uint32_t)BM, 1, 1};
        if (!encode_map(&ta, d.A, 5, dims, st, box, err)) return cudaErrorInvalidValue;
    }
    {
        // W: {ktot, N, plane}
        const uint64_t ktot = (uint64_t)d.nkb * BK;
        uint64_t dims[3] = {ktot, (uint64_t)d.N, (uint64_t)nplanes};
        uint64_t st[2] = {ktot * 2, (uint64_t)(nplanes > 1 ? d.w_plane_stride : (int64_t)ktot * d.N) * 2};
        uint32_t box[3] = {(uint32_t)BK, (uint32_t)d.BN, 1};
        if (!encode_map(&tb, d.W, 3, dims, st, box, err)) return cudaErrorInvalidValue;
    }
    GemmParams p;
    p.tiles_m_per_batch = (d.rows_per_batch + BM - 1) / BM;
    p.tiles_n = d.N / d.BN;
    p.num_tiles = p.tiles_m_per_batch * d.nbatch * p.tiles_n;
    p.nkb = d.nkb; p.kb_per_tap = d.kb_per_tap; p.ntaps_w = d.ntaps_w;
    p.shift0 = d.shift0; p.shift_h = d.shift_h; p.shift_w = d.shift_w;
    p.a_sub = d.a_sub; p.a_col_group_stride = d.a_col_group_stride; p.nseg = d.split ? 3 : 1;
    p.rows_per_batch = d.rows_per_batch; p.N = d.N;
    p.bias = d.bias; p.resid = d.resid; p.ld_resid = d.ld_resid;
    p.out32 = d.out32; p.ld32 = d.ld32; p.out16 = d.out16; p.ld16 = d.ld16;
    p.out16_plane = d.out16_plane_stride; p.gelu = d.gelu; p.row_valid = d.row_valid;
    if (p.num_tiles <= 0) return cudaSuccess;
    const int grid = p.num_tiles < num_sms ? p.num_tiles : num_sms;
    switch (d.BN) {
        case 64: return launch_t<64>(ta, tb, p, grid, s);
        case 128: return launch_t<128>(ta, tb, p, grid, s);
        default: return launch_t<256>(ta, tb, p, grid, s);
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
