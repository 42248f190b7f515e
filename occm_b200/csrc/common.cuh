// common.cuh -- sm_100a PTX wrappers (mbarrier, TMA, tcgen05/TMEM) and small device helpers.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

namespace occ {

// ----------------------------------------------------------------------------- math helpers
__device__ __forceinline__ float gelu_erf(float x) {
    // erf-GELU, the form torch.nn.functional.gelu(approximate='none') computes
    return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f));
}
// erf-GELU for the GEMM epilogues: x * Phi(x) with erfc(z) from Abramowitz-Stegun 7.1.26 (|error| <= 1.5e-7 in erf),
// MUFU.RCP + MUFU.EX2 and 9 FP32 ops instead of erff()'s ~30 (the FC1 epilogue was issue-bound, profiles/r1a).
__device__ __forceinline__ float gelu_erf_fast(float x) {
    const float z = fabsf(x) * 0.70710678118654752f;
    float t;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, z, 1.0f)));
    float p = fmaf(t, 1.061405429f, -1.453152027f);
    p = fmaf(p, t, 1.421413741f);
    p = fmaf(p, t, -0.284496736f);
    p = fmaf(p, t, 0.254829592f);
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(z * z * -1.4426950408889634f));
    const float q = 0.5f * p * t * e;              // = 0.5 * erfc(z) = Phi(-|x|)
    return x * (x > 0.f ? 1.0f - q : q);
}
// The packed fp32x2 pipe (FFMA2 / FMUL2 / FADD2, sm_100): two IEEE fp32 operations per issue slot.
__device__ __forceinline__ uint64_t f2_pack(float a, float b) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void f2_unpack(uint64_t v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ uint64_t f2_fma(uint64_t a, uint64_t b, uint64_t c) { uint64_t r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ uint64_t f2_mul(uint64_t a, uint64_t b) { uint64_t r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ uint64_t f2_add(uint64_t a, uint64_t b) { uint64_t r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ uint64_t f2_splat(float a) { return f2_pack(a, a); }
// erf-GELU on a packed pair with ONE special-function op per element (the previous form, below, needs a reciprocal and an exponential; at two
// per element the MUFU pipe alone would hold conv0 at 0.2 ms per 64-utterance step).  With u = |x| sqrt(log2(e)/2):
//   q = 0.5 erfc(|x| / sqrt 2) = 2^(-P(u)),  -P a degree-8 polynomial (weighted minimax fit on u in [0, 5.2], i.e. |x| <= 6.1, of
//   log2(0.5 erfc); negative leading coefficient and monotone beyond the interval, so q -> 0 for any larger input),
//   gelu(x) = x Phi(x) = 0.5 x + |x| (0.5 - q): no sign select.
// Error against the exact erf form: <= 1.5e-7 * max(|x|, 1) over the fp32 range (checked in numpy with fp32 Horner steps;
// the previous form sits at 4e-7).  15 issue slots per pair against 22.
__device__ __forceinline__ uint64_t gelu_erf_pair(uint64_t x) {
    float x0, x1;
    f2_unpack(x, x0, x1);
    const uint64_t u = f2_pack(fabsf(x0) * 0.84932180028801907f, fabsf(x1) * 0.84932180028801907f);
    uint64_t p = f2_fma(u, f2_splat(-7.644745892321225e-06f), f2_splat(9.486568160355091e-05f));
    p = f2_fma(p, u, f2_splat(-0.00038058895734138787f));
    p = f2_fma(p, u, f2_splat(-0.0005468587623909116f));
    p = f2_fma(p, u, f2_splat(0.013839146122336388f));
    p = f2_fma(p, u, f2_splat(-0.085844486951828f));
    p = f2_fma(p, u, f2_splat(-0.6365513205528259f));
    p = f2_fma(p, u, f2_splat(-1.3553282022476196f));
    p = f2_fma(p, u, f2_splat(-0.9999998807907104f));
    float a0, a1, q0, q1;
    f2_unpack(p, a0, a1);
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(q0) : "f"(a0));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(q1) : "f"(a1));
    const uint64_t h = f2_fma(f2_pack(q0, q1), f2_splat(-1.1774100225154747f), f2_splat(0.58870501125773733f));   // (0.5 - q) / sqrt(log2(e)/2)
    return f2_fma(u, h, f2_mul(x, f2_splat(0.5f)));
}

// N pairs in lock step (coefficient-major Horner: N independent dependency chains side by side, so one warp alone keeps the FMA
// pipe busy -- ptxas interleaves only two chains when handed the pairs one after the other).  Same arithmetic as gelu_erf_pair.
template <int NP>
__device__ __forceinline__ void gelu_erf_pairs(uint64_t (&x)[NP]) {
    uint64_t u[NP], p[NP];
#pragma unroll
    for (int i = 0; i < NP; ++i) {
        float x0, x1;
        f2_unpack(x[i], x0, x1);
        u[i] = f2_pack(fabsf(x0) * 0.84932180028801907f, fabsf(x1) * 0.84932180028801907f);
    }
#pragma unroll
    for (int i = 0; i < NP; ++i) p[i] = f2_fma(u[i], f2_splat(-7.644745892321225e-06f), f2_splat(9.486568160355091e-05f));
#pragma unroll
    for (int i = 0; i < NP; ++i) p[i] = f2_fma(p[i], u[i], f2_splat(-0.00038058895734138787f));
#pragma unroll
    for (int i = 0; i < NP; ++i) p[i] = f2_fma(p[i], u[i], f2_splat(-0.0005468587623909116f));
#pragma unroll
    for (int i = 0; i < NP; ++i) p[i] = f2_fma(p[i], u[i], f2_splat(0.013839146122336388f));
#pragma unroll
    for (int i = 0; i < NP; ++i) p[i] = f2_fma(p[i], u[i], f2_splat(-0.085844486951828f));
#pragma unroll
    for (int i = 0; i < NP; ++i) p[i] = f2_fma(p[i], u[i], f2_splat(-0.6365513205528259f));
#pragma unroll
    for (int i = 0; i < NP; ++i) p[i] = f2_fma(p[i], u[i], f2_splat(-1.3553282022476196f));
#pragma unroll
    for (int i = 0; i < NP; ++i) p[i] = f2_fma(p[i], u[i], f2_splat(-0.9999998807907104f));
#pragma unroll
    for (int i = 0; i < NP; ++i) {
        float a0, a1, q0, q1;
        f2_unpack(p[i], a0, a1);
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(q0) : "f"(a0));
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(q1) : "f"(a1));
        p[i] = f2_pack(q0, q1);
    }
#pragma unroll
    for (int i = 0; i < NP; ++i) p[i] = f2_fma(p[i], f2_splat(-1.1774100225154747f), f2_splat(0.58870501125773733f));
#pragma unroll
    for (int i = 0; i < NP; ++i) x[i] = f2_fma(u[i], p[i], f2_mul(x[i], f2_splat(0.5f)));
}

// gelu_erf_fast2: the pair form used by the GEMM epilogues, the LayerNorm kernels and conv0.  -DOCC_GELU_V1 builds the previous
// (reciprocal + exponential, Abramowitz-Stegun 7.1.26) form for A/B runs.
#ifdef OCC_GELU_V1
__device__ __forceinline__ void gelu_erf_fast2(float& x0, float& x1) {
    const uint64_t z = f2_pack(fabsf(x0) * 0.70710678118654752f, fabsf(x1) * 0.70710678118654752f);
    float d0, d1, t0, t1;
    f2_unpack(f2_fma(f2_splat(0.3275911f), z, f2_splat(1.0f)), d0, d1);
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t0) : "f"(d0));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t1) : "f"(d1));
    const uint64_t t = f2_pack(t0, t1);
    uint64_t p = f2_fma(t, f2_splat(1.061405429f), f2_splat(-1.453152027f));
    p = f2_fma(p, t, f2_splat(1.421413741f));
    p = f2_fma(p, t, f2_splat(-0.284496736f));
    p = f2_fma(p, t, f2_splat(0.254829592f));
    float a0, a1, e0, e1;
    f2_unpack(f2_mul(f2_mul(z, z), f2_splat(-1.4426950408889634f)), a0, a1);
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e0) : "f"(a0));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e1) : "f"(a1));
    const uint64_t q = f2_mul(f2_mul(f2_mul(f2_splat(0.5f), p), t), f2_pack(e0, e1));       // 0.5 erfc(z) = Phi(-|x|)
    float q0, q1, r0, r1;
    f2_unpack(q, q0, q1);
    f2_unpack(f2_fma(q, f2_splat(-1.0f), f2_splat(1.0f)), r0, r1);                            // 1 - q (exact same rounding as 1.0f - q)
    float y0, y1;
    f2_unpack(f2_mul(f2_pack(x0, x1), f2_pack(x0 > 0.f ? r0 : q0, x1 > 0.f ? r1 : q1)), y0, y1);
    x0 = y0; x1 = y1;
}
#else
__device__ __forceinline__ void gelu_erf_fast2(float& x0, float& x1) { f2_unpack(gelu_erf_pair(f2_pack(x0, x1)), x0, x1); }
#endif
__device__ __forceinline__ float selu_f(float x) {
    const float alpha = 1.6732632423543772848170429916717f;
    const float scale = 1.0507009873554804934193349852946f;
    return x > 0.f ? scale * x : scale * alpha * (expf(x) - 1.0f);   // expm1 precision matters little at fp32
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}
// hi = bf16(x); lo = bf16(x - hi): x ~= hi + lo to ~2^-17 relative
__device__ __forceinline__ void split_bf16(float x, __nv_bfloat16& hi, __nv_bfloat16& lo) {
    hi = __float2bfloat16_rn(x);
    lo = __float2bfloat16_rn(x - __bfloat162float(hi));
}

// ----------------------------------------------------------------------------- smem / mbarrier
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// Non-blocking probe for POLLING loops that watch several barriers.  mbarrier.try_wait may suspend the thread for a
// system-dependent time limit before it answers; measured on B200 (tools/attn_trace.py) that limit is long enough to park an
// MMA-issuer warp on one barrier for thousands of cycles while another barrier it should serve is already complete.
__device__ __forceinline__ bool mbar_test_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: a protocol bug (wrong tx count, missed arrive) traps instead of hanging the GPU box.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, int tag) {
    if (mbar_try_wait(bar, parity)) return;
    long long t0 = clock64();
    while (!mbar_try_wait(bar, parity)) {
        if (clock64() - t0 > 4000000000LL) {   // ~2 s
            printf("occ: mbarrier timeout tag=%d block=%d thread=%d parity=%u\n", tag, (int)blockIdx.x, (int)threadIdx.x, parity);
            __trap();
        }
    }
}

// ----------------------------------------------------------------------------- programmatic dependent launch (PDL)
// A kernel launched through launch_pdl may become resident while its predecessor on the stream still runs (as soon as the
// predecessor's CTAs have triggered or exited and the SM has room), so its prologue -- barrier init, TMEM allocation, descriptor
// prefetch, the launch latency itself -- hides under the predecessor's tail.  pdl_wait() blocks until every predecessor grid
// has completed and its memory is visible: NOTHING that reads or writes global memory may precede it.  Both instructions are
// no-ops in a kernel launched without the attribute.  OCC_PDL=0 launches everything the plain way (A/B).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
inline bool pdl_enabled() {
    static const bool on = [] { const char* e = getenv("OCC_PDL"); return !(e && e[0] == '0'); }();
    return on;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ----------------------------------------------------------------------------- TMA
__device__ __forceinline__ void tma_prefetch_desc(const void* desc) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(desc)) : "memory");
}
__device__ __forceinline__ void tma_load_3d(const void* desc, uint64_t* bar, void* dst, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(desc)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
__device__ __forceinline__ void tma_load_4d(const void* desc, uint64_t* bar, void* dst, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(desc)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}

__device__ __forceinline__ void tma_load_5d(const void* desc, uint64_t* bar, void* dst, int c0, int c1, int c2, int c3, int c4) {
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(desc)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
}

// plain (own-CTA) 2-D tile load, and the store direction: shared memory -> global through a tensor map (bulk async group)
__device__ __forceinline__ void tma_load_2d(const void* desc, uint64_t* bar, void* dst, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(desc)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tma_store_2d(const void* desc, const void* src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(desc)), "r"(smem_u32(src)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all bulk stores committed by this thread have finished READING shared memory (the buffers may be rewritten)
__device__ __forceinline__ void tma_store_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// ... have completed (global writes performed)
__device__ __forceinline__ void tma_store_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// ----------------------------------------------------------------------------- tcgen05 / TMEM
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// D[tmem] (+)= A[smem] * B[smem]^T, bf16 in / fp32 accumulate, issued by ONE thread.
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// mbarrier arrives once all previously issued tcgen05.mma of this thread have completed.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// 32 lanes x 32 columns of fp32: thread `lane` of the warp gets row (quadrant*32+lane), columns [col, col+32)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
// 32 lanes x 16 columns
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ----------------------------------------------------------------------------- CTA pair (cta_group::2)
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `p` (a shared::cta address of this CTA) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t p, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(p), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    // default (.release.cta) semantics as in CUTLASS' ClusterBarrier::arrive(cta_id): a .release.cluster arrive compiles
    // to MEMBAR/ERRBAR and a cluster-scope acquire wait to CCTL.IVALL (L1 invalidate) per poll (profiles/r1b), and the
    // data handed over here lives in TMEM / shared memory, ordered by tcgen05.fence + the mbarrier itself.
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_cl(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait on barriers that peer-CTA threads / the peer's TMA / multicast commits arrive on.
__device__ __forceinline__ void mbar_wait_cl(uint64_t* bar, uint32_t parity, int tag) {
    if (mbar_try_wait_cl(bar, parity)) return;
    long long t0 = clock64();
    while (!mbar_try_wait_cl(bar, parity)) {
        if (clock64() - t0 > 4000000000LL) {   // ~2 s
            printf("occ: mbarrier timeout tag=%d block=%d thread=%d parity=%u\n", tag, (int)blockIdx.x, (int)threadIdx.x, parity);
            __trap();
        }
    }
}
__device__ __forceinline__ void tmem_alloc_2cta(uint32_t* dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2cta(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem, 256 rows over the CTA pair] (+)= A * B^T; issued by ONE thread of the leader CTA.  Each CTA supplies its own
// 128 rows of A and its own half of the N rows of B at the same shared-memory offsets.
__device__ __forceinline__ void umma_bf16_2cta(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrives on the mbarrier at this shared-memory offset in BOTH CTAs of the pair once all prior MMAs completed
__device__ __forceinline__ void umma_commit_2cta(uint64_t* bar) {
    const uint16_t mask = 3;
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}
// TMA loads of a CTA pair: data lands in this CTA's shared memory, bytes are counted on the LEADER's mbarrier
__device__ __forceinline__ void tma_load_3d_2cta(const void* desc, uint32_t bar_cluster_addr, void* dst, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(desc)), "r"(bar_cluster_addr), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d_2cta(const void* desc, uint32_t bar_cluster_addr, void* dst, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(desc)), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tma_load_5d_2cta(const void* desc, uint32_t bar_cluster_addr, void* dst, int c0, int c1, int c2, int c3, int c4) {
    asm volatile(
        "cp.async.bulk.tensor.5d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(desc)), "r"(bar_cluster_addr), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
}

__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}

// K-major, 128-byte-swizzled operand tile: rows of 64 bf16 (128 B), 8-row groups 1024 B apart.
__device__ __forceinline__ uint64_t make_sw128_desc(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);        // start address  [0,14)
    d |= (uint64_t)0 << 16;                            // LBO (unused for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;                  // SBO = 1024 B   [32,46)
    d |= (uint64_t)1 << 46;                            // descriptor version (sm_100)
    d |= (uint64_t)2 << 61;                            // SWIZZLE_128B
    return d;
}
// kind::f16 instruction descriptor: bf16 x bf16 -> fp32, both operands K-major, M=128.
__host__ __device__ constexpr uint32_t make_idesc_bf16(int n) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}

// same for a CTA pair: M = 256 (128 rows in each CTA), N = n over the pair
__host__ __device__ constexpr uint32_t make_idesc_bf16_m256(int n) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
}

}  // namespace occ
