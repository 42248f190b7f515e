// eer.cu -- one-class scores and the GPU equal-error-rate.
//   score_l2 / score_cos / ref_embedding : /root/reference/oc_classifier.py:182-197, 261 (+ OC-Softmax cosine)
//   eer : /root/reference/evaluate_metrics.py:3-40, bit-identical to numpy:
//         stable ascending sort of concat(target, nontarget) (np.argsort kind='mergesort') -> cumulative target
//         count -> FRR/FAR in float64 -> first argmin |FRR-FAR| -> (mean(FRR,FAR), threshold).
// The sort is a hand-written stable LSD radix sort (8 passes x 8 bits over order-preserving uint64 keys);
// stability + the initial [targets..., nontargets...] order reproduces mergesort's tie behaviour exactly.
#include <mutex>
#include <string>

#include "../../include/occ_b200.h"
#include "common.cuh"
#include "kernels.h"

namespace occ {

// ------------------------------------------------------------------------------------------------ scores
__global__ void score_l2_kernel(const float* __restrict__ emb, const float* __restrict__ ref, int B, int D, float eps, float* __restrict__ out) {
    pdl_wait();
    const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (b >= B) return;
    float acc = 0.f;
    for (int d = lane; d < D; d += 32) {
        const float df = ref[d] - emb[(long long)b * D + d] + eps;   // eps added to the difference (torch pairwise_distance)
        acc = fmaf(df, df, acc);
    }
    acc = warp_sum(acc);
    if (lane == 0) out[b] = sqrtf(acc);
}
cudaError_t launch_score_l2(const float* emb, const float* ref, int B, int D, float eps, float* out, cudaStream_t s) {
    if (B <= 0) return cudaSuccess;
    (void)launch_pdl(score_l2_kernel, dim3((unsigned)((B + 3) / 4)), dim3(128), (size_t)0, s, emb, ref, B, D, eps, out);
    return cudaGetLastError();
}
__global__ void score_cos_kernel(const float* __restrict__ emb, const float* __restrict__ w, int B, int D, float* __restrict__ out) {
    const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (b >= B) return;
    float xx = 0.f, ww = 0.f, xw = 0.f;
    for (int d = lane; d < D; d += 32) {
        const float x = emb[(long long)b * D + d], c = w[d];
        xx = fmaf(x, x, xx); ww = fmaf(c, c, ww); xw = fmaf(x, c, xw);
    }
    xx = warp_sum(xx); ww = warp_sum(ww); xw = warp_sum(xw);
    // F.normalize: x / max(||x||, 1e-12)
    if (lane == 0) out[b] = xw / (fmaxf(sqrtf(xx), 1e-12f) * fmaxf(sqrtf(ww), 1e-12f));
}
cudaError_t launch_score_cos(const float* emb, const float* w, int B, int D, float* out, cudaStream_t s) {
    if (B <= 0) return cudaSuccess;
    score_cos_kernel<<<(B + 3) / 4, 128, 0, s>>>(emb, w, B, D, out);
    return cudaGetLastError();
}
// ref[d] = mean_i emb[i][d]; thr = max_i || ref - emb[i] + eps ||   (single CTA; n is the bona-fide train-set size)
__global__ void __launch_bounds__(256)
ref_embedding_kernel(const float* __restrict__ emb, int n, int D, float eps, float* __restrict__ ref, float* __restrict__ thr) {
    extern __shared__ float sref[];
    __shared__ float smax[8];
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        double acc = 0.0;
        for (int i = 0; i < n; ++i) acc += (double)emb[(long long)i * D + d];
        const float r = (float)(acc / (double)n);
        sref[d] = r;
        ref[d] = r;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float mx = 0.f;
    for (int i = warp; i < n; i += 8) {
        float acc = 0.f;
        for (int d = lane; d < D; d += 32) {
            const float df = sref[d] - emb[(long long)i * D + d] + eps;
            acc = fmaf(df, df, acc);
        }
        acc = warp_sum(acc);
        mx = fmaxf(mx, sqrtf(acc));
    }
    if (lane == 0) smax[warp] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        float m = smax[0];
        for (int i = 1; i < 8; ++i) m = fmaxf(m, smax[i]);
        thr[0] = m;
    }
}
cudaError_t launch_ref_embedding(const float* emb, int n, int D, float eps, float* ref, float* thr, cudaStream_t s) {
    if (n <= 0) return cudaErrorInvalidValue;
    ref_embedding_kernel<<<1, 256, D * sizeof(float), s>>>(emb, n, D, eps, ref, thr);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ radix sort
static constexpr int RS_THREADS = 256;
static constexpr int RS_CHUNK = 2048;     // elements per block per pass

__device__ __forceinline__ unsigned long long f64_key(double x) {
    x = x + 0.0;                                        // -0.0 -> +0.0 (numpy compares them equal)
    const unsigned long long u = (unsigned long long)__double_as_longlong(x);
    return (u >> 63) ? ~u : (u | 0x8000000000000000ULL);
}
__device__ __forceinline__ double key_f64(unsigned long long k) {
    const unsigned long long u = (k >> 63) ? (k & 0x7FFFFFFFFFFFFFFFULL) : ~k;
    return __longlong_as_double((long long)u);
}

__global__ void rs_init_kernel(const double* __restrict__ t, long long nt, const double* __restrict__ nn, long long n_n,
                               unsigned long long* __restrict__ keys, unsigned int* __restrict__ idx) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long n = nt + n_n;
    if (i >= n) return;
    const double v = i < nt ? t[i] : nn[i - nt];
    keys[i] = f64_key(v);
    idx[i] = (unsigned int)i;
}

__global__ void __launch_bounds__(RS_THREADS)
rs_hist_kernel(const unsigned long long* __restrict__ keys, long long n, int shift, unsigned int* __restrict__ hist, int G) {
    __shared__ unsigned int h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const long long beg = (long long)blockIdx.x * RS_CHUNK;
    const long long end = beg + RS_CHUNK < n ? beg + RS_CHUNK : n;
    for (long long i = beg + threadIdx.x; i < end; i += RS_THREADS) atomicAdd(&h[(keys[i] >> shift) & 255ULL], 1u);
    __syncthreads();
    hist[(long long)threadIdx.x * G + blockIdx.x] = h[threadIdx.x];       // digit-major
}

// exclusive scan of hist[256*G] (digit-major) in place; single CTA of 1024 threads
__global__ void __launch_bounds__(1024)
rs_scan_kernel(unsigned int* __restrict__ hist, int total) {
    __shared__ unsigned int part[1024];
    const int tid = threadIdx.x;
    const int per = (total + 1023) / 1024;
    const int beg = tid * per, end = min(beg + per, total);
    unsigned int s = 0;
    for (int i = beg; i < end; ++i) s += hist[i];
    part[tid] = s;
    __syncthreads();
    for (int off = 1; off < 1024; off <<= 1) {
        unsigned int v = tid >= off ? part[tid - off] : 0u;
        __syncthreads();
        part[tid] += v;
        __syncthreads();
    }
    unsigned int run = tid == 0 ? 0u : part[tid - 1];
    for (int i = beg; i < end; ++i) { const unsigned int v = hist[i]; hist[i] = run; run += v; }
}

__global__ void __launch_bounds__(RS_THREADS)
rs_scatter_kernel(const unsigned long long* __restrict__ kin, const unsigned int* __restrict__ iin, long long n, int shift,
                  const unsigned int* __restrict__ hist, int G, unsigned long long* __restrict__ kout, unsigned int* __restrict__ iout) {
    __shared__ unsigned int base[256];
    __shared__ unsigned int wh[RS_THREADS / 32][256];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    base[tid] = hist[(long long)tid * G + blockIdx.x];
    const long long beg = (long long)blockIdx.x * RS_CHUNK;
    const long long end = beg + RS_CHUNK < n ? beg + RS_CHUNK : n;
    for (long long t0 = beg; t0 < end; t0 += RS_THREADS) {
        for (int w = 0; w < RS_THREADS / 32; ++w) wh[w][tid] = 0;
        __syncthreads();
        const long long i = t0 + tid;
        const bool ok = i < end;
        unsigned long long k = 0;
        unsigned int id = 0, d = 0;
        if (ok) { k = kin[i]; id = iin[i]; d = (unsigned int)((k >> shift) & 255ULL); }
        // rank among the lanes of this warp that hold the same digit (inactive lanes use a digit-out-of-range tag)
        const unsigned int tag = ok ? d : 0xFFFFFFFFu;
        const unsigned int peers = __match_any_sync(0xffffffffu, tag);
        const unsigned int rank = __popc(peers & ((1u << lane) - 1u));
        if (ok && rank == 0) wh[warp][d] = __popc(peers);
        __syncthreads();
        if (ok) {
            unsigned int pre = 0;
            for (int w = 0; w < warp; ++w) pre += wh[w][d];
            const unsigned int pos = base[d] + pre + rank;
            kout[pos] = k;
            iout[pos] = id;
        }
        __syncthreads();
        unsigned int tot = 0;
        for (int w = 0; w < RS_THREADS / 32; ++w) tot += wh[w][tid];
        base[tid] += tot;
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------ DET / EER
// Single CTA; thread t owns a contiguous slice of the sorted order.
__global__ void __launch_bounds__(1024)
eer_kernel(const unsigned long long* __restrict__ keys, const unsigned int* __restrict__ idx, long long n, long long nt, long long nn,
           double* __restrict__ result) {
    __shared__ long long cnt[1024];
    __shared__ double bestv[1024];
    __shared__ long long besti[1024];
    const int tid = threadIdx.x;
    const long long per = (n + 1023) / 1024;
    const long long beg = (long long)tid * per, end = beg + per < n ? beg + per : n;
    long long c = 0;
    for (long long i = beg; i < end; ++i) c += (idx[i] < (unsigned long long)nt) ? 1 : 0;
    cnt[tid] = c;
    __syncthreads();
    for (int off = 1; off < 1024; off <<= 1) {
        long long v = tid >= off ? cnt[tid - off] : 0;
        __syncthreads();
        cnt[tid] += v;
        __syncthreads();
    }
    long long run = tid == 0 ? 0 : cnt[tid - 1];
    // candidate 0 of the DET curve is (frr, far) = (0, 1): |diff| = 1 at curve index 0
    double bv = 2.0;
    long long bi = -1;
    if (tid == 0) { bv = 1.0; bi = 0; }
    const double dnt = (double)nt, dnn = (double)nn;
    for (long long i = beg; i < end; ++i) {
        run += (idx[i] < (unsigned long long)nt) ? 1 : 0;
        const double frr = (double)run / dnt;
        const double far = (double)(nn - ((i + 1) - run)) / dnn;
        const double df = fabs(frr - far);
        if (df < bv) { bv = df; bi = i + 1; }            // strict: keeps the first minimum
    }
    bestv[tid] = bv;
    besti[tid] = bi;
    __syncthreads();
    if (tid == 0) {
        double v = bestv[0];
        long long ix = besti[0];
        for (int t = 1; t < 1024; ++t) {
            if (besti[t] >= 0 && bestv[t] < v) { v = bestv[t]; ix = besti[t]; }   // slices are in index order
        }
        // recompute (frr, far) at ix
        double frr = 0.0, far = 1.0, thr;
        if (ix == 0) {
            thr = key_f64(keys[0]) - 0.001;
        } else {
            // cumulative target count at sorted position ix-1
            const long long pos = ix - 1;
            const int owner = (int)(pos / per);
            long long r = owner == 0 ? 0 : cnt[owner - 1];
            for (long long i = (long long)owner * per; i <= pos; ++i) r += (idx[i] < (unsigned long long)nt) ? 1 : 0;
            frr = (double)r / dnt;
            far = (double)(nn - ((pos + 1) - r)) / dnn;
            thr = key_f64(keys[pos]);
        }
        result[0] = (frr + far) / 2.0;
        result[1] = thr;
    }
}

int64_t eer_workspace_bytes(int64_t n) {
    const int64_t G = (n + RS_CHUNK - 1) / RS_CHUNK;
    const int64_t a = ((n * 8 + 255) / 256) * 256;
    const int64_t b = ((n * 4 + 255) / 256) * 256;
    const int64_t h = ((256 * (G > 0 ? G : 1) * 4 + 255) / 256) * 256;
    return 2 * a + 2 * b + h + 256;
}

// the stable sort shared by the EER and the DET-curve export: on return *ks / *is point at the sorted keys / original indices
static cudaError_t sort_scores(const double* target, int64_t nt, const double* nontarget, int64_t nn, void* work, cudaStream_t s,
                               unsigned long long** ks, unsigned int** is) {
    const int64_t n = nt + nn;
    if (n <= 0 || n >= (1LL << 31)) return cudaErrorInvalidValue;
    const int G = (int)((n + RS_CHUNK - 1) / RS_CHUNK);
    if ((int64_t)256 * G > (1 << 24)) return cudaErrorInvalidValue;
    char* w = reinterpret_cast<char*>(work);
    const int64_t a = ((n * 8 + 255) / 256) * 256;
    const int64_t b = ((n * 4 + 255) / 256) * 256;
    unsigned long long* k0 = reinterpret_cast<unsigned long long*>(w);
    unsigned long long* k1 = reinterpret_cast<unsigned long long*>(w + a);
    unsigned int* i0 = reinterpret_cast<unsigned int*>(w + 2 * a);
    unsigned int* i1 = reinterpret_cast<unsigned int*>(w + 2 * a + b);
    unsigned int* hist = reinterpret_cast<unsigned int*>(w + 2 * a + 2 * b);
    rs_init_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(target, nt, nontarget, nn, k0, i0);
    for (int pass = 0; pass < 8; ++pass) {
        const int shift = pass * 8;
        rs_hist_kernel<<<G, RS_THREADS, 0, s>>>(k0, n, shift, hist, G);
        rs_scan_kernel<<<1, 1024, 0, s>>>(hist, 256 * G);
        rs_scatter_kernel<<<G, RS_THREADS, 0, s>>>(k0, i0, n, shift, hist, G, k1, i1);
        unsigned long long* tk = k0; k0 = k1; k1 = tk;
        unsigned int* ti = i0; i0 = i1; i1 = ti;
    }
    *ks = k0; *is = i0;
    return cudaGetLastError();
}

cudaError_t launch_eer(const double* target, int64_t nt, const double* nontarget, int64_t nn, double* result, void* work, cudaStream_t s) {
    unsigned long long* k0 = nullptr;
    unsigned int* i0 = nullptr;
    cudaError_t e = sort_scores(target, nt, nontarget, nn, work, s, &k0, &i0);
    if (e != cudaSuccess) return e;
    eer_kernel<<<1, 1024, 0, s>>>(k0, i0, nt + nn, nt, nn, result);
    return cudaGetLastError();
}

// Full DET curve (evaluate_metrics.py:3-21): frr / far / thresholds, n + 1 points each, the same float64 operations as numpy
// (integer cumulative counts, one division per point).  Single CTA; thread t owns a contiguous slice of the sorted order.
__global__ void __launch_bounds__(1024)
det_kernel(const unsigned long long* __restrict__ keys, const unsigned int* __restrict__ idx, long long n, long long nt, long long nn,
           double* __restrict__ frr, double* __restrict__ far, double* __restrict__ thr) {
    __shared__ long long cnt[1024];
    const int tid = threadIdx.x;
    const long long per = (n + 1023) / 1024;
    const long long beg = (long long)tid * per, end = beg + per < n ? beg + per : n;
    long long c = 0;
    for (long long i = beg; i < end; ++i) c += (idx[i] < (unsigned long long)nt) ? 1 : 0;
    cnt[tid] = c;
    __syncthreads();
    for (int off = 1; off < 1024; off <<= 1) {
        long long v = tid >= off ? cnt[tid - off] : 0;
        __syncthreads();
        cnt[tid] += v;
        __syncthreads();
    }
    long long run = tid == 0 ? 0 : cnt[tid - 1];
    const double dnt = (double)nt, dnn = (double)nn;
    if (tid == 0) { frr[0] = 0.0; far[0] = 1.0; thr[0] = key_f64(keys[0]) - 0.001; }
    for (long long i = beg; i < end; ++i) {
        run += (idx[i] < (unsigned long long)nt) ? 1 : 0;
        frr[i + 1] = (double)run / dnt;
        far[i + 1] = (double)(nn - ((i + 1) - run)) / dnn;
        thr[i + 1] = key_f64(keys[i]);
    }
}

cudaError_t launch_det_curve(const double* target, int64_t nt, const double* nontarget, int64_t nn, double* frr, double* far, double* thr,
                             void* work, cudaStream_t s) {
    unsigned long long* k0 = nullptr;
    unsigned int* i0 = nullptr;
    cudaError_t e = sort_scores(target, nt, nontarget, nn, work, s, &k0, &i0);
    if (e != cudaSuccess) return e;
    det_kernel<<<1, 1024, 0, s>>>(k0, i0, nt + nn, nt, nn, frr, far, thr);
    return cudaGetLastError();
}


// ------------------------------------------------------------------------------------------------ no-allocation EER
// occ_eer_noalloc: occ_eer for callers that cannot allocate (CUDA-graph capture, tight serving loops, harnesses that forbid
// cudaMalloc).  No counterpart in the reference.  The sort scratch (two key
// planes, two index planes, the digit histogram) lives in the file-scope device array below, sized for
// EER_STATIC_MAX_N = 131 072 scores (3.06 MiB; the 71 237-score ASVspoof2019-LA eval set fits).  Every GPU that loads the
// library holds it until the process exits.
// In HOST code `g_eer_scratch` is a host-side stand-in for the symbol, not a device address: the address the kernels get
// comes from cudaGetSymbolAddress, asked once per device with that device current (eer_static_scratch below).
// Nothing in the scratch has to start at zero: rs_init writes every key / index the first pass reads, rs_hist writes all
// 256 * G counters the scan reads, and each scatter pass writes all n slots the next pass reads -- so the contents a previous
// call left behind are never observed.  All calls on one device SHARE the array: two calls that overlap on the GPU
// (different streams, or one graph launched while another call runs) race; serialise them on one stream or with events.
static constexpr long long EER_STATIC_MAX_N = 131072;
static constexpr long long eer_ws_bytes_c(long long n) {
    const long long G = (n + RS_CHUNK - 1) / RS_CHUNK;
    const long long a = ((n * 8 + 255) / 256) * 256;
    const long long b = ((n * 4 + 255) / 256) * 256;
    const long long h = ((256 * (G > 0 ? G : 1) * 4 + 255) / 256) * 256;
    return 2 * a + 2 * b + h + 256;
}
static constexpr long long EER_STATIC_BYTES = eer_ws_bytes_c(EER_STATIC_MAX_N);
static_assert(EER_STATIC_BYTES <= (4LL << 20), "the static EER scratch is meant to stay modest");

static __device__ __align__(256) unsigned char g_eer_scratch[EER_STATIC_BYTES];

static constexpr int EER_STATIC_MAX_DEVICES = 64;
static void* g_eer_scratch_ptr[EER_STATIC_MAX_DEVICES];      // per device; filled once under g_eer_scratch_mu
static std::mutex g_eer_scratch_mu;

// device address of g_eer_scratch on the CURRENT device (set-up: first call per device)
static cudaError_t eer_static_scratch(void** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= EER_STATIC_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_eer_scratch_mu);
    if (!g_eer_scratch_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_eer_scratch);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_eer_scratch_ptr[dev] = p;
    }
    *out = g_eer_scratch_ptr[dev];
    return cudaSuccess;
}

}  // namespace occ

extern "C" int occ_eer_noalloc(const double* target_dev, int64_t n_t, const double* nontarget_dev, int64_t n_n, double* result_dev,
                               void* stream) {
    using namespace occ;
    if (n_t < 1 || n_n < 1 || !target_dev || !nontarget_dev || !result_dev)
        return report_error(OCC_ERR_ARG, "occ_eer_noalloc: need >= 1 target and >= 1 nontarget score and non-null buffers");
    if (n_t > EER_STATIC_MAX_N || n_n > EER_STATIC_MAX_N || n_t + n_n > EER_STATIC_MAX_N)
        return report_error(OCC_ERR_ARG, "occ_eer_noalloc: " + std::to_string((long long)n_t) + " + " + std::to_string((long long)n_n) +
                                             " scores exceed the static scratch (OCC_EER_NOALLOC_MAX_SCORES = " +
                                             std::to_string(EER_STATIC_MAX_N) + "); use occ_eer with a caller-provided workspace");
    static_assert(EER_STATIC_MAX_N == OCC_EER_NOALLOC_MAX_SCORES, "header and kernel file disagree on the limit");
    void* work = nullptr;
    cudaError_t e = eer_static_scratch(&work);
    if (e != cudaSuccess)
        return report_error(OCC_ERR_CUDA, std::string("occ_eer_noalloc: cudaGetSymbolAddress(g_eer_scratch): ") + cudaGetErrorString(e));
    e = launch_eer(target_dev, n_t, nontarget_dev, n_n, result_dev, work, reinterpret_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return report_error(OCC_ERR_CUDA, std::string("occ_eer_noalloc: launch: ") + cudaGetErrorString(e));
    return OCC_OK;
}
