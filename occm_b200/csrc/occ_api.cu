// occ_api.cu -- the C ABI of libocc_b200.so (include/occ_b200.h): handle, weight staging / folding,
// and the stage orchestration of the scoring path
//   waveform -> conv feature encoder -> feature LN + projection -> grouped positional conv ->
//   24 pre-LN transformer layers -> final LN -> AASIST backend -> (emb, out) -> one-class score.
// Follows /root/reference/models/sslassist.py:506-597 (AModel.forward) with the fairseq wav2vec 2.0
// arithmetic restated in torchaudio/models/wav2vec2/components.py (cited per stage below).
// No CPU path exists in this library: every stage is a CUDA kernel launch on the caller's stream.
#include "../../include/occ_b200.h"
#include "kernels.h"
#include <map>
#include <string>
#include <vector>
#include <cstring>
#include <cstdlib>
#include <cstdio>
#include <cmath>
#include <dlfcn.h>
#include <mutex>

using namespace occ;

namespace {

struct Tensor {
    float* d = nullptr;
    std::vector<int64_t> shape;
    int64_t numel = 0;
};

struct ConvW {
    bf16* w = nullptr;          // [512][k*512] K-major (+ lo plane)
    int64_t plane = 0;
    const float *bias = nullptr, *g = nullptr, *b = nullptr;
    int k = 0;
};
struct LayerW {
    const float *ln1g, *ln1b, *ln2g, *ln2b;
    bf16 *wqkv, *wo, *w1, *w2;
    float* bqkv;
    const float *bo, *b1, *b2;
    // LayerNorm folded into the weights (bf16 mode, GemmDesc::ln_stat_in): column sums s and folded biases c; nullptr = not folded
    float *sqkv = nullptr, *cqkv = nullptr, *s1 = nullptr, *c1 = nullptr;
};
struct ResW {
    int ci, co;
    // tcgen05 B operands (bf16 hi/lo planes, [64][taps*64]) and 64-padded per-channel vectors; block 0 keeps the raw fp32
    // conv1 / conv_downsample tensors (single input channel, SIMT front kernel)
    bf16 *c1w = nullptr, *c2w = nullptr, *dw = nullptr;
    float *c1b = nullptr, *c2b = nullptr, *db = nullptr;
    float *bn2s = nullptr, *bn2t = nullptr;
    const float *raw_c1w = nullptr, *raw_c1b = nullptr, *raw_dw = nullptr, *raw_db = nullptr;
    float *raw_bn2s = nullptr, *raw_bn2t = nullptr;
};
struct HtrgW {
    int din;
    const float *W1, *b1, *W2, *b2;
    GatW g;
};

static const int kConvK[7] = {10, 3, 3, 3, 3, 2, 2};
static const int kConvS[7] = {5, 2, 2, 2, 2, 2, 2};
static const int kResFilts[6][2] = {{1, 32}, {32, 32}, {32, 64}, {64, 64}, {64, 64}, {64, 64}};
static const int kHp = 45;             // padded rows of the residual-encoder grid: 1 zero row + up to 43 rows + 1 zero row
static const int64_t kSlack = 65536;   // elements of tail padding on activation buffers read through TMA boxes

}  // namespace

struct occ_handle {
    int device = 0, dtype = 0, max_batch = 0, max_samples = 0, num_sms = 148;
    int P = 1;                      // bf16 planes per GEMM operand (2 = hi/lo split, fp32-accurate)
    bool simt = false;              // OCC_GEMM_SIMT=1: run every GEMM on the SIMT check kernel (debug)
    int ln_fold = 0;                // bf16 mode: transformer LayerNorms folded into the GEMMs (decided at occ_create): bit 0 = LN1 into QKV
                                    // (layers 1..23, statistics from FC2's epilogue), bit 1 = LN2 into FC1 (statistics from out-proj's)
    float* lnstat = nullptr;        // per-row partial sums of the residual stream [Mmax][12][2]
    bool finalized = false;
    bool has_lcnn = false;          // lcnn.* weights loaded: occ_forward_lcnn / occ_lcnn_backend available (models/lcnn.py ssl_lcnn)
    LcnnW lcnn;
    float *l_y1 = nullptr, *l_a2 = nullptr, *l_y2 = nullptr, *l_a3 = nullptr, *l_y3 = nullptr, *l_out = nullptr;
    bool has_resnet = false;        // resnet34.* weights loaded: occ_forward_resnet34 / occ_resnet34_backend (models/senet.py ssl_resnet34)
    ResnetW resnet;
    float *r_buf[5] = {nullptr, nullptr, nullptr, nullptr, nullptr}, *r_vec = nullptr;
    bool has_backend = false;       // false: frontend-only handle (stand-alone SSLModel, models/xlsr.py): occ_extract_feat / stage entry points only
    std::string err;
    std::map<std::string, Tensor> raw;
    std::vector<void*> owned;       // device allocations freed in occ_destroy
    int64_t launches = 0;

    // ---- optional per-range CUDA-event profiling (occ_profile / occ_profile_read)
    struct ProfRange { int tag; double flops; cudaEvent_t a, b; };
    bool prof_on = false;
    std::vector<ProfRange> prof;
    std::vector<cudaEvent_t> prof_pool;

    // ---- packed weights
    const float *c0w = nullptr, *c0b = nullptr, *c0g = nullptr, *c0e = nullptr;
    float* c0frag = nullptr;        // conv0 on mma.sync: centred, split weights in A-fragment order + splatted gamma / beta (frontend.cu)
    ConvW conv[7];
    const float *flng = nullptr, *flnb = nullptr;
    bf16* wproj = nullptr; const float* bproj = nullptr;
    bf16* wpos = nullptr; const float* bpos = nullptr;
    bf16* wpos_toep = nullptr;      // Toeplitz-expanded pos-conv weight [4096][131*64] (bf16 mode): four output frames per N = 256 tile
    bf16* xpad16 = nullptr;         // flat padded frame sequence the Toeplitz pos-conv reads: utterance b at frames [b*Tp, b*Tp + T)
    int64_t xpad_elems = 0;
    LayerW layer[24];
    const float *elng = nullptr, *elnb = nullptr;
    bf16* wll = nullptr; const float* bll = nullptr;     // always two planes
    float bn0_scale = 1.f, bn0_shift = 0.f;
    ResW res[6];
    float *bn1s = nullptr, *bn1t = nullptr;
    const float *attW1 = nullptr, *attb1 = nullptr, *attb2 = nullptr;
    float *attbs = nullptr, *attbt = nullptr;
    const float* attW2 = nullptr;
    const float* posS = nullptr;
    GatW gatS, gatT;
    HtrgW htrg[4];                  // ST11, ST12, ST21, ST22
    const float *poolw[6], *poolb[6];   // S, T, hS1, hT1, hS2, hT2
    const float *outW = nullptr, *outb = nullptr;

    // ---- workspace
    int L[7] = {0, 0, 0, 0, 0, 0, 0};
    int Tmax = 0, Wmax = 0;
    bf16 *actA = nullptr, *actB = nullptr;
    int64_t actA_plane = 0, actB_plane = 0;
    float* c32 = nullptr;
    float *x32 = nullptr, *feat32 = nullptr, *qkv32 = nullptr, *ll32 = nullptr;
    bf16 *h16 = nullptr, *qkv16 = nullptr, *ctx16 = nullptr, *f16 = nullptr, *feat16 = nullptr, *y16 = nullptr;
    float* y32 = nullptr;           // sub-block output awaiting its residual add (fp32 mode; y16 in bf16 mode)
    int64_t pl1024 = 0, pl4096 = 0;     // plane strides of [Mmax,1024] / [Mmax,4096] bf16 buffers
    float *bx0 = nullptr, *wmap = nullptr, *xf32 = nullptr;     // xf32: residual-encoder output, channel-last padded fp32
    bf16 *cX[2] = {nullptr, nullptr}, *cA = nullptr, *cI = nullptr;   // channel-last padded hi/lo planes [B][45][Wmax+2][64]
    int64_t cplane = 0;
    float *bn1s64 = nullptr, *bn1t64 = nullptr;
    float *eS = nullptr, *eT = nullptr, *yS = nullptr, *yT = nullptr, *oS = nullptr, *oT = nullptr;
    float *xc1[2], *y1[2], *m1[2], *S1[2], *T1[2], *xc2[2], *y2[2], *maug[2];
    int *klen = nullptr, *wlen = nullptr, *kT1 = nullptr, *kT2 = nullptr;
    float *wav_dev = nullptr, *emb_dev = nullptr, *out_dev = nullptr, *score_dev = nullptr, *ref_dev = nullptr;
    std::vector<int> hbuf;
};

namespace {

thread_local std::string g_err;

int fail(occ_handle* h, int code, const std::string& msg) {
    if (h) h->err = msg;
    g_err = msg;
    return code;
}
int cuda_fail(occ_handle* h, cudaError_t e, const char* what) {
    std::string m = std::string(what) + ": " + cudaGetErrorString(e);
    if (h && !h->err.empty() && h->err.find("cuTensorMap") != std::string::npos) m += " [" + h->err + "]";
    return fail(h, OCC_ERR_CUDA, m);
}
#define CK(call)                                                      \
    do {                                                              \
        cudaError_t _e = (call);                                      \
        if (_e != cudaSuccess) return cuda_fail(h, _e, #call);        \
    } while (0)
#define LAUNCH(n, call)                                               \
    do {                                                              \
        cudaError_t _e = (call);                                      \
        if (_e != cudaSuccess) return cuda_fail(h, _e, #call);        \
        h->launches += (n);                                           \
    } while (0)

cudaEvent_t prof_event(occ_handle* h) {
    if (!h->prof_pool.empty()) { cudaEvent_t e = h->prof_pool.back(); h->prof_pool.pop_back(); return e; }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}
// Range timing on the launch stream: events are recorded only while profiling is enabled.
struct ProfScope {
    occ_handle* h; cudaStream_t s; int idx = -1;
    ProfScope(occ_handle* h_, int tag, double flops, cudaStream_t s_) : h(h_), s(s_) {
        if (!h->prof_on) return;
        occ_handle::ProfRange r{tag, flops, prof_event(h), prof_event(h)};
        cudaEventRecord(r.a, s);
        h->prof.push_back(r);
        idx = (int)h->prof.size() - 1;
    }
    ~ProfScope() { if (idx >= 0) cudaEventRecord(h->prof[idx].b, s); }
};

template <typename T>
int dalloc(occ_handle* h, T** p, int64_t n) {
    void* q = nullptr;
    if (n <= 0) n = 1;
    cudaError_t e = cudaMalloc(&q, (size_t)n * sizeof(T));
    if (e != cudaSuccess) return fail(h, OCC_ERR_NOMEM, std::string("cudaMalloc failed: ") + cudaGetErrorString(e));
    e = cudaMemset(q, 0, (size_t)n * sizeof(T));
    if (e != cudaSuccess) return cuda_fail(h, e, "cudaMemset");
    h->owned.push_back(q);
    *p = reinterpret_cast<T*>(q);
    return OCC_OK;
}
#define DA(p, n)                                  \
    do {                                          \
        int _r = dalloc(h, &(p), (int64_t)(n));   \
        if (_r != OCC_OK) return _r;              \
    } while (0)

int frames_after(int n, int layers) {
    long long L = n;
    for (int i = 0; i < layers; ++i) {
        L = L < kConvK[i] ? 0 : (L - kConvK[i]) / kConvS[i] + 1;
    }
    return (int)L;
}
int frames_layer(long long Lin, int i) {
    if (Lin < kConvK[i]) return 0;
    return (int)((Lin - kConvK[i]) / kConvS[i] + 1);
}

const Tensor* find(occ_handle* h, const std::string& k) {
    auto it = h->raw.find(k);
    return it == h->raw.end() ? nullptr : &it->second;
}

// -------------------------------------------------------------------------------------------- GEMM helpers
int run_gemm(occ_handle* h, const GemmDesc& d, cudaStream_t s) {
    cudaError_t e;
    if (h->simt) e = launch_gemm_simt(d, s);
    else e = launch_gemm_tc(d, s, h->num_sms, &h->err);
    if (e != cudaSuccess) return cuda_fail(h, e, "gemm launch");
    h->launches += 1;
    return OCC_OK;
}

GemmDesc lin_desc(const bf16* A, int64_t a_plane, int M, int K, const bf16* W, int64_t w_plane, int N, int split) {
    GemmDesc d;
    d.A = A; d.a_cols = K; d.a_sub = 1; d.a_sub_stride = 0; d.a_row_stride = K; d.a_rows = M;
    d.a_batch_stride = 0; d.nbatch = 1; d.a_plane_stride = a_plane;
    d.W = W; d.N = N; d.w_plane_stride = w_plane;
    d.nkb = K / 64; d.kb_per_tap = K / 64; d.ntaps_w = 1;
    d.split = split; d.rows_per_batch = M;
    d.BN = (N % 256 == 0) ? 256 : (N % 128 == 0 ? 128 : 64);
    return d;
}

struct Key {
    occ_handle* h;
    std::string prefix;
    bool ok = true;
    std::string missing;
    const Tensor* t(const std::string& k, int64_t numel) {
        const Tensor* x = find(h, prefix + k);
        if (!x || x->numel != numel) {
            if (ok) missing = prefix + k + (x ? " (wrong size)" : " (missing)");
            ok = false;
            static Tensor dummy;
            return &dummy;
        }
        return x;
    }
    const float* p(const std::string& k, int64_t numel) { return t(k, numel)->d; }
};

}  // namespace

// ================================================================================================ basic API
extern "C" int occ_version(void) { return 100; }

extern "C" const char* occ_last_error(occ_handle* h) { return h ? h->err.c_str() : g_err.c_str(); }

extern "C" int occ_num_frames(int n_samples) { return frames_after(n_samples, 7); }

extern "C" int64_t occ_kernel_launches(occ_handle* h) { return h ? h->launches : 0; }

extern "C" int occ_profile(occ_handle* h, int enable) {
    if (!h) return fail(h, OCC_ERR_ARG, "occ_profile: null handle");
    if (enable) {
        for (auto& r : h->prof) { h->prof_pool.push_back(r.a); h->prof_pool.push_back(r.b); }
        h->prof.clear();
    }
    h->prof_on = enable != 0;
    return OCC_OK;
}

extern "C" int occ_profile_read(occ_handle* h, int tag, double* ms_total, int64_t* ranges, double* flops_total) {
    if (!h || tag < 0 || tag >= OCC_PROF_NTAGS) return fail(h, OCC_ERR_ARG, "occ_profile_read: bad argument");
    double ms = 0.0, fl = 0.0;
    int64_t n = 0;
    for (auto& r : h->prof) {
        const bool xf = r.tag >= OCC_PROF_XF_QKV && r.tag <= OCC_PROF_XF_FC2;
        if (r.tag != tag && !(tag == OCC_PROF_XF_GEMM && xf)) continue;
        CK(cudaEventSynchronize(r.b));
        float t = 0.f;
        CK(cudaEventElapsedTime(&t, r.a, r.b));
        ms += t; fl += r.flops; ++n;
    }
    if (ms_total) *ms_total = ms;
    if (ranges) *ranges = n;
    if (flops_total) *flops_total = fl;
    return OCC_OK;
}

extern "C" void occ_destroy(occ_handle* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    for (auto& kv : h->raw) if (kv.second.d) cudaFree(kv.second.d);
    for (void* p : h->owned) cudaFree(p);
    for (auto& r : h->prof) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
    for (cudaEvent_t e : h->prof_pool) cudaEventDestroy(e);
    delete h;
}

extern "C" int occ_create(occ_handle** out, int device, int dtype, int max_batch, int max_samples) {
    if (!out) return fail(nullptr, OCC_ERR_ARG, "occ_create: null handle pointer");
    *out = nullptr;
    if (dtype != OCC_DTYPE_FP32 && dtype != OCC_DTYPE_BF16) return fail(nullptr, OCC_ERR_ARG, "occ_create: dtype must be 0 (fp32) or 1 (bf16)");
    if (max_batch < 1 || max_samples < 400) return fail(nullptr, OCC_ERR_ARG, "occ_create: max_batch >= 1 and max_samples >= 400 required");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(nullptr, OCC_ERR_CUDA, std::string("occ_create: no CUDA device (this library has no CPU path): ") + cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return fail(nullptr, OCC_ERR_ARG, "occ_create: bad device index");
    e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail(nullptr, OCC_ERR_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) return fail(nullptr, OCC_ERR_CUDA, std::string("cudaGetDeviceProperties: ") + cudaGetErrorString(e));
    if (prop.major != 10) return fail(nullptr, OCC_ERR_CUDA, "occ_create: libocc_b200 is built for sm_100a (B200) only");
    occ_handle* h = new occ_handle();
    h->device = device; h->dtype = dtype; h->max_batch = max_batch; h->max_samples = max_samples;
    h->num_sms = prop.multiProcessorCount;
    h->P = dtype == OCC_DTYPE_FP32 ? 2 : 1;
    const char* sv = getenv("OCC_GEMM_SIMT");
    h->simt = sv && sv[0] == '1';
    {
        const char* lf = getenv("OCC_LN_FOLD");     // development A/B switch (0 = LayerNorm kernels); default: both folds
        h->ln_fold = (dtype == OCC_DTYPE_BF16 && !h->simt) ? (lf ? atoi(lf) & 3 : 3) : 0;
    }
    long long Lc = max_samples;
    for (int i = 0; i < 7; ++i) { Lc = frames_layer(Lc, i); h->L[i] = (int)Lc; }
    h->Tmax = h->L[6];
    h->Wmax = h->Tmax / 3;
    if (h->Wmax < 4) { delete h; return fail(nullptr, OCC_ERR_ARG, "occ_create: max_samples too short"); }
    const int64_t B = max_batch, P = h->P, M = B * h->Tmax;
    int r = OCC_OK;
    auto go = [&]() -> int {
        h->actA_plane = B * h->L[0] * 512 + kSlack;
        h->actB_plane = B * h->L[1] * 512 + kSlack;
        DA(h->actA, h->actA_plane * P);
        DA(h->actB, h->actB_plane * P);
        DA(h->c32, B * h->L[1] * 512);
        h->pl1024 = M * 1024 + kSlack;
        h->pl4096 = M * 4096 + kSlack;
        DA(h->x32, M * 1024); DA(h->feat32, M * 1024); DA(h->ll32, M * 128);
        DA(h->h16, h->pl1024 * P); DA(h->ctx16, h->pl1024 * P); DA(h->f16, h->pl4096 * P);
        DA(h->feat16, h->pl1024 * 2);
        if (h->ln_fold) DA(h->lnstat, M * 24);
        if (dtype == OCC_DTYPE_BF16 && !h->simt) {
            h->xpad_elems = B * (int64_t)(h->Tmax + 64 + 4) * 1024 + kSlack;
            DA(h->xpad16, h->xpad_elems);
        }
        if (dtype == OCC_DTYPE_FP32) { DA(h->qkv32, M * 3072); DA(h->y32, M * 1024); } else { DA(h->qkv16, M * 3072 + kSlack); DA(h->y16, M * 1024); }
        const int64_t Wm = h->Wmax;
        DA(h->bx0, B * 42 * Wm);
        h->cplane = B * kHp * (Wm + 2) * 64 + kSlack;
        DA(h->cX[0], h->cplane * 2); DA(h->cX[1], h->cplane * 2); DA(h->cA, h->cplane * 2); DA(h->cI, h->cplane * 2);
        DA(h->xf32, B * kHp * (Wm + 2) * 64);
        DA(h->wmap, B * 64 * 42 * Wm);
        DA(h->eS, B * 42 * 64); DA(h->eT, B * Wm * 64); DA(h->yS, B * 42 * 64); DA(h->yT, B * Wm * 64);
        DA(h->oS, B * 21 * 64); DA(h->oT, B * (Wm / 2 + 1) * 64);
        const int64_t k1 = Wm / 2 + 1, k2 = k1 / 2 + 1;
        for (int z = 0; z < 2; ++z) {
            DA(h->xc1[z], B * (k1 + 21) * 64); DA(h->y1[z], B * (k1 + 21) * 32); DA(h->m1[z], B * 32);
            DA(h->S1[z], B * 10 * 32); DA(h->T1[z], B * k2 * 32);
            DA(h->xc2[z], B * (k2 + 10) * 32); DA(h->y2[z], B * (k2 + 10) * 32); DA(h->maug[z], B * 32);
        }
        DA(h->klen, B); DA(h->wlen, B); DA(h->kT1, B); DA(h->kT2, B);
        DA(h->wav_dev, B * (int64_t)max_samples); DA(h->emb_dev, B * 160); DA(h->out_dev, B * 2);
        DA(h->score_dev, B); DA(h->ref_dev, 160);
        return OCC_OK;
    };
    r = go();
    if (r != OCC_OK) { g_err = h->err; occ_destroy(h); return r; }
    *out = h;
    return OCC_OK;
}

// ================================================================================================ weights
static std::string normalise_key(const char* name) {
    std::string k(name);
    if (k.rfind("module.", 0) == 0) k = k.substr(7);
    if (k.rfind("frontend.model.", 0) == 0) k = "ssl_model.model." + k.substr(15);
    if (k.rfind("senet34_branch.", 0) == 0) k = "resnet34." + k.substr(15);            // OCCM's names for its two branches (models/occm.py:51-52)
    if (k.rfind("lcnn_branch.", 0) == 0) k = "lcnn." + k.substr(12);      // ssl_lcnn / OCCM name the extractor `frontend` (models/lcnn.py:239)
    const std::string a = "parametrizations.weight.original0", b = "parametrizations.weight.original1";
    size_t p = k.find(a);
    if (p != std::string::npos) k = k.substr(0, p) + "weight_g";
    p = k.find(b);
    if (p != std::string::npos) k = k.substr(0, p) + "weight_v";
    return k;
}
static bool dropped_key(const std::string& k) {
    const char* pre = "ssl_model.model.";
    if (k.rfind(pre, 0) == 0) {
        const std::string r = k.substr(strlen(pre));
        if (r == "mask_emb" || r.rfind("quantizer.", 0) == 0 || r.rfind("project_q.", 0) == 0 || r.rfind("final_proj.", 0) == 0 ||
            r.rfind("target_glu", 0) == 0 || r.rfind("layer_norm_q", 0) == 0)
            return true;
    }
    if (k.find("num_batches_tracked") != std::string::npos) return true;
    // Residual_block.bn1 output is discarded at models/sslassist.py:408-415 (conv1 sees the raw input)
    if (k.rfind("encoder.", 0) == 0 && k.find(".bn1.") != std::string::npos) return true;
    return false;
}

extern "C" int occ_load_weights(occ_handle* h, const char* name, const void* ptr, const int64_t* shape, int ndim, int dtype) {
    if (!h || !name) return fail(h, OCC_ERR_ARG, "occ_load_weights: null argument");
    if (h->finalized) return fail(h, OCC_ERR_STATE, "occ_load_weights: weights already finalized (create a new handle)");
    const std::string k = normalise_key(name);
    if (dropped_key(k)) return OCC_OK;
    if (dtype != 0) return fail(h, OCC_ERR_WEIGHT, "occ_load_weights: " + k + ": only float32 tensors are accepted");
    if (!ptr || ndim < 0 || ndim > 8) return fail(h, OCC_ERR_ARG, "occ_load_weights: bad pointer / ndim for " + k);
    CK(cudaSetDevice(h->device));
    Tensor t;
    t.numel = 1;
    for (int i = 0; i < ndim; ++i) { t.shape.push_back(shape[i]); t.numel *= shape[i]; }
    if (t.numel <= 0) return fail(h, OCC_ERR_WEIGHT, "occ_load_weights: empty tensor " + k);
    auto it = h->raw.find(k);
    if (it != h->raw.end()) { cudaFree(it->second.d); h->raw.erase(it); }
    CK(cudaMalloc(&t.d, (size_t)t.numel * sizeof(float)));
    cudaError_t e = cudaMemcpy(t.d, ptr, (size_t)t.numel * sizeof(float), cudaMemcpyDefault);
    if (e != cudaSuccess) { cudaFree(t.d); return cuda_fail(h, e, "cudaMemcpy(weight)"); }
    h->raw[k] = t;
    return OCC_OK;
}

static int fold_bn(occ_handle* h, Key& K, const std::string& p, int C, float** sc, float** sh) {
    DA(*sc, C); DA(*sh, C);
    const float* w = K.p(p + "weight", C); const float* b = K.p(p + "bias", C);
    const float* m = K.p(p + "running_mean", C); const float* v = K.p(p + "running_var", C);
    if (!K.ok) return OCC_OK;
    LAUNCH(1, launch_fold_bn(w, b, m, v, C, *sc, *sh, 0));
    return OCC_OK;
}

extern "C" int occ_finalize_weights(occ_handle* h) {
    if (!h) return fail(h, OCC_ERR_ARG, "occ_finalize_weights: null handle");
    if (h->finalized) return OCC_OK;
    CK(cudaSetDevice(h->device));
    cudaStream_t s = 0;
    const int P = h->P;
    Key F{h, "ssl_model.model."};
    // ---- conv feature encoder (torchaudio components.py:53-143)
    h->c0w = F.p("feature_extractor.conv_layers.0.0.weight", 512 * 10);
    h->c0b = F.p("feature_extractor.conv_layers.0.0.bias", 512);
    h->c0g = F.p("feature_extractor.conv_layers.0.2.1.weight", 512);
    h->c0e = F.p("feature_extractor.conv_layers.0.2.1.bias", 512);
    DA(h->c0frag, (int64_t)(conv0_frag_bytes() / sizeof(float)));
    if (F.ok) LAUNCH(1, launch_conv0_pack_frag(h->c0w, h->c0b, h->c0g, h->c0e, h->c0frag, s));
    for (int i = 1; i < 7; ++i) {
        const std::string q = "feature_extractor.conv_layers." + std::to_string(i) + ".";
        ConvW& c = h->conv[i];
        c.k = kConvK[i];
        const int64_t n = 512LL * 512 * c.k;
        const float* w = F.p(q + "0.weight", n);
        c.bias = F.p(q + "0.bias", 512); c.g = F.p(q + "2.1.weight", 512); c.b = F.p(q + "2.1.bias", 512);
        c.plane = n;
        DA(c.w, n * P);
        if (F.ok) LAUNCH(1, launch_pack_conv1d(w, 512, 512, c.k, c.w, P == 2 ? c.plane : 0, s));
    }
    // ---- feature projection (components.py:146-182)
    h->flng = F.p("layer_norm.weight", 512); h->flnb = F.p("layer_norm.bias", 512);
    {
        const float* w = F.p("post_extract_proj.weight", 1024 * 512);
        h->bproj = F.p("post_extract_proj.bias", 1024);
        DA(h->wproj, 1024LL * 512 * P);
        if (F.ok) LAUNCH(1, launch_pack_linear(w, 1024, 512, 0, 1.f, h->wproj, P == 2 ? 1024LL * 512 : 0, s));
    }
    // ---- positional conv (components.py:185-234), weight_norm(dim=2) folded
    {
        const float* v = F.p("encoder.pos_conv.0.weight_v", 1024LL * 64 * 128);
        const float* g = F.p("encoder.pos_conv.0.weight_g", 128);
        h->bpos = F.p("encoder.pos_conv.0.bias", 1024);
        float* norms = nullptr;
        DA(norms, 128);
        DA(h->wpos, 1024LL * 8192 * P);
        if (F.ok) LAUNCH(2, launch_pack_posconv(v, g, norms, h->wpos, P == 2 ? 1024LL * 8192 : 0, s));
        static const bool toep_env = [] { const char* e = getenv("OCC_POS_TOEPLITZ"); return !(e && e[0] == '0'); }();     // 0: previous form, for A/B runs
        if (F.ok && h->xpad16 != nullptr && toep_env) {
            DA(h->wpos_toep, 4096LL * 131 * 64);
            LAUNCH(2, launch_pack_posconv_toeplitz(v, g, norms, h->wpos_toep, s));
        }
    }
    // ---- transformer layers (components.py:237-401); q scaled by 1/sqrt(64) (exact: power of two)
    for (int l = 0; l < 24; ++l) {
        const std::string q = "encoder.layers." + std::to_string(l) + ".";
        LayerW& w = h->layer[l];
        w.ln1g = F.p(q + "self_attn_layer_norm.weight", 1024); w.ln1b = F.p(q + "self_attn_layer_norm.bias", 1024);
        w.ln2g = F.p(q + "final_layer_norm.weight", 1024); w.ln2b = F.p(q + "final_layer_norm.bias", 1024);
        const int64_t sq = 1024LL * 1024;
        DA(w.wqkv, 3 * sq * P); DA(w.wo, sq * P); DA(w.w1, 4 * sq * P); DA(w.w2, 4 * sq * P); DA(w.bqkv, 3072);
        const float* wq = F.p(q + "self_attn.q_proj.weight", sq); const float* wk = F.p(q + "self_attn.k_proj.weight", sq);
        const float* wv = F.p(q + "self_attn.v_proj.weight", sq); const float* wo = F.p(q + "self_attn.out_proj.weight", sq);
        const float* bq = F.p(q + "self_attn.q_proj.bias", 1024); const float* bk = F.p(q + "self_attn.k_proj.bias", 1024);
        const float* bv = F.p(q + "self_attn.v_proj.bias", 1024);
        w.bo = F.p(q + "self_attn.out_proj.bias", 1024);
        const float* w1 = F.p(q + "fc1.weight", 4 * sq); const float* w2 = F.p(q + "fc2.weight", 4 * sq);
        w.b1 = F.p(q + "fc1.bias", 4096); w.b2 = F.p(q + "fc2.bias", 1024);
        if (!F.ok) continue;
        const int64_t pq = P == 2 ? 3 * sq : 0;
        LAUNCH(1, launch_pack_linear(wo, 1024, 1024, 0, 1.f, w.wo, P == 2 ? sq : 0, s));
        LAUNCH(1, launch_pack_linear(w2, 1024, 4096, 0, 1.f, w.w2, P == 2 ? 4 * sq : 0, s));
        LAUNCH(1, launch_scale_copy(bq, 1024, 1024, 0.125f, w.bqkv, s));
        LAUNCH(1, launch_scale_copy(bk, 1024, 0, 1.f, w.bqkv + 1024, s));
        LAUNCH(1, launch_scale_copy(bv, 1024, 0, 1.f, w.bqkv + 2048, s));
        if ((h->ln_fold & 1) && l > 0) {
            // LN1 folded into QKV (layers 1..23; layer 0's LN1 also adds the positional-conv term and stays a kernel)
            DA(w.sqkv, 3072); DA(w.cqkv, 3072);
            LAUNCH(1, launch_pack_linear_ln(wq, 1024, 1024, 1024, 0.125f, w.ln1g, w.ln1b, bq, w.wqkv, w.sqkv, w.cqkv, s));
            LAUNCH(1, launch_pack_linear_ln(wk, 1024, 1024, 0, 1.f, w.ln1g, w.ln1b, bk, w.wqkv + sq, w.sqkv + 1024, w.cqkv + 1024, s));
            LAUNCH(1, launch_pack_linear_ln(wv, 1024, 1024, 0, 1.f, w.ln1g, w.ln1b, bv, w.wqkv + 2 * sq, w.sqkv + 2048, w.cqkv + 2048, s));
        } else {
            LAUNCH(1, launch_pack_linear(wq, 1024, 1024, 1024, 0.125f, w.wqkv, pq, s));
            LAUNCH(1, launch_pack_linear(wk, 1024, 1024, 0, 1.f, w.wqkv + sq, pq, s));
            LAUNCH(1, launch_pack_linear(wv, 1024, 1024, 0, 1.f, w.wqkv + 2 * sq, pq, s));
        }
        if (h->ln_fold & 2) {
            DA(w.s1, 4096); DA(w.c1, 4096);
            LAUNCH(1, launch_pack_linear_ln(w1, 4096, 1024, 0, 1.f, w.ln2g, w.ln2b, w.b1, w.w1, w.s1, w.c1, s));
        } else {
            LAUNCH(1, launch_pack_linear(w1, 4096, 1024, 0, 1.f, w.w1, P == 2 ? 4 * sq : 0, s));
        }
    }
    h->elng = F.p("encoder.layer_norm.weight", 1024); h->elnb = F.p("encoder.layer_norm.bias", 1024);
    if (!F.ok) return fail(h, OCC_ERR_WEIGHT, "occ_finalize_weights: frontend weight " + F.missing);

    // ---- AASIST backend (models/sslassist.py:446-504).  A state dict without any backend tensor (the stand-alone SSLModel of
    // models/xlsr.py:25-52 hands over only ssl_model.* keys) yields a frontend-only handle.
    Key K{h, ""};
    h->has_backend = false;
    h->has_lcnn = false;
    h->has_resnet = false;
    for (auto& kv : h->raw) {
        if (kv.first.rfind("lcnn.", 0) == 0) h->has_lcnn = true;
        else if (kv.first.rfind("resnet34.", 0) == 0) h->has_resnet = true;
        else if (kv.first.rfind("ssl_model.", 0) != 0) h->has_backend = true;
    }
    if (h->has_resnet) {
        // ---- SE-ResNet34 backend (models/senet.py:64-156): conv weights in their checkpoint layout, every BatchNorm folded
        Key R{h, "resnet34."};
        ResnetW& w = h->resnet;
        std::memset(&w, 0, sizeof w);
        auto bn = [&](const std::string& p, int C, const float** sc, const float** sh) -> int {
            float *a = nullptr, *b = nullptr;
            int r = fold_bn(h, R, p, C, &a, &b);
            *sc = a; *sh = b;
            return r;
        };
        w.conv1 = R.p("conv1.weight", 16 * 49);
        if (R.ok) {
            float* f = nullptr;
            DA(f, (int64_t)(rs_stem_frag_bytes() / sizeof(float)));
            LAUNCH(1, launch_rs_pack_stem(w.conv1, f, s));
            w.conv1f = f;
        }
        { int r = bn("bn1.", 16, &w.bn1s, &w.bn1t); if (r) return r; }
        static const int planes[4] = {16, 32, 64, 128}, blocks[4] = {3, 4, 6, 3};
        int inpl = 16, nb = 0;
        for (int li = 0; li < 4; ++li) {
            for (int b = 0; b < blocks[li]; ++b, ++nb) {
                ResnetBlockW& k = w.blk[nb];
                const std::string p = "layer" + std::to_string(li + 1) + "." + std::to_string(b) + ".";
                const int C = planes[li], cin = b == 0 ? inpl : C;
                k.planes = C; k.stride = (b == 0 && li > 0) ? 2 : 1;
                k.c1 = R.p(p + "conv1.weight", (int64_t)C * cin * 9);
                { int r = bn(p + "bn1.", C, &k.b1s, &k.b1t); if (r) return r; }
                k.c2 = R.p(p + "conv2.weight", (int64_t)C * C * 9);
                { int r = bn(p + "bn2.", C, &k.b2s, &k.b2t); if (r) return r; }
                if (R.ok) {      // the 3x3 convs also in mma.sync B-fragment order (senet.cu: rs_conv3_mma)
                    float *f1 = nullptr, *f2 = nullptr;
                    DA(f2, (int64_t)(rs_frag_bytes(C, C) / sizeof(float)));
                    LAUNCH(1, launch_rs_pack_frag(k.c2, C, C, f2, s));
                    k.c2f = f2;
                    DA(f1, (int64_t)(rs_frag_bytes(cin, C) / sizeof(float)));
                    LAUNCH(1, launch_rs_pack_frag(k.c1, cin, C, f1, s));
                    k.c1f = f1;
                }
                k.se1 = R.p(p + "se.fc.0.weight", (int64_t)(C / 16) * C);
                k.se2 = R.p(p + "se.fc.2.weight", (int64_t)C * (C / 16));
                if (b == 0 && (k.stride != 1 || inpl != C)) {
                    k.dw = R.p(p + "downsample.0.weight", (int64_t)C * inpl);
                    int r = bn(p + "downsample.1.", C, &k.dbs, &k.dbt); if (r) return r;
                }
            }
            inpl = planes[li];
        }
        w.nblocks = nb;
        w.emb_w = R.p("embedding.weight", 128 * 128); w.emb_b = R.p("embedding.bias", 128);
        w.cls_w = R.p("classifier.weight", 2 * 128); w.cls_b = R.p("classifier.bias", 2);
        if (!R.ok) return fail(h, OCC_ERR_WEIGHT, "occ_finalize_weights: SE-ResNet34 weight " + R.missing);
        const int64_t B = h->max_batch, H1 = h->Tmax / 2 + 2, H2 = H1 / 2 + 2;
        DA(h->r_buf[0], B * 16 * H1 * 512);
        for (int i = 1; i < 5; ++i) DA(h->r_buf[i], B * 16 * H2 * 256);
        DA(h->r_vec, B * 128 * 2);
    }
    if (h->has_lcnn) {
        // ---- LCNN backend (models/lcnn.py:151-187): weights are used in their checkpoint layout; the two BatchNorm2d that the
        // forward really calls (layer2.2, layer3.2) are folded; group.bn (layerN.0.bn) is dead code in the reference (:146-150)
        Key L{h, "lcnn."};
        LcnnW& w = h->lcnn;
        w.c1w = L.p("layer1.0.filter.weight", 8 * 25); w.c1b = L.p("layer1.0.filter.bias", 8);
        w.a2w = L.p("layer2.0.conv_a.filter.weight", 8 * 4); w.a2b = L.p("layer2.0.conv_a.filter.bias", 8);
        w.k2w = L.p("layer2.0.conv.filter.weight", 16 * 4 * 9); w.k2b = L.p("layer2.0.conv.filter.bias", 16);
        w.a3w = L.p("layer3.0.conv_a.filter.weight", 16 * 8); w.a3b = L.p("layer3.0.conv_a.filter.bias", 16);
        w.k3w = L.p("layer3.0.conv.filter.weight", 32 * 8 * 9); w.k3b = L.p("layer3.0.conv.filter.bias", 32);
        w.f0w = L.p("fc0.0.filter.0.weight", 64 * 1024); w.f0b = L.p("fc0.0.filter.0.bias", 64);
        w.f1w = L.p("fc1.0.filter.0.weight", 64 * 32); w.f1b = L.p("fc1.0.filter.0.bias", 64);
        w.f2w = L.p("fc2.0.filter.0.weight", 16 * 32); w.f2b = L.p("fc2.0.filter.0.bias", 16);
        w.f3w = L.p("fc3.weight", 2 * 8); w.f3b = L.p("fc3.bias", 2);
        float *s2 = nullptr, *t2 = nullptr, *s3 = nullptr, *t3 = nullptr;
        { int r = fold_bn(h, L, "layer2.2.", 8, &s2, &t2); if (r) return r; }
        { int r = fold_bn(h, L, "layer3.2.", 16, &s3, &t3); if (r) return r; }
        w.bn2s = s2; w.bn2t = t2; w.bn3s = s3; w.bn3t = t3;
        if (!L.ok) return fail(h, OCC_ERR_WEIGHT, "occ_finalize_weights: LCNN weight " + L.missing);
        const int64_t B = h->max_batch, T = h->Tmax;
        DA(h->l_y1, B * 4 * (T / 2 + 1) * 512); DA(h->l_a2, B * 4 * (T / 2 + 1) * 512);
        DA(h->l_y2, B * 8 * (T / 4 + 1) * 256); DA(h->l_a3, B * 8 * (T / 4 + 1) * 256);
        DA(h->l_y3, B * 16 * (T / 8 + 1) * 128); DA(h->l_out, B * 2);
    }
    if (!h->has_backend) {
        CK(cudaDeviceSynchronize());
        for (auto it = h->raw.begin(); it != h->raw.end(); ++it)
            if (it->second.numel >= 100000 && it->second.d && it->first.rfind("ssl_model.", 0) == 0) { cudaFree(it->second.d); it->second.d = nullptr; }
        h->finalized = true;
        return OCC_OK;
    }
    {
        const float* w = K.p("LL.weight", 128 * 1024);
        h->bll = K.p("LL.bias", 128);
        DA(h->wll, 128LL * 1024 * 2);
        if (K.ok) LAUNCH(1, launch_pack_linear(w, 128, 1024, 0, 1.f, h->wll, 128LL * 1024, s));
    }
    float *b0s = nullptr, *b0t = nullptr;
    { int r = fold_bn(h, K, "first_bn.", 1, &b0s, &b0t); if (r) return r; }
    { int r = fold_bn(h, K, "first_bn1.", 64, &h->bn1s, &h->bn1t); if (r) return r; }
    auto pad64 = [&](const float* src, int n, float fill, float** dst) -> int {
        DA(*dst, 64);
        if (K.ok) LAUNCH(1, launch_pad_vec(src, n, fill, *dst, 64, s));
        return OCC_OK;
    };
    for (int n = 0; n < 6; ++n) {
        ResW& R = h->res[n];
        R.ci = kResFilts[n][0]; R.co = kResFilts[n][1];
        const std::string p = "encoder." + std::to_string(n) + ".0.";
        const float* w1 = K.p(p + "conv1.weight", (int64_t)R.co * R.ci * 6);
        const float* b1 = K.p(p + "conv1.bias", R.co);
        const float* w2 = K.p(p + "conv2.weight", (int64_t)R.co * R.co * 6);
        const float* b2 = K.p(p + "conv2.bias", R.co);
        float *sc = nullptr, *sh = nullptr;
        { int r = fold_bn(h, K, p + "bn2.", R.co, &sc, &sh); if (r) return r; }
        const float *wd = nullptr, *bd = nullptr;
        if (R.ci != R.co) {
            wd = K.p(p + "conv_downsample.weight", (int64_t)R.co * R.ci * 3);
            bd = K.p(p + "conv_downsample.bias", R.co);
        }
        if (!K.ok) continue;
        const int64_t pl2 = 64LL * 6 * 64, pl1 = 64LL * 3 * 64;
        DA(R.c2w, pl2 * 2);
        LAUNCH(1, launch_pack_conv2d_tc(w2, R.co, R.co, 6, R.c2w, pl2, s));
        { int r = pad64(b2, R.co, 0.f, &R.c2b); if (r) return r; }
        if (n == 0) {
            R.raw_c1w = w1; R.raw_c1b = b1; R.raw_dw = wd; R.raw_db = bd; R.raw_bn2s = sc; R.raw_bn2t = sh;
        } else {
            DA(R.c1w, pl2 * 2);
            LAUNCH(1, launch_pack_conv2d_tc(w1, R.co, R.ci, 6, R.c1w, pl2, s));
            { int r = pad64(b1, R.co, 0.f, &R.c1b); if (r) return r; }
            { int r = pad64(sc, R.co, 1.f, &R.bn2s); if (r) return r; }
            { int r = pad64(sh, R.co, 0.f, &R.bn2t); if (r) return r; }
            if (wd) {
                DA(R.dw, pl1 * 2);
                LAUNCH(1, launch_pack_conv2d_tc(wd, R.co, R.ci, 3, R.dw, pl1, s));
                { int r = pad64(bd, R.co, 0.f, &R.db); if (r) return r; }
            }
        }
    }
    h->attW1 = K.p("attention.0.weight", 128 * 64); h->attb1 = K.p("attention.0.bias", 128);
    { int r = fold_bn(h, K, "attention.2.", 128, &h->attbs, &h->attbt); if (r) return r; }
    {
        h->attW2 = K.p("attention.3.weight", 64 * 128);
        h->attb2 = K.p("attention.3.bias", 64);
    }
    h->posS = K.p("pos_S", 42 * 64);
    auto gat = [&](const std::string& p, GatW& g, float temp) -> int {
        std::memset(&g, 0, sizeof g);
        g.Wp = K.p(p + "att_proj.weight", 64 * 64); g.bp = K.p(p + "att_proj.bias", 64);
        g.aw11 = g.aw22 = g.aw12 = K.p(p + "att_weight", 64);
        g.Wa = K.p(p + "proj_with_att.weight", 64 * 64); g.ba = K.p(p + "proj_with_att.bias", 64);
        g.Wo = K.p(p + "proj_without_att.weight", 64 * 64); g.bo = K.p(p + "proj_without_att.bias", 64);
        float *sc, *sh;
        int r = fold_bn(h, K, p + "bn.", 64, &sc, &sh);
        g.bns = sc; g.bnt = sh; g.temp = temp;
        return r;
    };
    { int r = gat("GAT_layer_S.", h->gatS, 2.0f); if (r) return r; }
    { int r = gat("GAT_layer_T.", h->gatT, 2.0f); if (r) return r; }
    const char* hn[4] = {"HtrgGAT_layer_ST11.", "HtrgGAT_layer_ST12.", "HtrgGAT_layer_ST21.", "HtrgGAT_layer_ST22."};
    for (int i = 0; i < 4; ++i) {
        HtrgW& H = h->htrg[i];
        const int din = (i & 1) ? 32 : 64, dout = 32;
        const std::string p = hn[i];
        H.din = din;
        H.W1 = K.p(p + "proj_type1.weight", din * din); H.b1 = K.p(p + "proj_type1.bias", din);
        H.W2 = K.p(p + "proj_type2.weight", din * din); H.b2 = K.p(p + "proj_type2.bias", din);
        GatW& g = H.g;
        std::memset(&g, 0, sizeof g);
        g.Wp = K.p(p + "att_proj.weight", dout * din); g.bp = K.p(p + "att_proj.bias", dout);
        g.aw11 = K.p(p + "att_weight11", dout); g.aw22 = K.p(p + "att_weight22", dout); g.aw12 = K.p(p + "att_weight12", dout);
        g.Wa = K.p(p + "proj_with_att.weight", dout * din); g.ba = K.p(p + "proj_with_att.bias", dout);
        g.Wo = K.p(p + "proj_without_att.weight", dout * din); g.bo = K.p(p + "proj_without_att.bias", dout);
        g.WpM = K.p(p + "att_projM.weight", dout * din); g.bpM = K.p(p + "att_projM.bias", dout);
        g.awM = K.p(p + "att_weightM", dout);
        g.WaM = K.p(p + "proj_with_attM.weight", dout * din); g.baM = K.p(p + "proj_with_attM.bias", dout);
        g.WoM = K.p(p + "proj_without_attM.weight", dout * din); g.boM = K.p(p + "proj_without_attM.bias", dout);
        float *sc, *sh;
        { int r = fold_bn(h, K, p + "bn.", dout, &sc, &sh); if (r) return r; }
        g.bns = sc; g.bnt = sh; g.temp = 100.0f;
    }
    h->htrg[0].g.master_in = K.p("master1", 64);
    h->htrg[2].g.master_in = K.p("master2", 64);
    const char* pn[6] = {"pool_S", "pool_T", "pool_hS1", "pool_hT1", "pool_hS2", "pool_hT2"};
    for (int i = 0; i < 6; ++i) {
        const int D = i < 2 ? 64 : 32;
        h->poolw[i] = K.p(std::string(pn[i]) + ".proj.weight", D);
        h->poolb[i] = K.p(std::string(pn[i]) + ".proj.bias", 1);
    }
    h->outW = K.p("out_layer.weight", 2 * 160); h->outb = K.p("out_layer.bias", 2);
    if (!K.ok) return fail(h, OCC_ERR_WEIGHT, "occ_finalize_weights: backend weight " + K.missing);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(&h->bn0_scale, b0s, sizeof(float), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(&h->bn0_shift, b0t, sizeof(float), cudaMemcpyDeviceToHost));
    // the large fp32 originals are no longer needed (everything >= 100k elements was re-packed)
    for (auto it = h->raw.begin(); it != h->raw.end(); ++it) {
        // (only the frontend's and LL's: the LCNN / SE-ResNet34 kernels read their conv weights in the checkpoint layout)
        const bool repacked = it->first.rfind("ssl_model.", 0) == 0 || it->first == "LL.weight";
        if (it->second.numel >= 100000 && it->second.d && repacked) { cudaFree(it->second.d); it->second.d = nullptr; }
    }
    h->finalized = true;
    return OCC_OK;
}

// ================================================================================================ stages
namespace {

struct Ctx {
    int B, N, T, Wm;
    bool varlen;
    int L[7];
    const int *klen, *wlen, *kT1, *kT2;     // device arrays or nullptr (fixed length)
    int kT1c, kT2c, kT1max, kT2max;
};

int make_ctx(occ_handle* h, Ctx& c, const int32_t* lengths_host, int B, int N, cudaStream_t s, bool backend = true) {
    if (!h->finalized) return fail(h, OCC_ERR_STATE, "weights not finalized (call occ_finalize_weights)");
    if (backend && !h->has_backend)
        return fail(h, OCC_ERR_STATE, "frontend-only handle: no AASIST backend weights were loaded (use occ_extract_feat)");
    if (B < 1 || B > h->max_batch) return fail(h, OCC_ERR_ARG, "batch exceeds max_batch of the handle");
    if (N < 400 || N > h->max_samples) return fail(h, OCC_ERR_ARG, "n_samples outside [400, max_samples] of the handle");
    c.B = B; c.N = N;
    long long Lc = N;
    for (int i = 0; i < 7; ++i) { Lc = frames_layer(Lc, i); c.L[i] = (int)Lc; }
    c.T = c.L[6];
    c.Wm = c.T / 3;
    if (backend && c.Wm < 4) return fail(h, OCC_ERR_ARG, "utterance too short for the AASIST backend (need >= 12 frames)");
    if (c.T < 1) return fail(h, OCC_ERR_ARG, "utterance too short for the conv feature encoder");
    c.varlen = lengths_host != nullptr;
    c.kT1max = c.kT1c = std::max(c.Wm / 2, 1);
    c.kT2max = c.kT2c = std::max(c.kT1c / 2, 1);
    c.klen = c.wlen = c.kT1 = c.kT2 = nullptr;
    if (c.varlen) {
        h->hbuf.resize(4 * (size_t)B);
        for (int b = 0; b < B; ++b) {
            const int n = lengths_host[b];
            if (n < 400 || n > N) return fail(h, OCC_ERR_ARG, "lengths[b] outside [400, N]");
            const int t = frames_after(n, 7), w = t / 3;
            if (backend && w < 4) return fail(h, OCC_ERR_ARG, "utterance too short for the AASIST backend (need >= 12 frames)");
            h->hbuf[b] = t; h->hbuf[B + b] = w;
            h->hbuf[2 * B + b] = std::max(w / 2, 1);
            h->hbuf[3 * B + b] = std::max(std::max(w / 2, 1) / 2, 1);
        }
        CK(cudaMemcpyAsync(h->klen, h->hbuf.data(), sizeof(int) * B, cudaMemcpyHostToDevice, s));
        CK(cudaMemcpyAsync(h->wlen, h->hbuf.data() + B, sizeof(int) * B, cudaMemcpyHostToDevice, s));
        CK(cudaMemcpyAsync(h->kT1, h->hbuf.data() + 2 * B, sizeof(int) * B, cudaMemcpyHostToDevice, s));
        CK(cudaMemcpyAsync(h->kT2, h->hbuf.data() + 3 * B, sizeof(int) * B, cudaMemcpyHostToDevice, s));
        c.klen = h->klen; c.wlen = h->wlen; c.kT1 = h->kT1; c.kT2 = h->kT2;
    }
    return OCC_OK;
}

// conv feature encoder + feature LN + projection: wav [B,N] -> x32 fp32 [B*T,1024] and h16 bf16 (same values),
// rows beyond each utterance's valid frames zeroed (varlen).   components.py:53-182
int stage_conv_encoder(occ_handle* h, const Ctx& c, const float* wav, cudaStream_t s) {
    const int B = c.B, P = h->P;
    {
        ProfScope ps(h, OCC_PROF_CONV0, (double)B * c.N * 4.0 + (double)B * c.L[0] * 512.0 * 2.0 * P, s);      // memory family: algorithmic BYTES
        LAUNCH(1, launch_conv0(wav, B, c.N, c.L[0], h->c0w, h->c0b, h->c0g, h->c0e, h->actA, P == 2 ? h->actA_plane : 0, s, h->c0frag, h->num_sms));
    }
    bf16* in = h->actA; int64_t in_plane = h->actA_plane;
    bf16* out = h->actB; int64_t out_plane = h->actB_plane;
    for (int i = 1; i < 7; ++i) {
        const ConvW& w = h->conv[i];
        const int Lin = c.L[i - 1], Lout = c.L[i];
        GemmDesc d;
        d.A = in; d.a_cols = 512; d.a_sub = 2; d.a_sub_stride = 512; d.a_row_stride = 1024;
        d.a_rows = (Lin + 1) / 2; d.a_batch_stride = (int64_t)Lin * 512; d.nbatch = B; d.a_plane_stride = in_plane;
        d.W = w.w; d.N = 512; d.w_plane_stride = w.plane;
        d.nkb = w.k * 8; d.kb_per_tap = 8; d.ntaps_w = w.k; d.shift0 = 0; d.shift_h = 0; d.shift_w = 1;
        d.split = P == 2; d.rows_per_batch = Lout;
        d.bias = w.bias; d.BN = 256;
        const bool last = i == 6;
        // bf16 mode, layers 1-5: LayerNorm(512) + GELU run in the GEMM epilogue (the whole 512-wide row sits in the pair's two
        // accumulator stages; gemm_tc.cu epilogue_ln_tiles) and the layer output is written once, in bf16.  Layer 6 (201 frames:
        // 0.3 % of the conv work) also needs the feature LayerNorm on top and keeps the fp32 round trip, as does the fp32-accurate
        // mode (hi/lo output planes).
        static const int fused_env = [] { const char* e = getenv("OCC_CONV_LN_FUSED"); return e ? atoi(e) : 1; }();     // 0: previous form, for A/B runs
        if (P == 1 && !last && !h->simt && fused_env) {
            d.ln_gamma = w.g; d.ln_beta = w.b; d.out16 = out; d.ld16 = 512;
            ProfScope ps(h, OCC_PROF_CONV_GEMM, 2.0 * B * Lout * 512.0 * 512 * w.k, s);
            int r = run_gemm(h, d, s); if (r) return r;
            std::swap(in, out); std::swap(in_plane, out_plane);
            continue;
        }
        d.out32 = h->c32; d.ld32 = 512;
        {
            ProfScope ps(h, OCC_PROF_CONV_GEMM, 2.0 * B * Lout * 512.0 * 512 * w.k, s);
            int r = run_gemm(h, d, s); if (r) return r;
        }
        ProfScope ps(h, OCC_PROF_NORM, (double)B * Lout * 512.0 * (4.0 + 2.0 * P), s);                           // bytes: fp32 in, bf16 (hi+lo) out
        // per-frame LayerNorm over 512 channels (fp32) + erf-GELU; the last layer also applies the feature LayerNorm
        LAUNCH(1, launch_rows_ln(h->c32, (int64_t)B * Lout, 512, true, w.g, w.b, last ? h->flng : nullptr, last ? h->flnb : nullptr,
                                 nullptr, out, P == 2 ? out_plane : 0, s));
        std::swap(in, out); std::swap(in_plane, out_plane);
    }
    // `in` now holds LN(features) bf16 [B*T,512]
    const int M = B * c.T;
    GemmDesc d;
    d.A = in; d.a_cols = 512; d.a_sub = 1; d.a_row_stride = 512; d.a_rows = c.T; d.a_batch_stride = (int64_t)c.T * 512;
    d.nbatch = B; d.a_plane_stride = in_plane;
    d.W = h->wproj; d.N = 1024; d.w_plane_stride = 1024LL * 512; d.nkb = 8; d.kb_per_tap = 8; d.ntaps_w = 1;
    d.split = P == 2; d.rows_per_batch = c.T; d.bias = h->bproj;
    d.out32 = h->x32; d.ld32 = 1024; d.out16 = h->h16; d.ld16 = 1024; d.out16_plane_stride = P == 2 ? h->pl1024 : 0;
    d.row_valid = c.klen; d.BN = 256;
    ProfScope ps(h, OCC_PROF_PROJ_GEMM, 2.0 * M * 512.0 * 1024, s);
    return run_gemm(h, d, s);
}

// y = GELU(pos_conv(h16) + bias)  (components.py:185-234); the sum x + y of components.py:422 is formed by the LayerNorm
// kernel that consumes it (stage_transformer, pending_y) or by stage_add_pending.  In: h16 = bf16 copy of x.
int stage_pos_conv(occ_handle* h, const Ctx& c, cudaStream_t s) {
    if (h->wpos_toep != nullptr) {
        // Toeplitz form (bf16 mode): four adjacent output frames share one N = 256 tile, so the 128-tap, 64-channel groups run on
        // full-width tcgen05.mma instead of N = 64 ones that sit on the per-instruction floor.  The GEMM walks the FLAT
        // frame sequence with one M row per 4 frames; utterances are copied into a padded layout (>= 64 zero frames between
        // them = the conv's own zero padding) so that no tap reaches into a neighbour.
        const int T = c.T, Tp = ((T + 64 + 3) / 4) * 4;
        ProfScope ps(h, OCC_PROF_POS_GEMM, 2.0 * c.B * c.T * 1024.0 * 64 * 128, s);       // includes the padded-layout copy
        CK(cudaMemset2DAsync(h->xpad16 + (int64_t)T * 1024, (size_t)Tp * 2048, 0, (size_t)(Tp - T) * 2048, (size_t)c.B, s));
        CK(cudaMemcpy2DAsync(h->xpad16, (size_t)Tp * 2048, h->h16, (size_t)T * 2048, (size_t)T * 2048, (size_t)c.B, cudaMemcpyDeviceToDevice, s));
        GemmDesc d;
        const int rows = c.B * Tp / 4;
        d.A = h->xpad16; d.a_cols = 1024; d.a_sub = 4; d.a_sub_stride = 1024; d.a_row_stride = 4096; d.a_rows = rows;
        d.a_batch_stride = 0; d.nbatch = 1; d.a_plane_stride = 0;
        d.W = h->wpos_toep; d.N = 4096; d.w_plane_stride = 0;
        d.nkb = 131; d.kb_per_tap = 1; d.ntaps_w = 131; d.shift0 = -64; d.shift_h = 0; d.shift_w = 1; d.a_col_group_stride = 64;
        d.split = 0; d.rows_per_batch = rows;
        d.bias = h->bpos; d.gelu = 1; d.BN = 256;
        d.out16 = h->y16; d.ld16 = 1024;
        d.toep_T = T; d.toep_Tp = Tp; d.toep_B = c.B;
        return run_gemm(h, d, s);
    }
    GemmDesc d;
    d.A = h->h16; d.a_cols = 1024; d.a_sub = 1; d.a_row_stride = 1024; d.a_rows = c.T; d.a_batch_stride = (int64_t)c.T * 1024;
    d.nbatch = c.B; d.a_plane_stride = h->pl1024;
    d.W = h->wpos; d.N = 1024; d.w_plane_stride = 1024LL * 8192;
    d.nkb = 128; d.kb_per_tap = 1; d.ntaps_w = 128; d.shift0 = -64; d.shift_h = 0; d.shift_w = 1; d.a_col_group_stride = 64;
    d.split = h->P == 2; d.rows_per_batch = c.T;
    d.bias = h->bpos; d.gelu = 1; d.BN = 64;
    if (h->P == 2) { d.out32 = h->y32; d.ld32 = 1024; } else { d.out16 = h->y16; d.ld16 = 1024; }
    ProfScope ps(h, OCC_PROF_POS_GEMM, 2.0 * c.B * c.T * 1024.0 * 64 * 128, s);
    return run_gemm(h, d, s);
}

// x32 += y (the pending sub-block output), no normalisation
int stage_add_pending(occ_handle* h, const Ctx& c, cudaStream_t s) {
    LAUNCH(1, launch_rows_ln(h->x32, (int64_t)c.B * c.T, 1024, false, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, s,
                             h->y16, h->y32, h->x32));
    return OCC_OK;
}

// n_layers pre-LN encoder layers on x32 in place (components.py:345-401), then the final LayerNorm
// (components.py:441-442) into feat32 / feat16 (hi+lo planes).  hidden: optional fp32 [n_layers, B*T, 1024].
// Residual adds sit in the out-proj / FC2 epilogues: x <- x + (A W^T + b) in place on the fp32 residual stream, the residual
// prefetched one chunk ahead (gemm_tc.cu, RESID_PF), so the LayerNorm kernels only read x and write the bf16 operand.  Only the
// positional-conv term of layer 0 is still added by the LayerNorm kernel that consumes it (pending_y).
// pending_y: y already holds the positional-conv term on entry.
int stage_transformer(occ_handle* h, const Ctx& c, int n_layers, float* hidden, bool pending_y, cudaStream_t s) {
    const int M = c.B * c.T, P = h->P, split = P == 2;
    const int64_t p1 = split ? h->pl1024 : 0, p4 = split ? h->pl4096 : 0;
    const int64_t sq = 1024LL * 1024;
    const bf16* add16 = pending_y ? h->y16 : nullptr;
    const float* add32 = pending_y ? h->y32 : nullptr;
    // Ragged batch: attention writes ctx only for rows inside each utterance, so rows beyond it would otherwise keep whatever an
    // earlier forward left there (possibly NaN/Inf from a corrupt utterance) and reach later layers' padded K/V rows through
    // out-proj (0 x NaN = NaN in the PV MMA).  Zeroed once per forward: padded rows are then a function of this batch only.
    if (c.klen != nullptr) CK(cudaMemsetAsync(h->ctx16, 0, sizeof(bf16) * (size_t)(h->pl1024 * P), s));
    // out-proj / FC2 add their result into the fp32 residual stream in their epilogue (x32 in place, residual prefetched)
    // LayerNorm fold (bf16 mode): out-proj / FC2 also leave bf16(x) in h16 and the row partial sums in lnstat (fold = true), and
    // the next QKV / FC1 read them instead of a LayerNorm kernel's output -- 47 of the 49 LayerNorm launches of a forward vanish.
    auto sub_out = [&](GemmDesc& d, bool fold) {
        d.resid = h->x32; d.ld_resid = 1024; d.out32 = h->x32; d.ld32 = 1024;
        if (fold) { d.ln_x16_out = h->h16; d.ln_stat_out = h->lnstat; }
    };
    bool stats_ready = false;          // h16 / lnstat describe the current residual stream
    for (int l = 0; l < n_layers; ++l) {
        const LayerW& w = h->layer[l];
        const double fl = 2.0 * M * 1024.0 * 1024.0;
        const bool fold_qkv = w.sqkv != nullptr && stats_ready;
        if (!fold_qkv) {
            ProfScope ps(h, OCC_PROF_NORM, (double)M * 1024.0 * (4.0 + 2.0 * P + (l == 0 && pending_y ? 4.0 + 2.0 * (split ? 2 : 1) : 0.0)), s);
            LAUNCH(1, launch_rows_ln(h->x32, M, 1024, false, w.ln1g, w.ln1b, nullptr, nullptr, nullptr, h->h16, p1, s, add16, add32, h->x32));
        }
        if (w.sqkv != nullptr && !stats_ready) return fail(h, OCC_ERR_STATE, "internal: folded QKV weights without row statistics");
        {
            GemmDesc d = lin_desc(h->h16, h->pl1024, M, 1024, w.wqkv, 3 * sq, 3072, split);
            d.bias = w.bqkv;
            if (fold_qkv) { d.bias = w.cqkv; d.ln_stat_in = h->lnstat; d.ln_colsum = w.sqkv; }
            if (split) { d.out32 = h->qkv32; d.ld32 = 3072; } else { d.out16 = h->qkv16; d.ld16 = 3072; }
            ProfScope ps(h, OCC_PROF_XF_QKV, 3 * fl, s);
            int r = run_gemm(h, d, s); if (r) return r;
        }
        {
            ProfScope ps(h, OCC_PROF_ATTN, 4.0 * c.B * c.T * (double)c.T * 1024.0, s);
            if (split) LAUNCH(1, launch_attn_f32(h->qkv32, c.B, c.T, c.klen, h->ctx16, p1, s));
            else LAUNCH(1, launch_attn_bf16(h->qkv16, c.B, c.T, c.klen, h->ctx16, h->num_sms, s));
        }
        {
            GemmDesc d = lin_desc(h->ctx16, h->pl1024, M, 1024, w.wo, sq, 1024, split);
            const bool fold_fc1 = w.s1 != nullptr;
            d.bias = w.bo; sub_out(d, fold_fc1);
            ProfScope ps(h, OCC_PROF_XF_OUT, fl, s);
            int r = run_gemm(h, d, s); if (r) return r;
        }
        if (w.s1 == nullptr) {
            ProfScope ps(h, OCC_PROF_NORM, (double)M * 1024.0 * (4.0 + 2.0 * P), s);
            LAUNCH(1, launch_rows_ln(h->x32, M, 1024, false, w.ln2g, w.ln2b, nullptr, nullptr, nullptr, h->h16, p1, s));
        }
        {
            GemmDesc d = lin_desc(h->h16, h->pl1024, M, 1024, w.w1, 4 * sq, 4096, split);
            d.bias = w.b1; d.gelu = 1; d.out16 = h->f16; d.ld16 = 4096; d.out16_plane_stride = p4;
            if (w.s1 != nullptr) { d.bias = w.c1; d.ln_stat_in = h->lnstat; d.ln_colsum = w.s1; }
            ProfScope ps(h, OCC_PROF_XF_FC1, 4 * fl, s);
            int r = run_gemm(h, d, s); if (r) return r;
        }
        {
            GemmDesc d = lin_desc(h->f16, h->pl4096, M, 4096, w.w2, 4 * sq, 1024, split);
            const bool fold_next = l + 1 < n_layers && l + 1 < 24 && h->layer[l + 1].sqkv != nullptr;
            d.bias = w.b2; sub_out(d, fold_next);
            stats_ready = fold_next;
            // layer output (pre final LN) = the residual stream after this add: the hidden-state export is a second store of the same
            // epilogue registers (fairseq layer_results / torchaudio extract_features, components.py:445-463)
            if (hidden) d.out32_copy = hidden + (int64_t)l * M * 1024;
            ProfScope ps(h, OCC_PROF_XF_FC2, 4 * fl, s);
            int r = run_gemm(h, d, s); if (r) return r;
        }
        add16 = nullptr; add32 = nullptr;
    }
    {
        ProfScope ps(h, OCC_PROF_NORM, (double)M * 1024.0 * (4.0 + 4.0 + 4.0), s);                               // fp32 in, fp32 + bf16 hi/lo out
        LAUNCH(1, launch_rows_ln(h->x32, M, 1024, false, h->elng, h->elnb, nullptr, nullptr, h->feat32, h->feat16, h->pl1024, s, add16, add32, h->x32));
    }
    return OCC_OK;
}

// AASIST backend on feat16 (hi/lo planes of the frontend output): models/sslassist.py:509-597
int stage_backend(occ_handle* h, const Ctx& c, float* emb, float* out, cudaStream_t s) {
    const int B = c.B, T = c.T, Wm = c.Wm, M = B * T;
    ProfScope ps_backend(h, OCC_PROF_BACKEND, 1.42e9 * B * (T / 201.0), s);
    {   // LL 1024 -> 128, fp32-accurate split GEMM in both modes (top-k pooling downstream is precision sensitive)
        GemmDesc d = lin_desc(h->feat16, h->pl1024, M, 1024, h->wll, 128LL * 1024, 128, 1);
        d.bias = h->bll; d.out32 = h->ll32; d.ld32 = 128;
        int r = run_gemm(h, d, s); if (r) return r;
    }
    LAUNCH(1, launch_ll_pool(h->ll32, B, T, Wm, c.wlen, h->bn0_scale, h->bn0_shift, h->bx0, s));
    // ---- residual encoder (sslassist.py:373-429, :518): tcgen05 shift-GEMMs over channel-last zero-padded activations.
    // Padded grid: kHp rows x Wp = Wm + 2 columns; activation (h, w) lives at (h + 1, w + 1); position index t = hp*Wp + wp.
    const int Wp = Wm + 2;
    const int64_t grid_rows = (int64_t)kHp * Wp;
    auto conv_desc = [&](const bf16* in, const bf16* w, int taps, int shift0, const float* bias, int h1, int cin) {
        GemmDesc d;
        d.k_real = cin;
        d.A = in; d.a_cols = 64; d.a_sub = 1; d.a_row_stride = 64; d.a_rows = (int)grid_rows; d.a_batch_stride = grid_rows * 64;
        d.nbatch = B; d.a_plane_stride = h->cplane;
        d.W = w; d.N = 64; d.w_plane_stride = 64LL * taps * 64;
        d.nkb = taps; d.kb_per_tap = 1; d.ntaps_w = 3; d.shift0 = shift0; d.shift_h = Wp; d.shift_w = 1;
        d.split = 1; d.rows_per_batch = (int)grid_rows; d.bias = bias; d.BN = 64;
        d.conv_mode = 1; d.conv_wp = Wp; d.conv_h0 = 1; d.conv_h1 = h1; d.conv_wm = Wm; d.conv_wlen = c.wlen;
        d.ld16 = 64; d.ld32 = 64; d.out16_plane_stride = h->cplane;
        return d;
    };
    {
        const ResW& R0 = h->res[0];
        LAUNCH(1, launch_res0_front(h->bx0, B, Wm, c.wlen, R0.raw_c1w, R0.raw_c1b, R0.raw_bn2s, R0.raw_bn2t, R0.raw_dw, R0.raw_db,
                                    kHp, h->cA, h->cI, h->cplane, s));
    }
    const bf16* xin = nullptr;
    int cur = 0;
    for (int n = 0; n < 6; ++n) {
        const ResW& R = h->res[n];
        const bf16* idn = xin;
        if (n > 0) {
            {   // conv1 (2,3) pad (1,1) on the raw block input, + bn2 + SELU: 43 rows
                GemmDesc d = conv_desc(xin, R.c1w, 6, -Wp - 1, R.c1b, 44, R.ci);
                d.cs1 = R.bn2s; d.ct1 = R.bn2t; d.out16 = h->cA;
                int r = run_gemm(h, d, s); if (r) return r;
            }
            if (R.dw) {   // conv_downsample (1,3) pad (0,1)
                GemmDesc d = conv_desc(xin, R.dw, 3, -1, R.db, 43, R.ci);
                d.out16 = h->cI;
                int r = run_gemm(h, d, s); if (r) return r;
                idn = h->cI;
            }
        } else {
            idn = h->cI;
        }
        {   // conv2 (2,3) pad (0,1) + identity; the last block also applies first_bn1 + SELU (sslassist.py:520-521)
            GemmDesc d = conv_desc(h->cA, R.c2w, 6, -1, R.c2b, 43, R.co);
            d.cresid = idn; d.cresid_plane = h->cplane;
            if (n == 5) { d.cs2 = h->bn1s; d.ct2 = h->bn1t; d.out32 = h->xf32; d.out16_plane_stride = 0; }
            else d.out16 = h->cX[cur];
            int r = run_gemm(h, d, s); if (r) return r;
        }
        xin = h->cX[cur];
        cur ^= 1;
    }
    XView x{h->xf32 + ((int64_t)Wp + 1) * 64, grid_rows * 64, 1, (int64_t)Wp * 64, 64};
    LAUNCH(1, launch_att_map(x, B, Wm, h->attW1, h->attb1, h->attbs, h->attbt, h->attW2, h->attb2, h->wmap, h->num_sms, s));
    LAUNCH(2, launch_st_pool(x, h->wmap, B, Wm, c.wlen, h->posS, h->eS, h->eT, s));
    {   // GAT_layer_S / GAT_layer_T (one launch)
        GatArgs a;
        std::memset(&a, 0, sizeof a);
        a.w[0] = h->gatS; a.w[1] = h->gatT;
        a.x[0] = h->eS; a.x[1] = h->eT; a.y[0] = h->yS; a.y[1] = h->yT;
        a.n1[0] = nullptr; a.n1c[0] = 42; a.n1[1] = c.wlen; a.n1c[1] = Wm;
        a.n2[0] = a.n2[1] = nullptr; a.n2c[0] = a.n2c[1] = 0;
        a.nmax[0] = 42; a.nmax[1] = Wm; a.has_master = 0;
        LAUNCH(1, launch_gat(a, B, 2, 64, 64, s));
    }
    {   // pool_S, pool_T
        PoolArgs p;
        std::memset(&p, 0, sizeof p);
        p.h[0] = h->yS; p.h_bstride[0] = 42 * 64; p.n[0] = nullptr; p.nc[0] = 42; p.w[0] = h->poolw[0]; p.bias[0] = h->poolb[0];
        p.out[0] = h->oS; p.kmax[0] = 21;
        p.h[1] = h->yT; p.h_bstride[1] = (long long)Wm * 64; p.n[1] = c.wlen; p.nc[1] = Wm; p.w[1] = h->poolw[1]; p.bias[1] = h->poolb[1];
        p.out[1] = h->oT; p.kmax[1] = c.kT1max;
        LAUNCH(1, launch_graph_pool(p, B, 2, 64, std::max(42, Wm), s));
    }
    const int nm1 = c.kT1max + 21, nm2 = c.kT2max + 10;
    {   // HS-GAL 1 of both branches: ST11, ST21
        HprojArgs a;
        std::memset(&a, 0, sizeof a);
        GatArgs g;
        std::memset(&g, 0, sizeof g);
        for (int z = 0; z < 2; ++z) {
            const HtrgW& H = h->htrg[2 * z];
            a.x1[z] = h->oT; a.x2[z] = h->oS; a.n1max[z] = c.kT1max; a.n2max[z] = 21;
            a.n1[z] = c.kT1; a.n1c[z] = c.kT1c; a.n2[z] = nullptr; a.n2c[z] = 21;
            a.W1[z] = H.W1; a.b1[z] = H.b1; a.W2[z] = H.W2; a.b2[z] = H.b2; a.xc[z] = h->xc1[z];
            g.w[z] = H.g; g.w[z].master_stride = 0; g.w[z].master_out = h->m1[z];
            g.x[z] = h->xc1[z]; g.y[z] = h->y1[z];
            g.n1[z] = c.kT1; g.n1c[z] = c.kT1c; g.n2[z] = nullptr; g.n2c[z] = 21; g.nmax[z] = nm1;
        }
        g.has_master = 1;
        LAUNCH(1, launch_hproj(a, B, 2, 64, s));
        LAUNCH(1, launch_gat(g, B, 2, 64, 32, s));
    }
    {   // pool_hS1, pool_hT1, pool_hS2, pool_hT2
        PoolArgs p;
        std::memset(&p, 0, sizeof p);
        for (int z = 0; z < 2; ++z) {
            const int iS = 2 * z, iT = 2 * z + 1;
            p.h[iS] = h->y1[z]; p.h_bstride[iS] = (long long)nm1 * 32; p.n[iS] = nullptr; p.nc[iS] = 21;
            p.noff[iS] = c.kT1; p.noffc[iS] = c.kT1c; p.w[iS] = h->poolw[2 + 2 * z]; p.bias[iS] = h->poolb[2 + 2 * z];
            p.out[iS] = h->S1[z]; p.kmax[iS] = 10;
            p.h[iT] = h->y1[z]; p.h_bstride[iT] = (long long)nm1 * 32; p.n[iT] = c.kT1; p.nc[iT] = c.kT1c;
            p.noff[iT] = nullptr; p.noffc[iT] = 0; p.w[iT] = h->poolw[3 + 2 * z]; p.bias[iT] = h->poolb[3 + 2 * z];
            p.out[iT] = h->T1[z]; p.kmax[iT] = c.kT2max;
        }
        LAUNCH(1, launch_graph_pool(p, B, 4, 32, std::max(21, c.kT1max), s));
    }
    {   // HS-GAL 2: ST12, ST22 (master = the first layer's updated master)
        HprojArgs a;
        std::memset(&a, 0, sizeof a);
        GatArgs g;
        std::memset(&g, 0, sizeof g);
        for (int z = 0; z < 2; ++z) {
            const HtrgW& H = h->htrg[2 * z + 1];
            a.x1[z] = h->T1[z]; a.x2[z] = h->S1[z]; a.n1max[z] = c.kT2max; a.n2max[z] = 10;
            a.n1[z] = c.kT2; a.n1c[z] = c.kT2c; a.n2[z] = nullptr; a.n2c[z] = 10;
            a.W1[z] = H.W1; a.b1[z] = H.b1; a.W2[z] = H.W2; a.b2[z] = H.b2; a.xc[z] = h->xc2[z];
            g.w[z] = H.g; g.w[z].master_in = h->m1[z]; g.w[z].master_stride = 32; g.w[z].master_out = h->maug[z];
            g.x[z] = h->xc2[z]; g.y[z] = h->y2[z];
            g.n1[z] = c.kT2; g.n1c[z] = c.kT2c; g.n2[z] = nullptr; g.n2c[z] = 10; g.nmax[z] = nm2;
        }
        g.has_master = 1;
        LAUNCH(1, launch_hproj(a, B, 2, 32, s));
        LAUNCH(1, launch_gat(g, B, 2, 32, 32, s));
    }
    {
        ReadoutArgs r;
        std::memset(&r, 0, sizeof r);
        for (int z = 0; z < 2; ++z) { r.T1[z] = h->T1[z]; r.S1[z] = h->S1[z]; r.y2[z] = h->y2[z]; r.m[z] = h->m1[z]; r.maug[z] = h->maug[z]; }
        r.kTmax = c.kT2max; r.kSmax = 10; r.kT = c.kT2; r.kS = nullptr; r.kTc = c.kT2c; r.kSc = 10;
        r.Wout = h->outW; r.bout = h->outb; r.emb = emb ? emb : h->emb_dev; r.out = out ? out : h->out_dev;
        LAUNCH(1, launch_readout(r, B, s));
    }
    return OCC_OK;
}

}  // namespace

// ================================================================================================ forward API
extern "C" int occ_extract_feat(occ_handle* h, const float* wav_dev, const int32_t* lengths_host, int B, int N,
                                float* feat_dev, float* hidden_dev, void* stream) {
    if (!h || !wav_dev) return fail(h, OCC_ERR_ARG, "occ_extract_feat: null argument");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(h->device));
    Ctx c;
    int r = make_ctx(h, c, lengths_host, B, N, s, false); if (r) return r;
    r = stage_conv_encoder(h, c, wav_dev, s); if (r) return r;
    r = stage_pos_conv(h, c, s); if (r) return r;
    r = stage_transformer(h, c, 24, hidden_dev, true, s); if (r) return r;
    if (feat_dev) CK(cudaMemcpyAsync(feat_dev, h->feat32, sizeof(float) * (size_t)B * c.T * 1024, cudaMemcpyDeviceToDevice, s));
    return OCC_OK;
}

extern "C" int occ_forward(occ_handle* h, const float* wav_dev, const int32_t* lengths_host, int B, int N,
                           float* emb_dev, float* out_dev, float* hidden_dev, void* stream) {
    if (!h || !wav_dev) return fail(h, OCC_ERR_ARG, "occ_forward: null argument");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(h->device));
    Ctx c;
    int r = make_ctx(h, c, lengths_host, B, N, s); if (r) return r;
    r = stage_conv_encoder(h, c, wav_dev, s); if (r) return r;
    r = stage_pos_conv(h, c, s); if (r) return r;
    r = stage_transformer(h, c, 24, hidden_dev, true, s); if (r) return r;
    return stage_backend(h, c, emb_dev, out_dev, s);
}

// ssl_lcnn.forward (models/lcnn.py:242-260): waveform -> XLS-R features -> LCNN -> logits [B,2]
extern "C" int occ_forward_lcnn(occ_handle* h, const float* wav_dev, int B, int N, float* out_dev, void* stream) {
    if (!h || !wav_dev || !out_dev) return fail(h, OCC_ERR_ARG, "occ_forward_lcnn: null argument");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(h->device));
    if (h->finalized && !h->has_lcnn) return fail(h, OCC_ERR_STATE, "no LCNN weights (lcnn.*) were loaded into this handle");
    Ctx c;
    int r = make_ctx(h, c, nullptr, B, N, s, false); if (r) return r;
    if (c.T < 8) return fail(h, OCC_ERR_ARG, "utterance too short for the LCNN backend (need >= 8 frames)");
    r = stage_conv_encoder(h, c, wav_dev, s); if (r) return r;
    r = stage_pos_conv(h, c, s); if (r) return r;
    r = stage_transformer(h, c, 24, nullptr, true, s); if (r) return r;
    LAUNCH(6, launch_lcnn(h->lcnn, h->feat32, B, c.T, h->l_y1, h->l_a2, h->l_y2, h->l_a3, h->l_y3, out_dev, s));
    return OCC_OK;
}

// ssl_resnet34.forward (models/senet.py:168-185): waveform -> XLS-R features -> SE-ResNet34 -> (com [B,128], des [B,2])
static int run_resnet34(occ_handle* h, const float* feat, int B, int T, float* com, float* des, cudaStream_t s) {
    int n = 0;
    cudaError_t e = launch_resnet34(h->resnet, feat, B, T, h->r_buf[0], h->r_buf[1], h->r_buf[2], h->r_buf[3], h->r_buf[4], h->r_vec, com, des, s, &n, h->num_sms);
    if (e != cudaSuccess) return cuda_fail(h, e, "launch_resnet34");
    h->launches += n;
    return OCC_OK;
}
extern "C" int occ_forward_resnet34(occ_handle* h, const float* wav_dev, int B, int N, float* com_dev, float* des_dev, void* stream) {
    if (!h || !wav_dev || !com_dev || !des_dev) return fail(h, OCC_ERR_ARG, "occ_forward_resnet34: null argument");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(h->device));
    if (h->finalized && !h->has_resnet) return fail(h, OCC_ERR_STATE, "no SE-ResNet34 weights (resnet34.*) were loaded into this handle");
    Ctx c;
    int r = make_ctx(h, c, nullptr, B, N, s, false); if (r) return r;
    r = stage_conv_encoder(h, c, wav_dev, s); if (r) return r;
    r = stage_pos_conv(h, c, s); if (r) return r;
    r = stage_transformer(h, c, 24, nullptr, true, s); if (r) return r;
    return run_resnet34(h, h->feat32, B, c.T, com_dev, des_dev, s);
}
// ResNet.forward (models/senet.py:121-143) alone: feat_dev fp32 [B,T,1024] -> com_dev [B,128], des_dev [B,2]
extern "C" int occ_resnet34_backend(occ_handle* h, const float* feat_dev, int B, int T, float* com_dev, float* des_dev, void* stream) {
    if (!h || !feat_dev || !com_dev || !des_dev) return fail(h, OCC_ERR_ARG, "occ_resnet34_backend: null argument");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(h->device));
    if (!h->finalized) return fail(h, OCC_ERR_STATE, "weights not finalized (call occ_finalize_weights)");
    if (!h->has_resnet) return fail(h, OCC_ERR_STATE, "no SE-ResNet34 weights (resnet34.*) were loaded into this handle");
    if (B < 1 || B > h->max_batch || T < 1 || T > h->Tmax) return fail(h, OCC_ERR_ARG, "B / T outside the handle's workspace");
    return run_resnet34(h, feat_dev, B, T, com_dev, des_dev, s);
}

// LCNN.forward (models/lcnn.py:189-213) alone: feat_dev fp32 [B,T,1024] -> out_dev [B,2]
extern "C" int occ_lcnn_backend(occ_handle* h, const float* feat_dev, int B, int T, float* out_dev, void* stream) {
    if (!h || !feat_dev || !out_dev) return fail(h, OCC_ERR_ARG, "occ_lcnn_backend: null argument");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(h->device));
    if (!h->finalized) return fail(h, OCC_ERR_STATE, "weights not finalized (call occ_finalize_weights)");
    if (!h->has_lcnn) return fail(h, OCC_ERR_STATE, "no LCNN weights (lcnn.*) were loaded into this handle");
    if (B < 1 || B > h->max_batch || T < 8 || T > h->Tmax) return fail(h, OCC_ERR_ARG, "B / T outside the handle's workspace (LCNN needs >= 8 frames)");
    LAUNCH(6, launch_lcnn(h->lcnn, feat_dev, B, T, h->l_y1, h->l_a2, h->l_y2, h->l_a3, h->l_y3, out_dev, s));
    return OCC_OK;
}

extern "C" int occ_conv_encoder(occ_handle* h, const float* wav_dev, int B, int N, float* proj_dev, void* stream) {
    if (!h || !wav_dev || !proj_dev) return fail(h, OCC_ERR_ARG, "occ_conv_encoder: null argument");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(h->device));
    Ctx c;
    int r = make_ctx(h, c, nullptr, B, N, s, false); if (r) return r;
    r = stage_conv_encoder(h, c, wav_dev, s); if (r) return r;
    CK(cudaMemcpyAsync(proj_dev, h->x32, sizeof(float) * (size_t)B * c.T * 1024, cudaMemcpyDeviceToDevice, s));
    return OCC_OK;
}

// Stage entry points take fp32 [B,T,1024] tensors; T must equal occ_num_frames(N) for some N <= max_samples only in the
// sense that B*T fits the workspace.
static int ctx_from_T(occ_handle* h, Ctx& c, int B, int T) {
    if (!h->finalized) return fail(h, OCC_ERR_STATE, "weights not finalized (call occ_finalize_weights)");
    if (B < 1 || B > h->max_batch || T < 12 || T > h->Tmax) return fail(h, OCC_ERR_ARG, "B / T outside the handle's workspace");
    std::memset(&c, 0, sizeof c);
    c.B = B; c.T = T; c.Wm = T / 3; c.N = 0; c.varlen = false;
    c.kT1max = c.kT1c = std::max(c.Wm / 2, 1);
    c.kT2max = c.kT2c = std::max(c.kT1c / 2, 1);
    return OCC_OK;
}

extern "C" int occ_pos_conv(occ_handle* h, const float* x_dev, int B, int T, float* y_dev, void* stream) {
    if (!h || !x_dev || !y_dev) return fail(h, OCC_ERR_ARG, "occ_pos_conv: null argument");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(h->device));
    Ctx c;
    int r = ctx_from_T(h, c, B, T); if (r) return r;
    const int64_t n = (int64_t)B * T * 1024;
    CK(cudaMemcpyAsync(h->x32, x_dev, sizeof(float) * n, cudaMemcpyDeviceToDevice, s));
    LAUNCH(1, launch_f32_to_bf16(x_dev, n, h->h16, h->P == 2 ? h->pl1024 : 0, s));
    r = stage_pos_conv(h, c, s); if (r) return r;
    r = stage_add_pending(h, c, s); if (r) return r;
    CK(cudaMemcpyAsync(y_dev, h->x32, sizeof(float) * n, cudaMemcpyDeviceToDevice, s));
    return OCC_OK;
}

extern "C" int occ_transformer(occ_handle* h, const float* x_dev, int B, int T, int n_layers, float* y_dev, float* hidden_dev, void* stream) {
    if (!h || !x_dev) return fail(h, OCC_ERR_ARG, "occ_transformer: null argument");
    if (n_layers < 0 || n_layers > 24) return fail(h, OCC_ERR_ARG, "occ_transformer: n_layers outside [0,24]");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(h->device));
    Ctx c;
    int r = ctx_from_T(h, c, B, T); if (r) return r;
    const int64_t n = (int64_t)B * T * 1024;
    CK(cudaMemcpyAsync(h->x32, x_dev, sizeof(float) * n, cudaMemcpyDeviceToDevice, s));
    r = stage_transformer(h, c, n_layers, hidden_dev, false, s); if (r) return r;
    if (y_dev) CK(cudaMemcpyAsync(y_dev, h->feat32, sizeof(float) * n, cudaMemcpyDeviceToDevice, s));
    return OCC_OK;
}

extern "C" int occ_aasist_backend(occ_handle* h, const float* feat_dev, int B, int T, float* emb_dev, float* out_dev, void* stream) {
    if (!h || !feat_dev) return fail(h, OCC_ERR_ARG, "occ_aasist_backend: null argument");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(h->device));
    Ctx c;
    int r = ctx_from_T(h, c, B, T); if (r) return r;
    if (!h->has_backend) return fail(h, OCC_ERR_STATE, "frontend-only handle: no AASIST backend weights were loaded");
    LAUNCH(1, launch_f32_to_bf16(feat_dev, (int64_t)B * T * 1024, h->feat16, h->pl1024, s));
    return stage_backend(h, c, emb_dev, out_dev, s);
}

// ================================================================================================ scores / EER
extern "C" int occ_score_l2(const float* emb_dev, const float* ref_dev, int B, int D, float eps, float* score_dev, void* stream) {
    occ_handle* h = nullptr;
    if (!emb_dev || !ref_dev || !score_dev || B < 0 || D < 1) return fail(h, OCC_ERR_ARG, "occ_score_l2: bad argument");
    CK(launch_score_l2(emb_dev, ref_dev, B, D, eps, score_dev, reinterpret_cast<cudaStream_t>(stream)));
    return OCC_OK;
}
extern "C" int occ_score_cos(const float* emb_dev, const float* center_dev, int B, int D, float* score_dev, void* stream) {
    occ_handle* h = nullptr;
    if (!emb_dev || !center_dev || !score_dev || B < 0 || D < 1) return fail(h, OCC_ERR_ARG, "occ_score_cos: bad argument");
    CK(launch_score_cos(emb_dev, center_dev, B, D, score_dev, reinterpret_cast<cudaStream_t>(stream)));
    return OCC_OK;
}
extern "C" int occ_ref_embedding(const float* emb_dev, int n, int D, float eps, float* ref_dev, float* thr_dev, void* stream) {
    occ_handle* h = nullptr;
    if (!emb_dev || !ref_dev || !thr_dev || n < 1 || D < 1 || D > 4096) return fail(h, OCC_ERR_ARG, "occ_ref_embedding: bad argument");
    CK(launch_ref_embedding(emb_dev, n, D, eps, ref_dev, thr_dev, reinterpret_cast<cudaStream_t>(stream)));
    return OCC_OK;
}
extern "C" int occ_ingest_pcm16(const int16_t* pcm_dev, const int64_t* offsets_dev, const int32_t* lens_dev, int B, int N, int mode,
                                float* wav_dev, void* stream) {
    occ_handle* h = nullptr;
    if (!pcm_dev || !offsets_dev || !lens_dev || !wav_dev || B < 0 || N < 1 || (mode != 0 && mode != 1))
        return fail(h, OCC_ERR_ARG, "occ_ingest_pcm16: bad argument");
    CK(launch_ingest_pcm16(pcm_dev, reinterpret_cast<const long long*>(offsets_dev), lens_dev, B, N, mode, wav_dev,
                           reinterpret_cast<cudaStream_t>(stream)));
    return OCC_OK;
}
extern "C" int occ_confusion_counts(const double* target_dev, int64_t n_t, const double* nontarget_dev, int64_t n_n, double threshold,
                                    uint64_t* counts_dev, void* stream) {
    occ_handle* h = nullptr;
    if (!counts_dev || n_t < 0 || n_n < 0 || (n_t > 0 && !target_dev) || (n_n > 0 && !nontarget_dev))
        return fail(h, OCC_ERR_ARG, "occ_confusion_counts: bad argument");
    CK(launch_confusion(target_dev, n_t, nontarget_dev, n_n, threshold, reinterpret_cast<unsigned long long*>(counts_dev),
                        reinterpret_cast<cudaStream_t>(stream)));
    return OCC_OK;
}
namespace occ {
int report_error(int code, const std::string& msg) { return fail(nullptr, code, msg); }
}  // namespace occ
extern "C" int64_t occ_eer_workspace_bytes(int64_t n) { return eer_workspace_bytes(n); }
extern "C" int occ_eer(const double* target_dev, int64_t n_t, const double* nontarget_dev, int64_t n_n, double* result_dev,
                       void* work_dev, void* stream) {
    occ_handle* h = nullptr;
    if (n_t < 1 || n_n < 1 || !target_dev || !nontarget_dev || !result_dev || !work_dev)
        return fail(h, OCC_ERR_ARG, "occ_eer: need >= 1 target and >= 1 nontarget score and non-null buffers");
    CK(launch_eer(target_dev, n_t, nontarget_dev, n_n, result_dev, work_dev, reinterpret_cast<cudaStream_t>(stream)));
    return OCC_OK;
}
extern "C" int occ_det_curve(const double* target_dev, int64_t n_t, const double* nontarget_dev, int64_t n_n, double* frr_dev,
                             double* far_dev, double* thr_dev, void* work_dev, void* stream) {
    occ_handle* h = nullptr;
    if (n_t < 1 || n_n < 1 || !target_dev || !nontarget_dev || !frr_dev || !far_dev || !thr_dev || !work_dev)
        return fail(h, OCC_ERR_ARG, "occ_det_curve: need >= 1 target and >= 1 nontarget score and non-null buffers");
    CK(launch_det_curve(target_dev, n_t, nontarget_dev, n_n, frr_dev, far_dev, thr_dev, work_dev, reinterpret_cast<cudaStream_t>(stream)));
    return OCC_OK;
}
extern "C" int occ_eer_host(const double* target, int64_t n_t, const double* nontarget, int64_t n_n, double* eer, double* thr, int device) {
    occ_handle* h = nullptr;
    if (n_t < 1 || n_n < 1 || !target || !nontarget || !eer || !thr)
        return fail(h, OCC_ERR_ARG, "occ_eer_host: need >= 1 target and >= 1 nontarget score");
    CK(cudaSetDevice(device));
    double *dt = nullptr, *dn = nullptr, *dr = nullptr;
    void* w = nullptr;
    int rc = OCC_OK;
    cudaError_t e = cudaSuccess;
    do {
        if ((e = cudaMalloc(&dt, n_t * 8)) != cudaSuccess) break;
        if ((e = cudaMalloc(&dn, n_n * 8)) != cudaSuccess) break;
        if ((e = cudaMalloc(&dr, 16)) != cudaSuccess) break;
        if ((e = cudaMalloc(&w, (size_t)eer_workspace_bytes(n_t + n_n))) != cudaSuccess) break;
        if ((e = cudaMemcpy(dt, target, n_t * 8, cudaMemcpyHostToDevice)) != cudaSuccess) break;
        if ((e = cudaMemcpy(dn, nontarget, n_n * 8, cudaMemcpyHostToDevice)) != cudaSuccess) break;
        if ((e = launch_eer(dt, n_t, dn, n_n, dr, w, 0)) != cudaSuccess) break;
        double res[2];
        if ((e = cudaMemcpy(res, dr, 16, cudaMemcpyDeviceToHost)) != cudaSuccess) break;
        *eer = res[0]; *thr = res[1];
    } while (0);
    if (e != cudaSuccess) rc = cuda_fail(h, e, "occ_eer_host");
    cudaFree(dt); cudaFree(dn); cudaFree(dr); cudaFree(w);
    return rc;
}

extern "C" int occ_forward_host(occ_handle* h, const float* wav_host, const int32_t* lengths_host, int B, int N,
                                const float* ref_host, float eps, float* emb_host, float* out_host, float* score_host, void* stream) {
    if (!h || !wav_host) return fail(h, OCC_ERR_ARG, "occ_forward_host: null argument");
    if (B < 1 || B > h->max_batch || N > h->max_samples) return fail(h, OCC_ERR_ARG, "occ_forward_host: B / N exceed the handle's workspace");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(h->device));
    CK(cudaMemcpyAsync(h->wav_dev, wav_host, sizeof(float) * (size_t)B * N, cudaMemcpyHostToDevice, s));
    int r = occ_forward(h, h->wav_dev, lengths_host, B, N, h->emb_dev, h->out_dev, nullptr, stream);
    if (r) return r;
    if (ref_host) {
        CK(cudaMemcpyAsync(h->ref_dev, ref_host, sizeof(float) * 160, cudaMemcpyHostToDevice, s));
        LAUNCH(1, launch_score_l2(h->emb_dev, h->ref_dev, B, 160, eps, h->score_dev, s));
        if (score_host) CK(cudaMemcpyAsync(score_host, h->score_dev, sizeof(float) * B, cudaMemcpyDeviceToHost, s));
    }
    if (emb_host) CK(cudaMemcpyAsync(emb_host, h->emb_dev, sizeof(float) * (size_t)B * 160, cudaMemcpyDeviceToHost, s));
    if (out_host) CK(cudaMemcpyAsync(out_host, h->out_dev, sizeof(float) * (size_t)B * 2, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return OCC_OK;
}

// ================================================================================================ multi-GPU exchange
// The path's ONE collective (replaces the gather of nn.DataParallel, oc_classifier.py:343): every rank contributes
// n_per_rank fp32 scores, every rank receives world * n_per_rank in rank order.  Plain ncclAllGather on the caller's
// communicator and stream (285 KB for ASVspoof2019-LA-eval: latency-bound, nothing to fuse with).  NCCL is resolved at run
// time from the process (the host framework has already loaded it), so the library itself does not link libnccl.
extern "C" int occ_allgather_scores(const float* local_dev, int64_t n_per_rank, float* all_dev, void* nccl_comm, void* stream) {
    occ_handle* h = nullptr;
    if (!local_dev || !all_dev || !nccl_comm || n_per_rank < 0) return fail(h, OCC_ERR_ARG, "occ_allgather_scores: bad argument");
    typedef int (*allgather_fn)(const void*, void*, size_t, int, void*, cudaStream_t);
    typedef const char* (*errstr_fn)(int);
    static std::mutex mu;
    static allgather_fn ag = nullptr;
    static errstr_fn es = nullptr;
    {
        std::lock_guard<std::mutex> lk(mu);
        if (!ag) {
            void* sym = dlsym(RTLD_DEFAULT, "ncclAllGather");
            void* lib = nullptr;
            if (!sym) {
                lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
                if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW);
                if (!lib) lib = dlopen("libnccl.so", RTLD_NOW);
                if (lib) sym = dlsym(lib, "ncclAllGather");
            }
            if (!sym) return fail(h, OCC_ERR_STATE, "occ_allgather_scores: ncclAllGather not found in this process (is NCCL loaded?)");
            ag = reinterpret_cast<allgather_fn>(sym);
            void* e2 = lib ? dlsym(lib, "ncclGetErrorString") : dlsym(RTLD_DEFAULT, "ncclGetErrorString");
            es = reinterpret_cast<errstr_fn>(e2);
        }
    }
    if (n_per_rank == 0) return OCC_OK;
    const int rc = ag(local_dev, all_dev, (size_t)n_per_rank, /*ncclFloat32*/ 7, nccl_comm, reinterpret_cast<cudaStream_t>(stream));
    if (rc != 0) return fail(h, OCC_ERR_CUDA, std::string("ncclAllGather failed: ") + (es ? es(rc) : "unknown NCCL error"));
    return OCC_OK;
}

// ================================================================================================ debug hooks
extern "C" int occ_debug_gemm(int impl, int M, int N, int K, int split, const float* A_dev, const float* W_dev, const float* bias_dev,
                              const float* resid_dev, int gelu, float* out_dev, void* stream) {
    occ_handle* h = nullptr;
    if (M < 1 || N % 64 != 0 || K % 64 != 0 || !A_dev || !W_dev || !out_dev) return fail(h, OCC_ERR_ARG, "occ_debug_gemm: N, K multiples of 64 required");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    bf16 *a = nullptr, *w = nullptr;
    const int64_t na = (int64_t)M * K + kSlack, nw = (int64_t)N * K + kSlack;
    CK(cudaMalloc(&a, sizeof(bf16) * na * 2));
    cudaError_t e = cudaMalloc(&w, sizeof(bf16) * nw * 2);
    if (e != cudaSuccess) { cudaFree(a); return cuda_fail(h, e, "cudaMalloc"); }
    cudaMemsetAsync(a, 0, sizeof(bf16) * na * 2, s);
    cudaMemsetAsync(w, 0, sizeof(bf16) * nw * 2, s);
    launch_f32_to_bf16(A_dev, (int64_t)M * K, a, na, s);
    launch_f32_to_bf16(W_dev, (int64_t)N * K, w, nw, s);
    GemmDesc d = lin_desc(a, na, M, K, w, nw, N, split);
    d.bias = bias_dev; d.resid = resid_dev; d.ld_resid = N; d.gelu = gelu; d.out32 = out_dev; d.ld32 = N;
    std::string err;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (impl == 0) e = launch_gemm_simt(d, s); else e = launch_gemm_tc(d, s, sms, &err);
    cudaError_t e2 = cudaStreamSynchronize(s);
    cudaFree(a); cudaFree(w);
    if (e != cudaSuccess) return fail(h, OCC_ERR_CUDA, std::string("occ_debug_gemm: ") + cudaGetErrorString(e) + " " + err);
    if (e2 != cudaSuccess) return cuda_fail(h, e2, "occ_debug_gemm sync");
    return OCC_OK;
}

// Development probe: average device time of `iters` back-to-back launches of one linear GEMM (bf16 out when out16 != 0,
// else fp32), optional GELU / residual.  probe bits: 1 = no TMA loads, 2 = no MMAs, 4 = no epilogue work.
extern "C" int occ_debug_gemm_bench(int M, int N, int K, int gelu, int resid, int out16, int probe, int iters, double* ms_avg) {
    occ_handle* h = nullptr;
    if (M < 1 || N % 64 != 0 || K % 64 != 0 || iters < 1 || !ms_avg) return fail(h, OCC_ERR_ARG, "occ_debug_gemm_bench: bad argument");
    bf16 *a = nullptr, *w = nullptr, *o16 = nullptr;
    float *o32 = nullptr, *bias = nullptr;
    const int64_t na = (int64_t)M * K + kSlack, nw = (int64_t)N * K + kSlack;
    CK(cudaMalloc(&a, sizeof(bf16) * na));
    CK(cudaMalloc(&w, sizeof(bf16) * nw));
    CK(cudaMalloc(&o16, sizeof(bf16) * (int64_t)M * N));
    CK(cudaMalloc(&o32, sizeof(float) * (int64_t)M * N));
    CK(cudaMalloc(&bias, sizeof(float) * N));
    CK(cudaMemset(a, 0, sizeof(bf16) * na)); CK(cudaMemset(w, 0, sizeof(bf16) * nw));
    CK(cudaMemset(o32, 0, sizeof(float) * (int64_t)M * N)); CK(cudaMemset(bias, 0, sizeof(float) * N));
    GemmDesc d = lin_desc(a, na, M, K, w, nw, N, 0);
    d.bias = bias; d.gelu = gelu;
    if (resid) { d.resid = o32; d.ld_resid = N; }
    if (out16) { d.out16 = o16; d.ld16 = N; } else { d.out32 = o32; d.ld32 = N; }
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    std::string err;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaError_t e = cudaSuccess;
    for (int i = 0; i < 3 && e == cudaSuccess; ++i) e = launch_gemm_tc(d, 0, sms, &err, probe);
    cudaEventRecord(e0, 0);
    for (int i = 0; i < iters && e == cudaSuccess; ++i) e = launch_gemm_tc(d, 0, sms, &err, probe);
    cudaEventRecord(e1, 0);
    cudaError_t e2 = cudaDeviceSynchronize();
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    *ms_avg = ms / iters;
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(a); cudaFree(w); cudaFree(o16); cudaFree(o32); cudaFree(bias);
    if (e != cudaSuccess) return fail(h, OCC_ERR_CUDA, std::string("occ_debug_gemm_bench: ") + cudaGetErrorString(e) + " " + err);
    if (e2 != cudaSuccess) return cuda_fail(h, e2, "occ_debug_gemm_bench sync");
    return OCC_OK;
}

extern "C" int occ_debug_attention(int impl, const float* qkv_dev, int B, int T, const int32_t* klen_dev, float* ctx_dev, void* stream) {
    // impl 0: fp32 SIMT attention, impl 1: bf16 mma.sync attention.  qkv fp32 [B*T,3072] (q already scaled), ctx fp32 [B*T,1024].
    occ_handle* h = nullptr;
    if (!qkv_dev || !ctx_dev || B < 1 || T < 1) return fail(h, OCC_ERR_ARG, "occ_debug_attention: bad argument");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const int64_t M = (int64_t)B * T;
    bf16 *q16 = nullptr, *c16 = nullptr;
    CK(cudaMalloc(&q16, sizeof(bf16) * M * 3072));
    cudaError_t e = cudaMalloc(&c16, sizeof(bf16) * M * 1024 * 2);
    if (e != cudaSuccess) { cudaFree(q16); return cuda_fail(h, e, "cudaMalloc"); }
    cudaMemsetAsync(c16, 0, sizeof(bf16) * M * 1024 * 2, s);
    if (impl == 1) {
        launch_f32_to_bf16(qkv_dev, M * 3072, q16, 0, s);
        int dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        e = launch_attn_bf16(q16, B, T, klen_dev, c16, sms, s);
    } else {
        e = launch_attn_f32(qkv_dev, B, T, klen_dev, c16, M * 1024, s);
    }
    // ctx fp32 = hi (+ lo)
    if (e == cudaSuccess) e = launch_bf16_planes_to_f32(c16, impl == 1 ? 0 : M * 1024, M * 1024, ctx_dev, s);
    cudaError_t e2 = cudaStreamSynchronize(s);
    cudaFree(q16); cudaFree(c16);
    if (e != cudaSuccess) return cuda_fail(h, e, "occ_debug_attention");
    if (e2 != cudaSuccess) return cuda_fail(h, e2, "occ_debug_attention sync");
    return OCC_OK;
}

extern "C" int occ_debug_attention_trace(int64_t* host_out, int n) {
    occ_handle* h = nullptr;
    if (!host_out || n < 1) return fail(h, OCC_ERR_ARG, "occ_debug_attention_trace: bad argument");
    CK(cudaDeviceSynchronize());
    CK(attn_tc2_trace_read(reinterpret_cast<long long*>(host_out), n));
    return OCC_OK;
}
