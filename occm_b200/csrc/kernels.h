// kernels.h -- launchers exported by the kernel translation units (internal).
#pragma once
#include <string>

#include "occ_internal.h"

namespace occ {

// frontend.cu
// frag != nullptr: the mma.sync form (tables written by launch_conv0_pack_frag, conv0_frag_bytes() bytes); else the fp32x2 SIMT form
cudaError_t launch_conv0(const float* wav, int B, int N, int L0, const float* w, const float* bias, const float* gamma,
                         const float* beta, bf16* out, long long plane_stride, cudaStream_t s, const void* frag = nullptr, int num_sms = 0);
cudaError_t launch_conv0_pack_frag(const float* w, const float* bias, const float* gamma, const float* beta, void* frag, cudaStream_t s);
size_t conv0_frag_bytes();
cudaError_t launch_rows_ln(const float* x, long long M, int W, bool gelu, const float* g1, const float* b1, const float* g2,
                           const float* b2, float* out32, bf16* out16, long long plane_stride, cudaStream_t s,
                           const bf16* add16 = nullptr, const float* add32 = nullptr, float* xsum = nullptr, const bf16* x16 = nullptr);
cudaError_t launch_f32_to_bf16(const float* x, long long n, bf16* hi, long long plane_stride, cudaStream_t s);
cudaError_t launch_bf16_planes_to_f32(const bf16* hi, long long plane_stride, long long n, float* out, cudaStream_t s);
cudaError_t launch_attn_bf16(const bf16* qkv, int B, int T, const int* klen, bf16* ctx, int num_sms, cudaStream_t s);
cudaError_t launch_attn_f32(const float* qkv, int B, int T, const int* klen, bf16* ctx, long long plane_stride, cudaStream_t s);

// backend.cu
struct GatW {
    const float *Wp, *bp;
    const float *aw11, *aw22, *aw12;
    const float *Wa, *ba, *Wo, *bo;
    const float *bns, *bnt;
    const float *WpM, *bpM, *awM, *WaM, *baM, *WoM, *boM;
    const float* master_in;
    long long master_stride;
    float* master_out;
    float temp;
};
struct GatArgs {
    GatW w[2];
    const float* x[2];
    float* y[2];
    const int* n1[2];
    const int* n2[2];
    int n1c[2], n2c[2];
    int nmax[2];
    int has_master;
};
struct PoolArgs {
    const float* h[4];
    long long h_bstride[4];
    const int* n[4];
    int nc[4];
    const float* w[4];
    const float* bias[4];
    float* out[4];
    int kmax[4];
    int* kout[4];
    const int* noff[4];
    int noffc[4];
};
struct HprojArgs {
    const float* x1[2]; const float* x2[2];
    int n1max[2], n2max[2];
    const int* n1[2]; const int* n2[2]; int n1c[2], n2c[2];
    const float *W1[2], *b1[2], *W2[2], *b2[2];
    float* xc[2];
};
struct ReadoutArgs {
    const float* T1[2]; const float* S1[2];
    const float* y2[2];
    const float* m[2]; const float* maug[2];
    int kTmax, kSmax;
    const int* kT; const int* kS; int kTc, kSc;
    const float* Wout; const float* bout;
    float* emb; float* out;
};
cudaError_t launch_ll_pool(const float* ll, int B, int T, int Wm, const int* wlen, float scale, float shift, float* x0, cudaStream_t s);
// first residual block front (Cin = 1): A = selu(bn2(conv1(x0))) and I = conv_downsample(x0), both written as channel-last
// zero-padded bf16 hi/lo planes [B][Hp][Wp][64] (channels 32..63 and the border are zeros)
cudaError_t launch_res0_front(const float* x0, int B, int Wm, const int* wlen, const float* w1, const float* b1, const float* s2,
                              const float* t2, const float* wd, const float* bd, int Hp, bf16* A, bf16* I, long long plane, cudaStream_t s);
// strided view of the residual-encoder output: element (b, c, h, w) at x[b*xb + c*xc + h*xh + w*xw]
struct XView { const float* x; long long xb, xc, xh, xw; };
// W1 [128][64], W2 [64][128] fp32 as stored in the checkpoint; wmap channel-last [B][42*Wm][64]   (backend_attmap.cu)
cudaError_t launch_att_map(XView x, int B, int Wm, const float* W1, const float* b1, const float* bs, const float* bt,
                           const float* W2, const float* b2, float* wmap, int num_sms, cudaStream_t s);
cudaError_t launch_st_pool(XView x, const float* wmap, int B, int Wm, const int* wlen, const float* posS, float* eS,
                           float* eT, cudaStream_t s);
cudaError_t launch_gat(const GatArgs& a, int B, int nz, int din, int dout, cudaStream_t s);
cudaError_t launch_graph_pool(const PoolArgs& a, int B, int nz, int D, int nmax, cudaStream_t s);
cudaError_t launch_hproj(const HprojArgs& a, int B, int nz, int D, cudaStream_t s);
cudaError_t launch_readout(const ReadoutArgs& a, int B, cudaStream_t s);

// lcnn.cu -- LCNN backend (models/lcnn.py): weights in their checkpoint layout, eval BatchNorm folded to scale / shift
struct LcnnW {
    const float *c1w, *c1b;                       // layer1.0.filter        [8][1][5][5]
    const float *a2w, *a2b, *k2w, *k2b;           // layer2.0.conv_a.filter [8][4][1][1], layer2.0.conv.filter [16][4][3][3]
    const float *bn2s, *bn2t;                     // layer2.2 (BatchNorm2d(8))
    const float *a3w, *a3b, *k3w, *k3b;           // layer3.0.conv_a.filter [16][8][1][1], layer3.0.conv.filter [32][8][3][3]
    const float *bn3s, *bn3t;                     // layer3.2 (BatchNorm2d(16))
    const float *f0w, *f0b, *f1w, *f1b, *f2w, *f2b, *f3w, *f3b;
};
// feat fp32 [B,T,1024] -> out [B,2]; scratch: y1 / a2 [B,4,T/2,512], y2 / a3 [B,8,T/4,256], y3 [B,16,T/8,128]
cudaError_t launch_lcnn(const LcnnW& w, const float* feat, int B, int T, float* y1, float* a2, float* y2, float* a3, float* y3,
                        float* out, cudaStream_t s);

// senet.cu -- SE-ResNet34 backend (models/senet.py); conv weights in their checkpoint layout, eval BatchNorm folded
struct ResnetBlockW {
    int planes, stride;
    const float *c1, *b1s, *b1t, *c2, *b2s, *b2t, *se1, *se2;
    const float *dw, *dbs, *dbt;                  // downsample 1x1 conv + BN, or nullptr
    const void *c1f, *c2f;                        // the 3x3 convs' weights in mma.sync B-fragment order (rs_pack_frag), or nullptr
};
struct ResnetW {
    const float *conv1, *bn1s, *bn1t;
    const void* conv1f;                           // the 7x7 stem's weights in mma.sync B-fragment order (rs_pack_stem), or nullptr
    ResnetBlockW blk[16];
    int nblocks;
    const float *emb_w, *emb_b, *cls_w, *cls_b;
};
// feat fp32 [B,T,1024] -> com [B,128], des [B,2]; buf0 [B,16,H1,512] (H1 = (T-1)/2+1), buf1..4 [B,16,H2,256] (H2 = (H1-1)/2+1), vec [2][B,128]
cudaError_t launch_resnet34(const ResnetW& w, const float* feat, int B, int T, float* buf0, float* buf1, float* buf2, float* buf3,
                            float* buf4, float* vec, float* com, float* des, cudaStream_t s, int* launches, int num_sms = 148);
cudaError_t launch_rs_pack_frag(const float* w /*[Co][Ci][3][3]*/, int Ci, int Co, void* frag, cudaStream_t s);
size_t rs_frag_bytes(int Ci, int Co);
cudaError_t launch_rs_pack_stem(const float* w /*[16][1][7][7]*/, void* frag, cudaStream_t s);
size_t rs_stem_frag_bytes();

// eer.cu
cudaError_t launch_score_l2(const float* emb, const float* ref, int B, int D, float eps, float* out, cudaStream_t s);
cudaError_t launch_score_cos(const float* emb, const float* w, int B, int D, float* out, cudaStream_t s);
cudaError_t launch_ref_embedding(const float* emb, int n, int D, float eps, float* ref, float* thr, cudaStream_t s);
int64_t eer_workspace_bytes(int64_t n);
cudaError_t launch_eer(const double* target, int64_t nt, const double* nontarget, int64_t nn, double* result, void* work, cudaStream_t s);
cudaError_t launch_det_curve(const double* target, int64_t nt, const double* nontarget, int64_t nn, double* frr, double* far, double* thr,
                             void* work, cudaStream_t s);

// ingest.cu
cudaError_t launch_ingest_pcm16(const int16_t* pcm, const long long* offsets, const int* lens, int B, int N, int mode, float* wav, cudaStream_t s);
cudaError_t launch_confusion(const double* target, long long nt, const double* nontarget, long long nn, double thr,
                             unsigned long long* counts, cudaStream_t s);

// weights.cu
cudaError_t launch_pack_linear(const float* in, long long rows, long long cols, long long scale_rows, float scale, bf16* out,
                               long long plane, cudaStream_t s);
cudaError_t launch_pack_linear_ln(const float* in, long long rows, long long cols, long long scale_rows, float scale, const float* gamma,
                                  const float* beta, const float* bias, bf16* out, float* colsum, float* cbias, cudaStream_t s);
cudaError_t launch_pack_conv1d(const float* in, int Cout, int Cin, int k, bf16* out, long long plane, cudaStream_t s);
cudaError_t launch_pack_posconv(const float* v, const float* g, float* norms_tmp, bf16* out, long long plane, cudaStream_t s);
cudaError_t launch_pack_posconv_toeplitz(const float* v, const float* g, float* norms_tmp, bf16* out, cudaStream_t s);
cudaError_t launch_fold_bn(const float* w, const float* b, const float* mean, const float* var, int C, float* scale, float* shift, cudaStream_t s);
cudaError_t launch_pack_conv2d(const float* in, int Cout, int Cin, int taps, float* out, cudaStream_t s);
cudaError_t launch_pack_conv2d_tc(const float* in, int Cout, int Cin, int taps, bf16* out, long long plane, cudaStream_t s);
cudaError_t launch_pad_vec(const float* in, int n, float fill, float* out, int total, cudaStream_t s);
cudaError_t launch_transpose(const float* in, int R, int C, float* out, cudaStream_t s);
cudaError_t launch_scale_copy(const float* in, int n, int scale_n, float scale, float* out, cudaStream_t s);

cudaError_t attn_tc2_trace_read(long long* host_out, int n);   // development trace of attn_tc2 (OCC_ATTN_FLAGS bit 2)
// sets the thread's occ_last_error(NULL) message and returns `code` (occ_api.cu); for entry points defined outside occ_api.cu
int report_error(int code, const std::string& msg);

}  // namespace occ
