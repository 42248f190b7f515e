/* occ_b200.h -- C ABI of libocc_b200.so: the B200 (sm_100a) scoring hot path of
 * nguyenvulong/occm (raw 16 kHz waveform -> XLS-R/wav2vec 2.0 -> AASIST -> one-class
 * score -> EER).  Plain pointers and sizes only; no torch types.  Every entry point
 * returns 0 on success and a negative code on failure (message: occ_last_error);
 * nothing inside the library calls exit()/abort().  All kernels are enqueued on the
 * cudaStream_t passed as `void* stream` (NULL = legacy default stream); the library
 * never synchronises the device unless the entry point says so.
 *
 * Reference interfaces replaced (paths relative to the reference repo):
 *   occ_create / occ_load_weights / occ_finalize_weights
 *        <- AModel.__init__ + load_state_dict      models/sslassist.py:433-504, oc_classifier.py:335-340
 *   occ_forward            <- AModel.forward       models/sslassist.py:506-597
 *   occ_extract_feat       <- SSLModel.extract_feat models/sslassist.py:31-49, models/xlsr.py:39-52
 *   occ_score_l2           <- F.pairwise_distance  oc_classifier.py:261
 *   occ_ref_embedding      <- create_reference_embedding2  oc_classifier.py:182-197
 *   occ_score_cos          <- OCSoftmax score (named by BASELINE.json; absent from the reference)
 *   occ_eer                <- compute_det_curve + compute_eer  evaluate_metrics.py:3-40
 *   occ_det_curve          <- compute_det_curve (full curve)   evaluate_metrics.py:3-21
 *   occ_confusion_counts   <- calculate_confusion_matrix       evaluate_metrics.py:23-32
 *   occ_allgather_scores   <- nn.DataParallel gather           oc_classifier.py:343
 *   occ_ingest_pcm16       <- librosa.load + data_utils_SSL.pad  oc_classifier.py:93-95, data_utils_SSL.py:47-54
 *   occ_forward_host       <- the per-utterance loop body of score_eval_set_1c2, oc_classifier.py:256-261
 */
#ifndef OCC_B200_H
#define OCC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct occ_handle occ_handle;

#define OCC_OK 0
#define OCC_ERR_ARG (-1)     /* bad argument / shape */
#define OCC_ERR_CUDA (-2)    /* CUDA runtime or driver error */
#define OCC_ERR_STATE (-3)   /* call order (weights not finalized, ...) */
#define OCC_ERR_NOMEM (-4)
#define OCC_ERR_WEIGHT (-5)  /* unknown / missing / mis-shaped weight */

#define OCC_DTYPE_FP32 0     /* fp32-accurate: split-bf16 (3 MMA) tensor-core GEMMs, fp32 everything else */
#define OCC_DTYPE_BF16 1     /* bf16 operands, fp32 accumulate / LayerNorm / softmax / residual / backend */

#define OCC_EMB_DIM 160
#define OCC_FEAT_DIM 1024

int occ_version(void);

/* One handle per GPU / host thread.  Allocates the activation workspace for up to
 * max_batch utterances of max_samples samples (no allocation on the forward path). */
int occ_create(occ_handle** h, int device, int dtype, int max_batch, int max_samples);
void occ_destroy(occ_handle* h);
const char* occ_last_error(occ_handle* h);

/* Called once per state-dict entry, with the reference's key names
 * ("ssl_model.model.encoder.layers.0.fc1.weight", "LL.weight", ...).  `ptr` may be a host
 * or a device pointer; data is copied.  dtype: 0 = float32, 2 = int64 (num_batches_tracked, ignored).
 * Unused fairseq keys (mask_emb, quantizer.*, project_q.*, final_proj.*) and bn1.* (dead code at
 * models/sslassist.py:408-415) are accepted and dropped.  Both pos-conv spellings are accepted
 * (weight_g/weight_v and parametrizations.weight.original0/1). */
int occ_load_weights(occ_handle* h, const char* name, const void* ptr, const int64_t* shape, int ndim, int dtype);
/* Folds weight-norm, eval BatchNorm and the 1/sqrt(64) query scale, re-lays conv weights K-major,
 * converts GEMM weights to bf16 (+ low-order plane in OCC_DTYPE_FP32).  Synchronises. */
int occ_finalize_weights(occ_handle* h);

/* wav_dev [B,N] fp32 device.  lengths_host: NULL (all N) or B valid sample counts (host memory).
 * emb_dev [B,160], out_dev [B,2] fp32 device (either may be NULL).
 * hidden_dev: NULL or fp32 [24,B,T,1024] receiving every encoder layer's output (pre final LN). */
int occ_forward(occ_handle* h, const float* wav_dev, const int32_t* lengths_host, int B, int N,
                float* emb_dev, float* out_dev, float* hidden_dev, void* stream);

/* Frontend only: feat_dev [B,T,1024] fp32 (final LayerNorm applied, = fairseq ['x']). */
int occ_extract_feat(occ_handle* h, const float* wav_dev, const int32_t* lengths_host, int B, int N,
                     float* feat_dev, float* hidden_dev, void* stream);

/* SURVEY 8(f) N4 -- the LCNN backend on the shared frontend (models/lcnn.py).  Available when lcnn.* weights were loaded
 * (the reference's ssl_lcnn state dict: frontend.model.* + lcnn.*; `frontend.model.` is accepted as an alias of `ssl_model.model.`).
 * occ_forward_lcnn  <- ssl_lcnn.forward  models/lcnn.py:242-260 : wav_dev [B,N] -> out_dev [B,2] logits
 * occ_lcnn_backend  <- LCNN.forward      models/lcnn.py:189-213 : feat_dev fp32 [B,T,1024] -> out_dev [B,2]   (T >= 8)
 * Likewise the SE-ResNet34 backend (models/senet.py), when resnet34.* weights were loaded (ssl_resnet34's layout; OCCM's
 * `senet34_branch.` / `lcnn_branch.` prefixes are accepted as aliases of `resnet34.` / `lcnn.`):
 * occ_forward_resnet34 <- ssl_resnet34.forward models/senet.py:168-185 : wav_dev [B,N] -> com_dev [B,128], des_dev [B,2]
 * occ_resnet34_backend <- ResNet.forward      models/senet.py:121-143 : feat_dev fp32 [B,T,1024] -> com_dev, des_dev */
int occ_forward_resnet34(occ_handle* h, const float* wav_dev, int B, int N, float* com_dev, float* des_dev, void* stream);
int occ_resnet34_backend(occ_handle* h, const float* feat_dev, int B, int T, float* com_dev, float* des_dev, void* stream);
/* LCNN entry points (see above). */
int occ_forward_lcnn(occ_handle* h, const float* wav_dev, int B, int N, float* out_dev, void* stream);
int occ_lcnn_backend(occ_handle* h, const float* feat_dev, int B, int T, float* out_dev, void* stream);

/* Number of frames T for N samples (seven conv layers). */
int occ_num_frames(int n_samples);

/* Stage entry points (tests / ncu).  All tensors fp32 device, shapes as documented in DESIGN.md. */
int occ_conv_encoder(occ_handle* h, const float* wav_dev, int B, int N, float* proj_dev /*[B,T,1024]*/, void* stream);
int occ_pos_conv(occ_handle* h, const float* x_dev /*[B,T,1024]*/, int B, int T, float* y_dev /*x+pos*/, void* stream);
int occ_transformer(occ_handle* h, const float* x_dev /*[B,T,1024]*/, int B, int T, int n_layers,
                    float* y_dev /*[B,T,1024] after final LN*/, float* hidden_dev, void* stream);
int occ_aasist_backend(occ_handle* h, const float* feat_dev /*[B,T,1024]*/, int B, int T,
                       float* emb_dev, float* out_dev, void* stream);

/* score[b] = || ref - emb[b] + eps ||_2   (torch pairwise_distance; eps = 1e-6 in the reference). */
int occ_score_l2(const float* emb_dev, const float* ref_dev, int B, int D, float eps, float* score_dev, void* stream);
/* score[b] = <emb[b]/||emb[b]||, w/||w||>   (OC-Softmax). */
int occ_score_cos(const float* emb_dev, const float* center_dev, int B, int D, float* score_dev, void* stream);
/* ref[d] = mean_i emb[i][d] (index order); thr = max_i ||ref - emb[i] + eps||.  ref_dev [D], thr_dev [1]. */
int occ_ref_embedding(const float* emb_dev, int n, int D, float eps, float* ref_dev, float* thr_dev, void* stream);

/* EER on the GPU, bit-identical to numpy's evaluate_metrics.compute_eer on the same float64 scores.
 * target_dev / nontarget_dev: device float64.  result_dev: device double[2] = {eer, threshold}.
 * work_dev: device scratch of occ_eer_workspace_bytes(n_t + n_n) bytes. */
int64_t occ_eer_workspace_bytes(int64_t n);
int occ_eer(const double* target_dev, int64_t n_t, const double* nontarget_dev, int64_t n_n,
            double* result_dev, void* work_dev, void* stream);
/* Full DET curve, compute_det_curve of evaluate_metrics.py:3-21: frr_dev / far_dev / thr_dev receive n_t + n_n + 1 float64
 * points each, bit-identical to numpy (same stable sort, integer cumulative counts, one float64 division per point).
 * work_dev as for occ_eer. */
int occ_det_curve(const double* target_dev, int64_t n_t, const double* nontarget_dev, int64_t n_n, double* frr_dev,
                  double* far_dev, double* thr_dev, void* work_dev, void* stream);
/* occ_eer without any allocation -- no counterpart in the reference; added for
 * callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments and result as occ_eer minus work_dev: the sort scratch is a file-scope `static __device__` array inside the
 * library (3.06 MiB, resident on every GPU that loads the library until the process exits), so n_t + n_n is limited to
 * OCC_EER_NOALLOC_MAX_SCORES; larger inputs return OCC_ERR_ARG with a message naming the limit (use occ_eer then).
 * Calls no cudaMalloc / cudaFree / allocating library routine.  Runs on the CURRENT device; the first call on each device
 * resolves the array's device address (cudaGetSymbolAddress) -- make that call outside stream capture.  All calls on one
 * device share the array: calls that overlap on the GPU (two streams, concurrent graph launches) race -- serialise them.
 * Nothing has to be cleared between calls (every pass writes what the next reads). */
#define OCC_EER_NOALLOC_MAX_SCORES 131072
int occ_eer_noalloc(const double* target_dev, int64_t n_t, const double* nontarget_dev, int64_t n_n, double* result_dev,
                    void* stream);
/* Convenience: host float64 arrays in, host results out (allocates, copies, synchronises). */
int occ_eer_host(const double* target, int64_t n_t, const double* nontarget, int64_t n_n, double* eer, double* thr,
                 int device);

/* The path's one exchange step at N > 1 GPUs (replaces nn.DataParallel's gather, oc_classifier.py:343): ncclAllGather of
 * n_per_rank fp32 scores per rank into all_dev [world * n_per_rank] on every rank, enqueued on `stream`.  nccl_comm is the
 * host's ncclComm_t passed as an opaque pointer (torch: ProcessGroupNCCL._comm_ptr()); NCCL is resolved from the process at
 * run time.  Ragged shards are padded to a common n_per_rank by the caller. */
int occ_allgather_scores(const float* local_dev, int64_t n_per_rank, float* all_dev, void* nccl_comm, void* stream);

/* Waveform ingest (the loader side of the path): ragged int16 PCM utterances -> fp32 batch wav_dev [B,N], sample / 32768
 * (librosa.load of a PCM16 file, oc_classifier.py:93-95).  Utterance b = pcm_dev[offsets_dev[b] .. + lens_dev[b]).
 * mode 0: zero-pad / truncate to N (variable-length batch; pass the lengths to occ_forward);
 * mode 1: data_utils_SSL.pad (data_utils_SSL.py:47-54): truncate, or repeat the utterance until N samples. */
int occ_ingest_pcm16(const int16_t* pcm_dev, const int64_t* offsets_dev, const int32_t* lens_dev, int B, int N, int mode,
                     float* wav_dev, void* stream);
/* calculate_confusion_matrix (evaluate_metrics.py:23-32) on device float64 scores:
 * counts_dev[4] = {tp = #(target > thr), tn = #(nontarget <= thr), fp = #(nontarget > thr), fn = #(target <= thr)}. */
int occ_confusion_counts(const double* target_dev, int64_t n_t, const double* nontarget_dev, int64_t n_n, double threshold,
                         uint64_t* counts_dev, void* stream);

/* End-to-end call with HOST buffers (pinned or pageable): H2D of wav, forward, one-class scores against
 * ref_host [160] (NULL = skip scoring), D2H of emb/out/scores.  Synchronises the stream before returning.
 * emb_host [B,160], out_host [B,2], score_host [B]; any may be NULL. */
int occ_forward_host(occ_handle* h, const float* wav_host, const int32_t* lengths_host, int B, int N,
                     const float* ref_host, float eps, float* emb_host, float* out_host, float* score_host,
                     void* stream);

/* Debug / test hooks. */
int occ_debug_gemm(int impl /*0 = SIMT check kernel, 1 = tcgen05*/, int M, int N, int K, int split,
                   const float* A_dev, const float* W_dev, const float* bias_dev, const float* resid_dev,
                   int gelu, float* out_dev, void* stream);
/* Development probe: average device ms of `iters` back-to-back launches of one [M,K]x[N,K]^T tcgen05 GEMM on zero data.
 * probe bits: 1 = skip the TMA loads, 2 = skip the MMAs, 4 = skip the epilogue work (isolates each pipeline). */
int occ_debug_gemm_bench(int M, int N, int K, int gelu, int resid, int out16, int probe, int iters, double* ms_avg);
/* impl 0 = fp32 SIMT attention, 1 = bf16 tensor-core attention.  qkv fp32 [B*T,3072] (q pre-scaled), klen_dev device int32[B] or NULL,
 * ctx fp32 [B*T,1024].  Synchronises. */
int occ_debug_attention(int impl, const float* qkv_dev, int B, int T, const int32_t* klen_dev, float* ctx_dev, void* stream);
/* Development trace of the T <= 256 attention kernel (enabled by OCC_ATTN_FLAGS bit 2): copies n int64 clock stamps
 * [cta < 160][item < 8][slot * 8 + event] to host_out.  Synchronises the device. */
int occ_debug_attention_trace(int64_t* host_out, int n);
int64_t occ_kernel_launches(occ_handle* h);   /* kernels launched by this handle so far */

/* Per-range device timing for bench.py's roofline: while enabled, every kernel family of the forward path is bracketed
 * by CUDA events on the launch stream.  occ_profile(h, 1) clears and starts, occ_profile(h, 0) stops (ranges kept);
 * occ_profile_read sums one family: elapsed ms, number of ranges (= launches for the GEMM families) and the
 * ALGORITHMIC work those launches performed: flops (valid rows only, 2*M*N*K) for the tensor families and the backend,
 * compulsory HBM BYTES (inputs read once + outputs written once) for the memory-bound families NORM and CONV0.
 * Synchronises on the recorded events. */
#define OCC_PROF_XF_GEMM 0    /* transformer QKV / out-proj / FC1 / FC2 (gemm_tc_kernel<256>) */
#define OCC_PROF_CONV_GEMM 1  /* conv feature encoder layers 1-6 (implicit GEMM) */
#define OCC_PROF_POS_GEMM 2   /* grouped positional conv */
#define OCC_PROF_ATTN 3       /* fused attention */
#define OCC_PROF_NORM 4       /* LayerNorm (+GELU) row kernels */
#define OCC_PROF_CONV0 5      /* conv layer 0 + LN + GELU */
#define OCC_PROF_BACKEND 6    /* whole AASIST backend */
#define OCC_PROF_PROJ_GEMM 7  /* post_extract_proj */
#define OCC_PROF_XF_QKV 8     /* the four transformer GEMMs separately (tag 0 is their sum: occ_profile_read(0) adds 8..11) */
#define OCC_PROF_XF_OUT 9
#define OCC_PROF_XF_FC1 10
#define OCC_PROF_XF_FC2 11
#define OCC_PROF_NTAGS 12
int occ_profile(occ_handle* h, int enable);
int occ_profile_read(occ_handle* h, int tag, double* ms_total, int64_t* ranges, double* flops_total);

#ifdef __cplusplus
}
#endif
#endif /* OCC_B200_H */
