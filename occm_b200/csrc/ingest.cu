// ingest.cu -- waveform ingest and threshold counts: the callers on either side of the scoring path.
//   occ_ingest_pcm16 : ragged int16 PCM utterances (what the FLAC / WAV files of ASVspoof hold; librosa / soundfile scale
//                      them by 1/32768 to fp32, oc_classifier.py:93-95) -> one fp32 batch [B, N] on the device.
//                      mode 0: zero-pad / truncate (variable-length batch, lengths handed to occ_forward);
//                      mode 1: data_utils_SSL.pad (/root/reference/data_utils_SSL.py:47-54): truncate to N, or tile the
//                              utterance (np.tile) until N samples are filled.
//   occ_confusion_counts : calculate_confusion_matrix (/root/reference/evaluate_metrics.py:23-32): tp = #(target > thr),
//                      tn = #(nontarget <= thr), fp = #(nontarget > thr), fn = #(target <= thr), float64 compares.
// Both are HBM-bound streaming kernels: 2 B read + 4 B written per sample; 8 B read per score.
#include "common.cuh"
#include "kernels.h"

namespace occ {

__global__ void __launch_bounds__(256)
ingest_pcm16_kernel(const int16_t* __restrict__ pcm, const long long* __restrict__ offsets, const int* __restrict__ lens,
                    int N, int mode, float* __restrict__ wav) {
    const int b = blockIdx.y;
    const long long off = offsets[b];
    const int len = lens[b];
    float* o = wav + (long long)b * N;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
        float v = 0.f;
        if (len > 0) {
            if (i < len) v = (float)pcm[off + i] * (1.0f / 32768.0f);
            else if (mode == 1) v = (float)pcm[off + (i % len)] * (1.0f / 32768.0f);
        }
        o[i] = v;
    }
}
cudaError_t launch_ingest_pcm16(const int16_t* pcm, const long long* offsets, const int* lens, int B, int N, int mode, float* wav, cudaStream_t s) {
    if (B <= 0 || N <= 0) return cudaSuccess;
    dim3 grid((N + 256 * 8 - 1) / (256 * 8), B);
    ingest_pcm16_kernel<<<grid, 256, 0, s>>>(pcm, offsets, lens, N, mode, wav);
    return cudaGetLastError();
}

// counts[0..3] = tp, tn, fp, fn (unsigned long long, zeroed by the launcher)
__global__ void __launch_bounds__(256)
confusion_kernel(const double* __restrict__ target, long long nt, const double* __restrict__ nontarget, long long nn, double thr,
                 unsigned long long* __restrict__ counts) {
    unsigned int tp = 0, fp = 0;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nt; i += stride) tp += target[i] > thr;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nn; i += stride) fp += nontarget[i] > thr;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { tp += __shfl_xor_sync(0xffffffffu, tp, o); fp += __shfl_xor_sync(0xffffffffu, fp, o); }
    if ((threadIdx.x & 31) == 0) {
        if (tp) atomicAdd(counts + 0, (unsigned long long)tp);
        if (fp) atomicAdd(counts + 2, (unsigned long long)fp);
    }
}
__global__ void confusion_finish_kernel(long long nt, long long nn, unsigned long long* counts) {
    counts[1] = (unsigned long long)nn - counts[2];      // tn = nontarget <= thr  (NaN scores compare false on both sides in
    counts[3] = (unsigned long long)nt - counts[0];      // numpy as well only for '>' ; the reference never produces NaN scores)
}
cudaError_t launch_confusion(const double* target, long long nt, const double* nontarget, long long nn, double thr,
                             unsigned long long* counts, cudaStream_t s) {
    cudaError_t e = cudaMemsetAsync(counts, 0, 4 * sizeof(unsigned long long), s);
    if (e != cudaSuccess) return e;
    const long long n = nt > nn ? nt : nn;
    int blocks = (int)((n + 255) / 256);
    if (blocks > 148 * 4) blocks = 148 * 4;
    if (blocks < 1) blocks = 1;
    confusion_kernel<<<blocks, 256, 0, s>>>(target, nt, nontarget, nn, thr, counts);
    confusion_finish_kernel<<<1, 1, 0, s>>>(nt, nn, counts);
    return cudaGetLastError();
}

}  // namespace occ
