// Micro-benchmark (development tool): issue rate of FFMA2 / FFMA forms on one SM sub-partition, sm_100a.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/ubench/fma_rate tools/ubench/fma_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint64_t pk(float a, float b) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) { uint64_t r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ float fma1(float a, float b, float c) { float r; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
template <int MODE, int CH>
__global__ void k(float* out, long long* cyc, int iters, float seed) {
    uint64_t p[CH], u[CH];
    float s[2 * CH], v[2 * CH];
    const float c0 = 0.999f;
    for (int i = 0; i < CH; ++i) { p[i] = pk(seed + i, seed - i); u[i] = pk(0.5f + seed * i, 0.25f); }
    for (int i = 0; i < 2 * CH; ++i) { s[i] = seed + i; v[i] = 0.5f + seed * i; }
    uint64_t c2 = pk(seed * 3.f, seed * 5.f);
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (MODE == 0) {           // FFMA2, constant (immediate) addend: p = p * u + imm
#pragma unroll
                for (int i = 0; i < CH; ++i) p[i] = fma2(p[i], u[i], pk(c0, c0));
            } else if (MODE == 1) {    // FFMA2, three register operands
#pragma unroll
                for (int i = 0; i < CH; ++i) p[i] = fma2(p[i], u[i], c2);
            } else if (MODE == 2) {    // scalar FFMA, immediate addend, 2*CH chains
#pragma unroll
                for (int i = 0; i < 2 * CH; ++i) s[i] = fma1(s[i], v[i], c0);
            } else {                   // scalar FFMA, three registers
#pragma unroll
                for (int i = 0; i < 2 * CH; ++i) s[i] = fma1(s[i], v[i], seed);
            }
        }
    }
    long long t1 = clock64();
    float acc = 0.f;
    for (int i = 0; i < CH; ++i) { float a, b; asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(p[i])); acc += a + b; }
    for (int i = 0; i < 2 * CH; ++i) acc += s[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int MODE, int CH>
void run(const char* name, int warps_per_smsp) {
    float* out; long long* cyc;
    const int threads = warps_per_smsp * 4 * 32, blocks = 148, iters = 2000;
    cudaMalloc(&out, sizeof(float) * threads * blocks); cudaMalloc(&cyc, sizeof(long long) * blocks);
    k<MODE, CH><<<blocks, threads>>>(out, cyc, iters, 1e-3f);
    k<MODE, CH><<<blocks, threads>>>(out, cyc, iters, 1e-3f);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    const double n_inst = (double)iters * 8 * (MODE < 2 ? CH : 2 * CH) * warps_per_smsp;    // warp-instructions per SMSP
    printf("%-34s chains %2d warps/SMSP %d: %.2f cycles per warp-instruction per SMSP (%.1f FMA/clk/SM)  err=%s\n", name, MODE < 2 ? CH : 2 * CH,
           warps_per_smsp, (double)h[0] / n_inst, n_inst * (MODE < 2 ? 64 : 32) * 4 / (double)h[0], cudaGetErrorString(cudaGetLastError()));
    cudaFree(out); cudaFree(cyc);
}
int main() {
    for (int w : {1, 2, 4}) {
        run<0, 8>("FFMA2 imm addend", w);
        run<1, 8>("FFMA2 3 registers", w);
        run<2, 8>("FFMA imm addend", w);
        run<3, 8>("FFMA 3 registers", w);
    }
    run<0, 4>("FFMA2 imm addend", 4);
    run<0, 2>("FFMA2 imm addend", 4);
    run<2, 4>("FFMA imm addend", 4);
    return 0;
}
