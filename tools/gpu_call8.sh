#!/bin/bash
# round-2 GPU call 8: bf16 TMA-store epilogue (MODE 6) correctness + A/B; LayerNorm fold now default
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_eer_kernels.py -q -k "gemm" > gpurun_out/t_gemm.log 2>&1; tail -4 gpurun_out/t_gemm.log
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/t_all.log 2>&1; tail -6 gpurun_out/t_all.log
for v in 0 1 0 1; do
OCC_EPI_TMA16=$v timeout 500 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_tma16_${v}_$RANDOM.json 2>> gpurun_out/bench.err
done
python tools/bf16_err.py > gpurun_out/bf16_err.txt 2>&1; cat gpurun_out/bf16_err.txt
