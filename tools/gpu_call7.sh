#!/bin/bash
# round-2 GPU call 7: TMA residual epilogue (MODE 5) correctness + A/B, with and without the LayerNorm fold
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_eer_kernels.py -q -k "gemm" > gpurun_out/t_gemm.log 2>&1; tail -4 gpurun_out/t_gemm.log
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/t_all.log 2>&1; tail -6 gpurun_out/t_all.log
OCC_LN_FOLD=3 timeout 1200 python -m pytest tests/test_gpu_model.py tests/test_gpu_round2.py -q --timeout 900 -k "not fp32 and not standalone" > gpurun_out/t_fold.log 2>&1; tail -4 gpurun_out/t_fold.log
for v in "0 0" "1 0" "1 3" "1 1" "0 0" "1 3"; do
set -- $v
OCC_EPI_TMA=$1 OCC_LN_FOLD=$2 timeout 500 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_tma$1_fold$2_$RANDOM.json 2>> gpurun_out/bench.err
done
OCC_LN_FOLD=3 python tools/bf16_err.py > gpurun_out/bf16_err_fold.txt 2>&1; cat gpurun_out/bf16_err_fold.txt
