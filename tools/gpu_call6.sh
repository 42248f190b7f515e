#!/bin/bash
# round-2 GPU call 6: LayerNorm fold variants, per-GEMM timing
mkdir -p gpurun_out
for v in 0 1 2 3 0 1; do
OCC_LN_FOLD=$v timeout 500 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_fold${v}_$RANDOM.json 2>> gpurun_out/bench.err
done
ls gpurun_out
