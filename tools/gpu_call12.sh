#!/bin/bash
# round-2 GPU call 12: Toeplitz pos-conv correctness + A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_model.py -q --timeout 600 -k "stage_entry_points or config1 or config2 or batch_invariance or config4 or ragged" > gpurun_out/t_toep.log 2>&1; tail -15 gpurun_out/t_toep.log
for v in 0 1 0 1; do
OCC_POS_TOEPLITZ=$v timeout 500 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_toep${v}_$RANDOM.json 2>> gpurun_out/bench.err
done
tail -3 gpurun_out/bench.err
