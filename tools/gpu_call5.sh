#!/bin/bash
# round-2 GPU call 5: LayerNorm fold A/B (tests + bench), fused conv LN now default
mkdir -p gpurun_out
OCC_LN_FOLD=1 timeout 1200 python -m pytest tests/test_gpu_model.py tests/test_gpu_round2.py -q --timeout 900 -k "not fp32 and not standalone" > gpurun_out/t_fold.log 2>&1
tail -6 gpurun_out/t_fold.log
OCC_LN_FOLD=1 python tools/bf16_err.py > gpurun_out/bf16_err_fold.txt 2>&1
for i in 1 2; do
timeout 500 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_plain$i.json 2> gpurun_out/bench_plain$i.err
OCC_LN_FOLD=1 timeout 500 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_fold$i.json 2> gpurun_out/bench_fold$i.err
done
cat gpurun_out/bf16_err_fold.txt
