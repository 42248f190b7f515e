#!/bin/bash
# round-2 GPU call 10: MODE 5 with 2-deep residual prefetch, MODE 6 with register double-buffering: correctness, probes, bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_eer_kernels.py -q -k "gemm" > gpurun_out/t_gemm.log 2>&1; tail -4 gpurun_out/t_gemm.log
python tools/gemm_probe.py 12864 qkv,fc1,outr,fc2r 0,4,3 > gpurun_out/gemm_probe_new.txt 2>&1; cat gpurun_out/gemm_probe_new.txt
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/t_all.log 2>&1; tail -6 gpurun_out/t_all.log
for v in 1 2 3; do
timeout 500 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_new_$v.json 2>> gpurun_out/bench.err
done
