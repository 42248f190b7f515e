#!/bin/bash
# round-2 GPU call 1: parity suite (new attention first, in its own process), bench, configs 4/5, reference eager bar
mkdir -p gpurun_out; rm -f gpurun_out/r2_parity_stats.jsonl
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/gpu.txt
timeout 600 python -m pytest tests/test_gpu_round2.py -q -k "tc3_kernel" > gpurun_out/t_tc3.log 2>&1
timeout 2000 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/t_all.log 2>&1
timeout 500 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
timeout 400 python tools/score_config4.py --utterances 1024 > gpurun_out/config4.json 2> gpurun_out/config4.err
timeout 400 python tools/extract_config5.py > gpurun_out/config5.json 2> gpurun_out/config5.err
timeout 500 python tools/ref_eager_gpu.py > gpurun_out/ref_eager.json 2> gpurun_out/ref_eager.err
timeout 300 python bench.py --dtype fp32 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_fp32.json 2> gpurun_out/bench_fp32.err
tail -3 gpurun_out/t_tc3.log; tail -5 gpurun_out/t_all.log
