#!/bin/bash
# round-2 GPU call 3: full suite with the final tolerances; fused conv LN v2 A/B; config 3 tool at N=1 (subset)
mkdir -p gpurun_out; rm -f gpurun_out/r2_parity_stats.jsonl
timeout 2400 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/t_all.log 2>&1
tail -8 gpurun_out/t_all.log
OCC_CONV_LN_FUSED=1 timeout 900 python -m pytest tests/test_gpu_model.py tests/test_gpu_round2.py -q -k "stage_entry_points or config2 or batch_invariance or config1 or long_utterance or config4" > gpurun_out/t_fused.log 2>&1
tail -5 gpurun_out/t_fused.log
OCC_CONV_LN_FUSED=1 python tools/bf16_err.py > gpurun_out/bf16_err_fused.txt 2>&1
for i in 1 2; do
timeout 500 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_plain$i.json 2> gpurun_out/bench_plain$i.err
OCC_CONV_LN_FUSED=1 timeout 500 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_fused$i.json 2> gpurun_out/bench_fused$i.err
done
timeout 600 python tools/score_config3.py --utterances 8192 > gpurun_out/config3_n1_8192.json 2> gpurun_out/config3.err
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
