#!/bin/bash
# round-2 GPU call 2: fixed long attention; full parity suite; fused conv LN A/B; bench; configs 4/5
mkdir -p gpurun_out; rm -f gpurun_out/r2_parity_stats.jsonl
timeout 600 python -m pytest tests/test_gpu_round2.py -q -k "tc3_kernel" > gpurun_out/t_tc3.log 2>&1
tail -3 gpurun_out/t_tc3.log
if grep -q "5 passed" gpurun_out/t_tc3.log; then
  timeout 2400 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/t_all.log 2>&1
  tail -8 gpurun_out/t_all.log
  timeout 400 python tools/score_config4.py --utterances 1024 > gpurun_out/config4.json 2> gpurun_out/config4.err
  timeout 400 python tools/extract_config5.py > gpurun_out/config5.json 2> gpurun_out/config5.err
else
  timeout 2400 python -m pytest tests -m gpu -q --timeout 900 --deselect "tests/test_gpu_eer_kernels.py::test_attention_kernels_vs_torch" -k "not tc3 and not long_utterance and not attention_paths and not config4_bucketed" > gpurun_out/t_all.log 2>&1
  tail -8 gpurun_out/t_all.log
fi
OCC_CONV_LN_FUSED=1 timeout 900 python -m pytest tests/test_gpu_model.py -q -k "stage_entry_points or config2_bf16 or batch_invariance or config1" > gpurun_out/t_fused.log 2>&1
tail -5 gpurun_out/t_fused.log
python tools/bf16_err.py > gpurun_out/bf16_err_plain.txt 2>&1
OCC_CONV_LN_FUSED=1 python tools/bf16_err.py > gpurun_out/bf16_err_fused.txt 2>&1
timeout 500 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_plain.json 2> gpurun_out/bench_plain.err
OCC_CONV_LN_FUSED=1 timeout 500 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_fused.json 2> gpurun_out/bench_fused.err
timeout 500 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_plain2.json 2> gpurun_out/bench_plain2.err
OCC_CONV_LN_FUSED=1 timeout 500 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_fused2.json 2> gpurun_out/bench_fused2.err
