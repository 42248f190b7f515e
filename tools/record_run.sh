#!/bin/bash
# round-2 GPU call 11: the record run -- parity suite, bench lines (bf16 / fp32 / reference arm), configs 4 / 5, reference eager bar,
# launch list + ncu captures (one transformer layer's GEMMs, long attention, HS-GAL)
mkdir -p gpurun_out; rm -f gpurun_out/r2_parity_stats.jsonl
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/t_all.log 2>&1; tail -4 gpurun_out/t_all.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_B64_bf16.json 2> gpurun_out/bench_B64_bf16.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
timeout 400 python bench.py --dtype fp32 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_B64_fp32.json 2> gpurun_out/bench_B64_fp32.err
timeout 400 python bench.py --dtype fp32 --batch 8 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_B8_fp32.json 2> gpurun_out/bench_B8_fp32.err
timeout 400 python tools/score_config4.py --utterances 2048 > gpurun_out/config4.json 2> gpurun_out/config4.err
timeout 400 python tools/extract_config5.py > gpurun_out/config5.json 2> gpurun_out/config5.err
timeout 500 python tools/ref_eager_gpu.py > gpurun_out/ref_eager.json 2> gpurun_out/ref_eager.err
python tools/prof_step.py > gpurun_out/prof_step_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python tools/prof_step.py > gpurun_out/ncu_launches.log 2>&1
python tools/prof_step.py > gpurun_out/prof_step_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:gemm_tc_kernel -s 12 -c 4 -f -o gpurun_out/layer_gemms python tools/prof_step.py > gpurun_out/ncu_layer.log 2>&1
python tools/prof_step.py > gpurun_out/prof_step_plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:gat_kernel -c 3 -f -o gpurun_out/gat python tools/prof_step.py > gpurun_out/ncu_gat.log 2>&1
python tools/prof_step.py --batch 16 --samples 160000 > gpurun_out/prof_step_plain4.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:attn_tc3 -c 1 -f -o gpurun_out/attn_tc3 python tools/prof_step.py --batch 16 --samples 160000 > gpurun_out/ncu_attn3.log 2>&1
ls -la gpurun_out | head -40
