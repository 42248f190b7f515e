#!/bin/bash
# "r4" record run on the final code: GPU suite, bench lines (bf16, reference arm), conv0 A/B, launch list, ncu captures of one
# transformer layer's GEMMs + attention and of conv0.  Outputs land in gpurun_out/ and are copied to profiles/r4_*.
mkdir -p gpurun_out; rm -f gpurun_out/r2_parity_stats.jsonl
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r4_t_all.log 2>&1; tail -4 gpurun_out/r4_t_all.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r4_bench_B64_bf16.json 2> gpurun_out/r4_bench_B64_bf16.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r4_bench_reference.json 2> gpurun_out/r4_bench_reference.err
OCC_CONV0_MMA=0 timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r4_bench_B64_bf16_conv0simt.json 2> /dev/null
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r4_bench_B64_bf16_again.json 2> /dev/null
python tools/prof_step.py > gpurun_out/prof_step_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r4_launches.csv python tools/prof_step.py > gpurun_out/ncu_launches.log 2>&1
python tools/prof_step.py > gpurun_out/prof_step_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'gemm_tc_kernel|attn_tc2' -s 15 -c 5 -f -o gpurun_out/r4_layer python tools/prof_step.py > gpurun_out/ncu_layer.log 2>&1
python tools/prof_step.py > gpurun_out/prof_step_plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'conv0' -c 1 -f -o gpurun_out/r4_conv0 python tools/prof_step.py > gpurun_out/ncu_conv0.log 2>&1
ls -la gpurun_out | head -40
