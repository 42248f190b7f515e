#!/bin/bash
# ncu --set full on one transformer layer's GEMMs + attention (after a plain run of the same command)
mkdir -p gpurun_out
python tools/prof_step.py > gpurun_out/prof_step_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'gemm_tc_kernel|attn_tc2' -s 15 -c 5 -f -o gpurun_out/r4_layer python tools/prof_step.py > gpurun_out/ncu_layer.log 2>&1
tail -2 gpurun_out/ncu_layer.log; ls -la gpurun_out | head
