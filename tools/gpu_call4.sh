#!/bin/bash
# round-2 GPU call 4: ncu of the conv-encoder GEMMs (fused LN epilogue and plain)
mkdir -p gpurun_out
OCC_CONV_LN_FUSED=1 python tools/prof_conv.py > gpurun_out/prof_conv_plain.log 2>&1 && \
OCC_CONV_LN_FUSED=1 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:gemm_tc_kernel -c 2 -f -o gpurun_out/conv_fused python tools/prof_conv.py > gpurun_out/ncu_conv_fused.log 2>&1
python tools/prof_conv.py > gpurun_out/prof_conv_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:gemm_tc_kernel -c 1 -f -o gpurun_out/conv_plain python tools/prof_conv.py > gpurun_out/ncu_conv_plain.log 2>&1
tail -3 gpurun_out/ncu_conv_fused.log gpurun_out/ncu_conv_plain.log
ls -la gpurun_out
