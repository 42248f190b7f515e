#!/bin/bash
# ncu --set full on one conv0 launch of the one-pass (1) and two-pass (3) mma.sync forms (after a plain run of the same command)
mkdir -p gpurun_out
for v in 3 1; do
OCC_CONV0_MMA=$v python tools/prof_step.py > gpurun_out/prof_step_plain_v$v.log 2>&1 && \
OCC_CONV0_MMA=$v ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'conv0' -c 1 -f -o gpurun_out/r4_conv0_v$v python tools/prof_step.py > gpurun_out/ncu_conv0_v$v.log 2>&1
tail -2 gpurun_out/ncu_conv0_v$v.log
done
ls -la gpurun_out
