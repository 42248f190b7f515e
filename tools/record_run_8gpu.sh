#!/bin/bash
# 8-GPU box: config 3 at N = 1, 2, 4, 8 (bf16) and N = 8 (fp32 mode) on ONE box; bench at N = 8
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 900 $TR --nproc-per-node 8 --master-port 29621 tools/score_config3.py > gpurun_out/config3_N8.json 2> gpurun_out/config3_N8.err
timeout 900 $TR --nproc-per-node 4 --master-port 29622 tools/score_config3.py > gpurun_out/config3_N4.json 2> gpurun_out/config3_N4.err
timeout 900 $TR --nproc-per-node 2 --master-port 29623 tools/score_config3.py > gpurun_out/config3_N2.json 2> gpurun_out/config3_N2.err
timeout 900 python tools/score_config3.py > gpurun_out/config3_N1.json 2> gpurun_out/config3_N1.err
timeout 900 $TR --nproc-per-node 8 --master-port 29624 tools/score_config3.py --dtype fp32 > gpurun_out/config3_N8_fp32.json 2> gpurun_out/config3_N8_fp32.err
timeout 600 $TR --nproc-per-node 8 --master-port 29625 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench_N8.json 2> gpurun_out/bench_N8.err
tail -q -n 1 gpurun_out/config3_N*.json | cut -c1-400
tail -n 2 gpurun_out/*.err
