#!/bin/bash
# 2-GPU box: config 3 at N = 1 and N = 2 (bf16), bench at N = 2 (exercises occ_allgather_scores over NCCL)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 900 python tools/score_config3.py > gpurun_out/config3_N1.json 2> gpurun_out/config3_N1.err
timeout 900 $TR --nproc-per-node 2 --master-port 29611 tools/score_config3.py > gpurun_out/config3_N2.json 2> gpurun_out/config3_N2.err
timeout 600 $TR --nproc-per-node 2 --master-port 29612 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_N2.json 2> gpurun_out/bench_N2.err
tail -2 gpurun_out/config3_N1.json gpurun_out/config3_N2.json gpurun_out/bench_N2.json | cut -c1-600
tail -3 gpurun_out/config3_N2.err gpurun_out/bench_N2.err
