#!/bin/bash
# 8-GPU box, final code: bench at N = 8, 4, 1 on ONE box (the driver runs its own 1/2/4/8 sweep at round end)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 $TR --nproc-per-node 8 --master-port 29625 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r4_bench_N8.json 2> gpurun_out/r4_bench_N8.err
timeout 600 $TR --nproc-per-node 4 --master-port 29626 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r4_bench_N4.json 2> gpurun_out/r4_bench_N4.err
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r4_bench_N1_samebox.json 2> gpurun_out/r4_bench_N1_samebox.err
python - <<'PY'
import json
for n in ("N8", "N4", "N1_samebox"):
    try:
        d = json.loads(open(f"gpurun_out/r4_bench_{n}.json").read().strip().splitlines()[-1])
        print(n, d["n_gpus"], round(d["ms_per_step"], 3), round(d["value"], 1), round(d["e2e"]["value"], 1), d["e2e"].get("passes_ms"), d["clocks"])
    except Exception as e:
        print(n, "ERR", e)
PY
tail -n 2 gpurun_out/r4_bench_N8.err
