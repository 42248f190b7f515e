#!/bin/bash
# r4b: conv0 variants on one box: 1 = one-pass mma (12 warps, prefetch), 3 = two-pass mma, 2 = SIMT on centred weights, 0 = previous
mkdir -p gpurun_out
OCC_CONV0_MMA=4 timeout 900 python -m pytest tests/test_gpu_model.py -m gpu -q -x --timeout 600 > gpurun_out/r4b_t_model_v3.log 2>&1; tail -3 gpurun_out/r4b_t_model_v3.log
OCC_CONV0_MMA=3 timeout 900 python -m pytest tests/test_gpu_model.py -m gpu -q -x --timeout 600 -k "stage or config1 or config2 or batch_invariance or config4" > gpurun_out/r4b_t_model_v2.log 2>&1; tail -3 gpurun_out/r4b_t_model_v2.log
for v in 1 3 4 3 4 1; do
  OCC_CONV0_MMA=$v timeout 400 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r4b_bench_v$v.json 2> /dev/null
  python - <<PY
import json
d = json.loads(open("gpurun_out/r4b_bench_v$v.json").read().strip().splitlines()[-1])
fm = d["roofline"]["family_ms_per_step"]
print("conv0 variant $v:", round(d["ms_per_step"], 3), "ms/step", round(d["value"], 1), "utt/s  conv0", fm["conv0"], "conv_gemm", fm["conv_gemm"], "clk", d["clocks"]["sm_mhz"])
PY
done
