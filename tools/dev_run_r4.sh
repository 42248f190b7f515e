#!/bin/bash
# r4 development run: GPU suite on the new conv0 (mma.sync) + one-MUFU GELU, then same-box A/B bench lines:
#   default | OCC_CONV0_MMA=0 (previous conv0) | library built with -DOCC_GELU_V1 (previous GELU in the epilogues) | default again
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/r4_t_all.log 2>&1; tail -5 gpurun_out/r4_t_all.log
timeout 400 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r4_bench_new.json 2> gpurun_out/r4_bench_new.err
OCC_CONV0_MMA=0 timeout 400 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r4_bench_conv0old.json 2> /dev/null
if [ -f build_v1/libocc_b200_geluv1.so ]; then
  cp occm_b200/libocc_b200.so /tmp/lib_new.so && cp build_v1/libocc_b200_geluv1.so occm_b200/libocc_b200.so
  timeout 400 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r4_bench_geluv1.json 2> /dev/null
  cp /tmp/lib_new.so occm_b200/libocc_b200.so
fi
timeout 400 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r4_bench_new2.json 2> /dev/null
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r4_bench_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        fm = d["roofline"]["family_ms_per_step"]
        print(f, round(d["ms_per_step"], 3), round(d["value"], 1), d["clocks"]["sm_mhz"], {k: fm[k] for k in ("conv0", "conv_gemm", "xf_fc1", "xf_qkv", "xf_gemm", "pos_gemm", "norm")})
    except Exception as e:
        print(f, "ERR", e)
PY
