#!/bin/bash
# round 2, visit j: block-scaled FP8 prefill on CTA pairs (tcgen05 cta_group::2)
mkdir -p gpurun_out
timeout -k 10 240 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout=100 -p no:cacheprovider \
  -k "hardware_block" > gpurun_out/r2j_pytest_pair.log 2>&1; rc=$?; echo "rc=$rc"; tail -25 gpurun_out/r2j_pytest_pair.log
if [ $rc -ne 0 ]; then exit 0; fi
for mode in 1 2; do
for T in 4096 8192; do
  echo "== bench fp8_pow2 T=$T GLM4MOE_FP8_MX=$mode"
  GLM4MOE_FP8_MX=$mode timeout -k 10 300 python bench.py --weights fp8_pow2 --tokens $T --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2j_bench_fp8_pow2_T${T}_mx$mode.json 2> gpurun_out/r2j_bench_fp8_pow2_T${T}_mx$mode.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2j_bench_fp8_pow2_T${T}_mx$mode.json").read().strip().splitlines()[-1])
    print({k: d.get(k) for k in ("value", "ms_per_step")}); print(d["roofline"].get("stage_ms")); print({k: d["roofline"].get(k) for k in ("achieved", "frac", "layer_tensor")})
except Exception as e:
    print("bench failed", e); print(open("gpurun_out/r2j_bench_fp8_pow2_T${T}_mx$mode.err").read()[-1500:])
PY
done; done
timeout -k 10 600 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout=300 -p no:cacheprovider -k "fp8_air_prefill" > gpurun_out/r2j_pytest_fp8.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/r2j_pytest_fp8.log
