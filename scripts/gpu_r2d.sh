#!/bin/bash
# round 2, visit d: hardware block-scaled FP8 prefill (umma_fp8_mx.cu) -- tests first under a hard limit
mkdir -p gpurun_out
echo "== pow2 mini tests"
timeout -k 10 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout=120 -p no:cacheprovider \
  -k "pow2 or hardware_block" > gpurun_out/r2d_pytest_pow2.log 2>&1; echo "rc=$?"; tail -12 gpurun_out/r2d_pytest_pow2.log
echo "== fp8 peak"
timeout -k 10 120 python scripts/fp8_peak.py gpurun_out/fp8_peak_r2.json 2>&1 | tail -2
for w in fp8 fp8_pow2; do
  echo "== bench $w T=4096"
  timeout -k 10 300 python bench.py --weights $w --tokens 4096 --steps 20 --warmup 5 > gpurun_out/r2d_bench_${w}_T4096.json 2> gpurun_out/r2d_bench_${w}_T4096.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2d_bench_${w}_T4096.json").read().strip().splitlines()[-1])
    print({k: d.get(k) for k in ("value", "ms_per_step", "e2e")}); print(d["roofline"].get("stage_ms")); print({k: d["roofline"].get(k) for k in ("kernel", "achieved", "peak", "frac", "bound")})
except Exception as e:
    print("bench failed", e); print(open("gpurun_out/r2d_bench_${w}_T4096.err").read()[-1500:])
PY
done
echo "== full fp8 + prefill tests"
timeout -k 10 900 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout=300 -p no:cacheprovider \
  -k "fp8 or prefill_tiles or glm45" > gpurun_out/r2d_pytest_fp8.log 2>&1; echo "rc=$?"; tail -12 gpurun_out/r2d_pytest_fp8.log
echo "== bench fp8_pow2 T=32 / T=8192"
timeout -k 10 200 python bench.py --weights fp8_pow2 --tokens 32 --steps 20 --warmup 5 > gpurun_out/r2d_bench_fp8_pow2_T32.json 2>/dev/null; tail -c 400 gpurun_out/r2d_bench_fp8_pow2_T32.json | head -c 400; echo
timeout -k 10 300 python bench.py --weights fp8_pow2 --tokens 8192 --steps 10 --warmup 3 > gpurun_out/r2d_bench_fp8_pow2_T8192.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/r2d_bench_fp8_pow2_T8192.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['roofline'].get('stage_ms'))"
