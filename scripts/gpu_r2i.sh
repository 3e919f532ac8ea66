#!/bin/bash
# round 2, visit i: block-scaled FP8 prefill kernel with tokens as the MMA N dimension (ragged tails cost their rows)
mkdir -p gpurun_out
timeout -k 10 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout=120 -p no:cacheprovider \
  -k "pow2 or hardware_block" > gpurun_out/r2i_pytest_pow2.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/r2i_pytest_pow2.log
for T in 4096 8192 1024; do
  echo "== bench fp8_pow2 T=$T"
  timeout -k 10 300 python bench.py --weights fp8_pow2 --tokens $T --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2i_bench_fp8_pow2_T$T.json 2> gpurun_out/r2i_bench_fp8_pow2_T$T.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2i_bench_fp8_pow2_T$T.json").read().strip().splitlines()[-1])
    print({k: d.get(k) for k in ("value", "ms_per_step")}); print(d["roofline"].get("stage_ms")); print({k: d["roofline"].get(k) for k in ("kernel", "achieved", "peak", "frac", "layer_tensor")})
except Exception as e:
    print("bench failed", e); print(open("gpurun_out/r2i_bench_fp8_pow2_T$T.err").read()[-1500:])
PY
done
timeout -k 10 600 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout=300 -p no:cacheprovider -k "fp8 or glm45" > gpurun_out/r2i_pytest_fp8.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/r2i_pytest_fp8.log
python scripts/ncu_traffic.py r2i_fp8mx_T4096 "umma_fp8_mx_kernel.*bool.1" --weights fp8_pow2 --tokens 4096 2>&1 | tail -1
