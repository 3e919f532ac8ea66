#!/bin/bash
# round 2, visit o: ragged token tiles through smaller TMA boxes (bf16 prefill); stage breakdown at T=4096
mkdir -p gpurun_out
timeout -k 10 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout=300 -p no:cacheprovider \
  -k "tail_boxes or prefill_tiles or air_prefill_properties or dense_mlp" > gpurun_out/r2o_pytest.log 2>&1; rc=$?; echo "rc=$rc"; tail -8 gpurun_out/r2o_pytest.log
if [ $rc -ne 0 ]; then exit 0; fi
for tb in 1 0; do for T in 4096 2048 8192; do
  timeout -k 10 300 python bench.py --tokens $T --steps 60 --warmup 10 --no-cpu-baseline --knob 10=$tb > gpurun_out/r2o_bench_T${T}_tail$tb.json 2>gpurun_out/r2o_err.log
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2o_bench_T${T}_tail$tb.json").read().strip().splitlines()[-1])
    print("T=$T tail=$tb", round(d["ms_per_step"], 4), round(d["ms_per_step_median"], 4), {k: round(v, 4) for k, v in d["roofline"]["stage_ms"].items()}, d["clocks"]["sm_mhz"])
except Exception as e:
    print("bench failed", e); print(open("gpurun_out/r2o_err.log").read()[-1500:])
PY
done; done
