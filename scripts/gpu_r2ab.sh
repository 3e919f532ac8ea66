#!/bin/bash
# round 2, visit ab (1 GPU): sweep T = 1..4096 on the final tree
mkdir -p gpurun_out
timeout -k 5 140 python bench.py --sweep --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/r2ab_bench_sweep.json 2> gpurun_out/r2ab_bench_sweep.err; echo "rc=$?"; python - <<PY
import json
d = json.loads(open("gpurun_out/r2ab_bench_sweep.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"])
for r in d["details"].get("sweep", d.get("sweep", [])) or []:
    print({k: r.get(k) for k in ("T", "ms", "ms_per_step", "tokens_per_s", "gbs", "achieved_gbs", "tflops")})
PY
