#!/bin/bash
# round 2, visit x (1 GPU): final regression of the round-2 tree: smoke, the whole GPU suite, the default bench
# (what the driver runs), the ncu launch list of the same command.
mkdir -p gpurun_out
timeout -k 10 300 python __graft_entry__.py smoke > gpurun_out/r2x_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r2x_smoke.log
timeout -k 10 900 python -m pytest tests -m gpu -q --timeout=600 -p no:cacheprovider > gpurun_out/r2x_pytest_gpu.log 2>&1; echo "full suite rc=$?"; tail -6 gpurun_out/r2x_pytest_gpu.log
timeout -k 10 600 python bench.py > gpurun_out/r2x_bench_T32.json 2> gpurun_out/r2x_bench_T32.err; echo "bench rc=$?"; python - <<PY
import json
d = json.loads(open("gpurun_out/r2x_bench_T32.json").read().strip().splitlines()[-1])
print({k: d.get(k) for k in ("metric", "value", "ms_per_step", "e2e", "gpu_launches", "clocks")}); print(d["roofline"].get("frac"), d["roofline"].get("frac_of_read_stream"), d["cpu_baseline"]["value"], d["cpu_baseline"]["kind"])
PY
ARGS="--steps 2 --warmup 3 --no-cpu-baseline"
timeout -k 10 200 python bench.py $ARGS > gpurun_out/r2x_plain.json 2> gpurun_out/r2x_plain.err &&
timeout -k 10 400 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -c 400 --csv \
    --log-file gpurun_out/launches_r2x_T32.csv python bench.py $ARGS > gpurun_out/r2x_ncu_launch.log 2>&1
echo "launch-list rc=$?"; tail -4 gpurun_out/launches_r2x_T32.csv | cut -c1-300
