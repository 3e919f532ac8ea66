#!/bin/bash
# round 2, visit n: router front-end merged into the persistent decode kernel at T <= 2; dense MLP tile choice
mkdir -p gpurun_out
timeout -k 10 400 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout=200 -p no:cacheprovider \
  -k "merged_front or direct_mode or route_small or cuda_graph or dense_mlp" > gpurun_out/r2n_pytest.log 2>&1; rc=$?; echo "rc=$rc"; tail -15 gpurun_out/r2n_pytest.log
if [ $rc -ne 0 ]; then exit 0; fi
timeout -k 10 400 python bench.py --steps 50 --warmup 5 --sweep > gpurun_out/r2n_bench_sweep.json 2> gpurun_out/r2n_bench_sweep.err
python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2n_bench_sweep.json").read().strip().splitlines()[-1])
    print({k: d.get(k) for k in ("value", "ms_per_step", "gpu_launches", "clocks")}, d["roofline"]["frac"], d["roofline"].get("traffic"))
    for s in d["sweep"]: print(s["T"], round(s["ms"], 4), round(s["GBps"]), s.get("launches"))
except Exception as e:
    print("bench failed", e); print(open("gpurun_out/r2n_bench_sweep.err").read()[-1500:])
PY
for T in 1 2; do timeout -k 10 200 python bench.py --tokens $T --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/r2n_bench_T$T.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/r2n_bench_T$T.json').read().strip().splitlines()[-1]); print($T, d['ms_per_step'], d['ms_per_step_median'], d['gpu_launches'], d['roofline']['layer'])"; done
timeout -k 10 200 python scripts/dense_probe.py 64 128 256 512 1024 > gpurun_out/r2n_dense_probe.jsonl 2>/dev/null; cat gpurun_out/r2n_dense_probe.jsonl
timeout -k 10 900 python -m pytest tests -m gpu -q --timeout=600 -p no:cacheprovider > gpurun_out/r2n_pytest_gpu.log 2>&1; echo "full suite rc=$?"; tail -5 gpurun_out/r2n_pytest_gpu.log
