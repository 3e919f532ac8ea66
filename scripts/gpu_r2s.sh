#!/bin/bash
# round 2, visit s (1 GPU): speculative expert prefetch A/B (one EP rank's local step, single-GPU block, bench),
# re-run of the tests changed after visit q, the all-reduce kernels at world 1 after the row-flag rewrite.
mkdir -p gpurun_out
timeout -k 10 500 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout=300 -p no:cacheprovider \
  -k "prenorm or fused_allreduce or rounding_error_budget or direct_mode or peer_memory" > gpurun_out/r2s_pytest.log 2>&1; echo "tests rc=$?"; tail -6 gpurun_out/r2s_pytest.log
timeout -k 10 300 python scripts/shard_probe.py --ranks 8,4,1 --tokens 32,16 --spec 0,-1,1,2,3 > gpurun_out/r2s_shard_bf16.jsonl 2> gpurun_out/r2s_shard_bf16.err; echo "shard rc=$?"; cat gpurun_out/r2s_shard_bf16.jsonl; tail -3 gpurun_out/r2s_shard_bf16.err
timeout -k 10 300 python scripts/shard_probe.py --ranks 8,1 --tokens 32 --spec 0,-1,2,4,6 --weights fp8 > gpurun_out/r2s_shard_fp8.jsonl 2> gpurun_out/r2s_shard_fp8.err; echo "shard fp8 rc=$?"; cat gpurun_out/r2s_shard_fp8.jsonl; tail -3 gpurun_out/r2s_shard_fp8.err
for k in 0 -1; do
  timeout -k 10 300 python bench.py --steps 200 --warmup 20 --no-cpu-baseline --knob 12=$k > gpurun_out/r2s_bench_T32_spec$k.json 2> gpurun_out/r2s_bench_T32_spec$k.err; echo "bench spec=$k rc=$?"; python - <<PY
import json
d = json.loads(open("gpurun_out/r2s_bench_T32_spec$k.json").read().strip().splitlines()[-1])
print({k_: d.get(k_) for k_ in ("value", "ms_per_step", "e2e", "clocks")}, d["roofline"].get("kernel_ms"))
PY
done
