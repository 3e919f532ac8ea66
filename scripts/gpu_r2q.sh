#!/bin/bash
# round 2, visit q (1 GPU): the fused front end (norm + block in one call, fused all-reduce + norm kernel at world 1),
# engine with fused_norm, full GPU suite, default bench.
mkdir -p gpurun_out
timeout -k 10 500 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout=300 -p no:cacheprovider \
  -k "prenorm or fused_allreduce or rmsnorm" > gpurun_out/r2q_pytest_front.log 2>&1; echo "front tests rc=$?"; tail -12 gpurun_out/r2q_pytest_front.log
timeout -k 10 300 python scripts/front_probe.py --tokens 1,2,8,32,256,4096 > gpurun_out/r2q_front_probe.jsonl 2> gpurun_out/r2q_front_probe.err; echo "probe rc=$?"; cat gpurun_out/r2q_front_probe.jsonl; tail -3 gpurun_out/r2q_front_probe.err
run() { tag=$1; shift; echo "== $tag: $*"; timeout -k 15 ${LIMIT:-300} python scripts/engine_run.py "$@" > gpurun_out/r2q_$tag.json 2> gpurun_out/r2q_$tag.err; echo "rc=$?"; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2q_$tag.json").read().strip().splitlines()[-1])
    print({k: v for k, v in d.items() if k not in ("token_ids", "what")})
except Exception as e:
    print("no result:", e); print(open("gpurun_out/r2q_$tag.err").read()[-1800:])
PY
}
run small_ref_norm --preset small --layers 4 --seqs 8
run small_fused_norm --preset small --layers 4 --seqs 8 --fused-norm
run small_fused_norm_eager --preset small --layers 4 --seqs 8 --fused-norm --eager
LIMIT=500 run air46_fp8_fused_norm --preset air --layers 46 --weights fp8 --seqs 32 --max-tokens 32 --fused-norm
LIMIT=500 run air46_fp8_ref_norm --preset air --layers 46 --weights fp8 --seqs 32 --max-tokens 32
timeout -k 10 900 python -m pytest tests -m gpu -q --timeout=600 -p no:cacheprovider > gpurun_out/r2q_pytest_gpu.log 2>&1; echo "full suite rc=$?"; tail -6 gpurun_out/r2q_pytest_gpu.log
timeout -k 10 600 python bench.py > gpurun_out/r2q_bench_T32.json 2> gpurun_out/r2q_bench_T32.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/r2q_bench_T32.json
