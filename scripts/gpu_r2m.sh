#!/bin/bash
# round 2, visit m: dense MLP through the tensor-core grouped GEMM (tests, timing, engine with the swap)
mkdir -p gpurun_out
timeout -k 10 400 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout=200 -p no:cacheprovider \
  -k "dense_mlp or simt_and_umma or golden or edge_cases" > gpurun_out/r2m_pytest_dense.log 2>&1; echo "rc=$?"; tail -12 gpurun_out/r2m_pytest_dense.log
timeout -k 10 200 python scripts/dense_probe.py 1 32 256 4096 > gpurun_out/r2m_dense_probe.jsonl 2> gpurun_out/r2m_dense_probe.err; cat gpurun_out/r2m_dense_probe.jsonl; tail -3 gpurun_out/r2m_dense_probe.err
run() { tag=$1; shift; echo "== $tag: $*"; timeout -k 15 ${LIMIT:-300} python scripts/engine_run.py "$@" > gpurun_out/r2m_$tag.json 2> gpurun_out/r2m_$tag.err; echo "rc=$?"; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2m_$tag.json").read().strip().splitlines()[-1])
    print({k: v for k, v in d.items() if k not in ("token_ids", "what")})
except Exception as e:
    print("no result:", e); print(open("gpurun_out/r2m_$tag.err").read()[-1800:])
PY
}
run small_dense_ours_eager --preset small --layers 4 --seqs 8 --dense ours --eager
run small_dense_ours --preset small --layers 4 --seqs 8 --dense ours
run small_dense_ref --preset small --layers 4 --seqs 8 --dense reference
LIMIT=500 run air46_fp8_dense_ours --preset air --layers 46 --weights fp8 --seqs 32 --max-tokens 32 --dense ours
timeout -k 10 900 python -m pytest tests -m gpu -q --timeout=600 -p no:cacheprovider > gpurun_out/r2m_pytest_gpu.log 2>&1; echo "full suite rc=$?"; tail -5 gpurun_out/r2m_pytest_gpu.log
