#!/bin/bash
# round 2, visit z (2 GPUs): the final all-reduce kernels inside the reference engine (CUDA-graph decode)
N=${1:-2}
mkdir -p gpurun_out
run() { tag=$1; shift; echo "== $tag: $*"; timeout -k 15 ${LIMIT:-200} python scripts/engine_run.py "$@" > gpurun_out/r2z_$tag.json 2> gpurun_out/r2z_$tag.err; rc=$?; echo "rc=$rc"; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2z_$tag.json").read().strip().splitlines()[-1])
    print({k: v for k, v in d.items() if k not in ("token_ids", "what")})
except Exception as e:
    print("no result:", e); print(open("gpurun_out/r2z_$tag.err").read()[-1800:])
PY
}
run small_tp${N}_fused --preset small --layers 4 --seqs 8 --tp $N --ep p2p --fused-norm
LIMIT=300 run air46_fp8_tp${N}_fused --preset air --layers 46 --weights fp8 --seqs 32 --max-tokens 32 --tp $N --ep p2p --max-batched-tokens 2048 --mem 0.9 --fused-norm --dense ours
