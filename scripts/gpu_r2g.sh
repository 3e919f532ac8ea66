#!/bin/bash
# round 2, visit g (2 GPUs): the reference engine at tensor_parallel_size=2 with expert-parallel sparse blocks
mkdir -p gpurun_out
run() { tag=$1; shift; echo "== $tag: $*"; timeout -k 15 ${LIMIT:-240} python scripts/engine_run.py "$@" > gpurun_out/r2g_$tag.json 2> gpurun_out/r2g_$tag.err; rc=$?; echo "rc=$rc"; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2g_$tag.json").read().strip().splitlines()[-1])
    print({k: v for k, v in d.items() if k not in ("token_ids", "what")})
except Exception as e:
    print("no result:", e); print(open("gpurun_out/r2g_$tag.err").read()[-1800:])
PY
}
run small_tp1 --preset small --layers 4 --seqs 8
run small_tp2_p2p_eager --preset small --layers 4 --seqs 8 --tp 2 --ep p2p --eager
run small_tp2_p2p --preset small --layers 4 --seqs 8 --tp 2 --ep p2p
run small_tp2_allreduce --preset small --layers 4 --seqs 8 --tp 2 --ep allreduce
LIMIT=600 run air46_fp8_tp2 --preset air --layers 46 --weights fp8 --seqs 32 --max-tokens 32 --tp 2 --ep p2p --max-batched-tokens 2048 --mem 0.9
LIMIT=600 run air46_bf16_tp2 --preset air --layers 46 --weights bf16 --seqs 32 --max-tokens 32 --tp 2 --ep p2p --max-batched-tokens 2048 --mem 0.9
nvidia-smi --query-gpu=memory.used --format=csv | tail -2
