#!/bin/bash
# round 2, visit r (N GPUs, default 2): fused all-reduce + add + norm over peer memory and the one-call front end:
# tests over N ranks, front-end timings against NCCL, the reference engine at tensor_parallel_size=N with / without.
N=${1:-2}
mkdir -p gpurun_out
echo "== EP tests world=$N"; GLM4MOE_TEST_WORLD=$N timeout -k 10 500 python -m pytest tests/test_gpu_ep.py -m gpu -q -x --timeout=480 -p no:cacheprovider > gpurun_out/r2r_pytest_ep$N.log 2>&1; echo "rc=$?"; tail -12 gpurun_out/r2r_pytest_ep$N.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
echo "== front end only"; timeout -k 10 240 $TR scripts/front_probe.py --no-block --tokens 1,8,32,128,512,2048,8192 > gpurun_out/r2r_front_only_n$N.jsonl 2> gpurun_out/r2r_front_only_n$N.err; echo "rc=$?"; cat gpurun_out/r2r_front_only_n$N.jsonl; tail -3 gpurun_out/r2r_front_only_n$N.err | cut -c1-400
echo "== front end only, GLM-4.5 width"; timeout -k 10 240 $TR scripts/front_probe.py --no-block --shape glm-4.5 --tokens 64,8192 > gpurun_out/r2r_front_only_glm45_n$N.jsonl 2> gpurun_out/r2r_front_only_glm45_n$N.err; echo "rc=$?"; cat gpurun_out/r2r_front_only_glm45_n$N.jsonl; tail -3 gpurun_out/r2r_front_only_glm45_n$N.err | cut -c1-400
echo "== front end + block"; timeout -k 10 300 $TR scripts/front_probe.py --tokens 1,32,1024 > gpurun_out/r2r_front_block_n$N.jsonl 2> gpurun_out/r2r_front_block_n$N.err; echo "rc=$?"; cat gpurun_out/r2r_front_block_n$N.jsonl; tail -3 gpurun_out/r2r_front_block_n$N.err | cut -c1-400
run() { tag=$1; shift; echo "== $tag: $*"; timeout -k 15 ${LIMIT:-240} python scripts/engine_run.py "$@" > gpurun_out/r2r_$tag.json 2> gpurun_out/r2r_$tag.err; rc=$?; echo "rc=$rc"; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2r_$tag.json").read().strip().splitlines()[-1])
    print({k: v for k, v in d.items() if k not in ("token_ids", "what")})
except Exception as e:
    print("no result:", e); print(open("gpurun_out/r2r_$tag.err").read()[-1800:])
PY
}
run small_tp${N}_fused --preset small --layers 4 --seqs 8 --tp $N --ep p2p --fused-norm
run small_tp${N}_fused_eager --preset small --layers 4 --seqs 8 --tp $N --ep p2p --fused-norm --eager
run small_tp${N}_ref --preset small --layers 4 --seqs 8 --tp $N --ep p2p
if [ -z "$SKIP_AIR" ]; then
LIMIT=600 run air46_fp8_tp${N}_fused --preset air --layers 46 --weights fp8 --seqs 32 --max-tokens 32 --tp $N --ep p2p --max-batched-tokens 2048 --mem 0.9 --fused-norm
LIMIT=600 run air46_fp8_tp${N}_ref --preset air --layers 46 --weights fp8 --seqs 32 --max-tokens 32 --tp $N --ep p2p --max-batched-tokens 2048 --mem 0.9
fi
