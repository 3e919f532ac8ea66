#!/bin/bash
# round 2, visit v (N GPUs, default 8): the fused all-reduce + norm at N ranks (tests, timings against NCCL),
# the layer bench at N, the reference engine at tensor_parallel_size=N with the fused front end.
N=${1:-8}
mkdir -p gpurun_out
echo "== front tests world=$N"; GLM4MOE_TEST_WORLD=$N timeout -k 10 300 python -m pytest tests/test_gpu_ep.py -m gpu -q -x --timeout=280 -p no:cacheprovider -k fused_allreduce > gpurun_out/r2v_pytest_front$N.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/r2v_pytest_front$N.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
echo "== front end only"; timeout -k 10 200 $TR --master-port 29517 scripts/front_probe.py --no-block --tokens 1,32,128,1024,8192 --reps 9 > gpurun_out/r2v_front_only_n$N.jsonl 2> gpurun_out/r2v_front_only_n$N.err; echo "rc=$?"; cat gpurun_out/r2v_front_only_n$N.jsonl; tail -3 gpurun_out/r2v_front_only_n$N.err | cut -c1-400
echo "== air T=32 p2p"; timeout -k 15 200 $TR --master-port 29601 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2v_n${N}_air_T32_p2p.json 2> gpurun_out/r2v_n${N}_air_T32_p2p.err; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2v_n${N}_air_T32_p2p.json").read().strip().splitlines()[-1])
    print({k: d.get(k) for k in ("value", "ms_per_step", "ms_per_step_median", "ep_bit_identical")}, d["e2e"]["value"], d["details"].get("ep_load"))
except Exception as e:
    print("failed:", e); print(open("gpurun_out/r2v_n${N}_air_T32_p2p.err").read()[-1500:])
PY
run() { tag=$1; shift; echo "== $tag: $*"; timeout -k 15 ${LIMIT:-300} python scripts/engine_run.py "$@" > gpurun_out/r2v_n${N}_$tag.json 2> gpurun_out/r2v_n${N}_$tag.err; echo "rc=$?"; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2v_n${N}_$tag.json").read().strip().splitlines()[-1])
    print({k: v for k, v in d.items() if k not in ("token_ids", "what")})
except Exception as e:
    print("no result:", e); print(open("gpurun_out/r2v_n${N}_$tag.err").read()[-1500:])
PY
}
run air46_${WEIGHTS:-bf16}_fused --preset air --layers 46 --weights ${WEIGHTS:-bf16} --seqs 32 --max-tokens 32 --tp $N --ep p2p --max-batched-tokens 2048 --mem 0.9 --fused-norm
if [ -n "$ALSO_REF" ]; then
run air46_${WEIGHTS:-bf16}_ref --preset air --layers 46 --weights ${WEIGHTS:-bf16} --seqs 32 --max-tokens 32 --tp $N --ep p2p --max-batched-tokens 2048 --mem 0.9
fi
