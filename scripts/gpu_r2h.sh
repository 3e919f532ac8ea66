#!/bin/bash
# round 2, visit h (N GPUs, N = $1): expert-parallel scaling of the layer bench + the reference engine at tensor_parallel_size=N
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
show() { python - <<PY
import json
try:
    d = json.loads(open("$1").read().strip().splitlines()[-1])
    print({k: d.get(k) for k in ("value", "ms_per_step", "ms_per_step_median", "ep_bit_identical")}, d["e2e"]["value"], d["details"].get("ep_load"), d["details"]["parallelism"])
except Exception as e:
    print("failed:", e)
PY
}
echo "== air T=32 p2p"; timeout -k 15 300 $TR --master-port 29601 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2h_n${N}_air_T32_p2p.json 2> gpurun_out/r2h_n${N}_air_T32_p2p.err; show gpurun_out/r2h_n${N}_air_T32_p2p.json
echo "== air T=32 p2p, 200 steps"; timeout -k 15 300 $TR --master-port 29602 bench.py --gpus $N --steps 200 --warmup 10 --no-ep-check > gpurun_out/r2h_n${N}_air_T32_p2p_200.json 2> gpurun_out/r2h_n${N}_air_T32_p2p_200.err; show gpurun_out/r2h_n${N}_air_T32_p2p_200.json
echo "== air T=32 allreduce"; timeout -k 15 300 $TR --master-port 29603 bench.py --gpus $N --steps 20 --warmup 5 --ep-layout allreduce > gpurun_out/r2h_n${N}_air_T32_allreduce.json 2> gpurun_out/r2h_n${N}_air_T32_allreduce.err; show gpurun_out/r2h_n${N}_air_T32_allreduce.json
echo "== glm-4.5 T=64 p2p"; timeout -k 15 300 $TR --master-port 29604 bench.py --gpus $N --steps 20 --warmup 5 --shape glm-4.5 --tokens 64 > gpurun_out/r2h_n${N}_glm45_T64_p2p.json 2> gpurun_out/r2h_n${N}_glm45_T64_p2p.err; show gpurun_out/r2h_n${N}_glm45_T64_p2p.json
echo "== glm-4.5 T=8192 p2p"; timeout -k 15 300 $TR --master-port 29605 bench.py --gpus $N --steps 10 --warmup 3 --shape glm-4.5 --tokens 8192 > gpurun_out/r2h_n${N}_glm45_T8192_p2p.json 2> gpurun_out/r2h_n${N}_glm45_T8192_p2p.err; show gpurun_out/r2h_n${N}_glm45_T8192_p2p.json
run() { tag=$1; shift; echo "== $tag: $*"; timeout -k 15 ${LIMIT:-400} python scripts/engine_run.py "$@" > gpurun_out/r2h_n${N}_$tag.json 2> gpurun_out/r2h_n${N}_$tag.err; echo "rc=$?"; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2h_n${N}_$tag.json").read().strip().splitlines()[-1])
    print({k: v for k, v in d.items() if k not in ("token_ids", "what")})
except Exception as e:
    print("no result:", e); print(open("gpurun_out/r2h_n${N}_$tag.err").read()[-1500:])
PY
}
run air46_bf16 --preset air --layers 46 --weights bf16 --seqs 32 --max-tokens 32 --tp $N --ep p2p --max-batched-tokens 2048 --mem 0.9
run air46_fp8 --preset air --layers 46 --weights fp8 --seqs 32 --max-tokens 32 --tp $N --ep p2p --max-batched-tokens 2048 --mem 0.9
