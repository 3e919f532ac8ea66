#!/bin/bash
# round 2, visit u (4 GPUs): the N = 4 point of the expert-parallel layer bench
N=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout -k 15 150 $TR --master-port 29601 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2u_n${N}_air_T32_p2p.json 2> gpurun_out/r2u_n${N}_air_T32_p2p.err; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2u_n${N}_air_T32_p2p.json").read().strip().splitlines()[-1])
    print({k: d.get(k) for k in ("value", "ms_per_step", "ms_per_step_median", "ep_bit_identical")}, d["e2e"]["value"], d["details"].get("ep_load"))
except Exception as e:
    print("failed:", e); print(open("gpurun_out/r2u_n${N}_air_T32_p2p.err").read()[-1500:])
PY
