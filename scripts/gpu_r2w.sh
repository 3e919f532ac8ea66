#!/bin/bash
# round 2, visit w (2 GPUs): the two-shot reduce-scatter push as flat unrolled copies: tests + timings
N=${1:-2}
mkdir -p gpurun_out
GLM4MOE_TEST_WORLD=$N timeout -k 10 300 python -m pytest tests/test_gpu_ep.py -m gpu -q -x --timeout=280 -p no:cacheprovider -k fused_allreduce > gpurun_out/r2w_pytest_front$N.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/r2w_pytest_front$N.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout -k 10 200 $TR --master-port 29517 scripts/front_probe.py --no-block --tokens 32,512,2048,8192 --reps 9 > gpurun_out/r2w_front_only_n$N.jsonl 2> gpurun_out/r2w_front_only_n$N.err; echo "rc=$?"; cat gpurun_out/r2w_front_only_n$N.jsonl; tail -3 gpurun_out/r2w_front_only_n$N.err | cut -c1-400
