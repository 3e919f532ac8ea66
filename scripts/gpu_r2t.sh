#!/bin/bash
# round 2, visit t (N GPUs, default 2): the all-reduce kernels after the row-flag / single-plane rewrite:
# tests, front-end timings against NCCL.
N=${1:-2}
mkdir -p gpurun_out
echo "== EP tests world=$N"; GLM4MOE_TEST_WORLD=$N timeout -k 10 500 python -m pytest tests/test_gpu_ep.py -m gpu -q -x --timeout=480 -p no:cacheprovider > gpurun_out/r2t_pytest_ep$N.log 2>&1; echo "rc=$?"; tail -12 gpurun_out/r2t_pytest_ep$N.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
echo "== front end only"; timeout -k 10 240 $TR scripts/front_probe.py --no-block --tokens 1,8,32,128,512,2048,8192 > gpurun_out/r2t_front_only_n$N.jsonl 2> gpurun_out/r2t_front_only_n$N.err; echo "rc=$?"; cat gpurun_out/r2t_front_only_n$N.jsonl; tail -3 gpurun_out/r2t_front_only_n$N.err | cut -c1-400
echo "== front end + block"; timeout -k 10 300 $TR scripts/front_probe.py --tokens 1,32,1024 > gpurun_out/r2t_front_block_n$N.jsonl 2> gpurun_out/r2t_front_block_n$N.err; echo "rc=$?"; cat gpurun_out/r2t_front_block_n$N.jsonl; tail -3 gpurun_out/r2t_front_block_n$N.err | cut -c1-400
