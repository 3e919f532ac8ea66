#!/bin/bash
# r1d 4/8-GPU visit, trimmed to what the GPU-minute budget allows: EP test (pull + sliced), config 5 benches.
N=${1:-8}; TAG=${2:-r1d}
mkdir -p gpurun_out
run() {
  local name=$1; shift
  echo "== bench N=$N $name"
  timeout -k 10 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
     bench.py --gpus $N --no-cpu-baseline "$@" > gpurun_out/bench_${TAG}_n${N}_$name.json 2> gpurun_out/bench_${TAG}_n${N}_$name.err
  echo "rc=$?"; grep '^{' gpurun_out/bench_${TAG}_n${N}_$name.json | cut -c1-200; tail -2 gpurun_out/bench_${TAG}_n${N}_$name.err | cut -c1-300
}
if [ -z "$SKIP_TESTS" ]; then
echo "== EP tests world=$N"; GLM4MOE_TEST_WORLD=$N timeout -k 10 300 python -m pytest tests/test_gpu_ep.py -m gpu -q -x --timeout=280 > gpurun_out/pytest_ep${N}_$TAG.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/pytest_ep${N}_$TAG.log
fi
run glm45_T8192_p2p_sliced --shape glm-4.5 --tokens 8192 --steps 20 --warmup 3 --ep-layout p2p --ep-sliced 1
run glm45_T8192_p2p_pull   --shape glm-4.5 --tokens 8192 --steps 20 --warmup 3 --ep-layout p2p --ep-sliced 0
run glm45_T8192_allreduce  --shape glm-4.5 --tokens 8192 --steps 20 --warmup 3 --ep-layout allreduce
run glm45_T64_p2p          --shape glm-4.5 --tokens 64 --steps 200 --warmup 20 --ep-layout p2p
run glm45_T64_allreduce    --shape glm-4.5 --tokens 64 --steps 200 --warmup 20 --ep-layout allreduce
