#!/bin/bash
# r1d multi-GPU visit (N = $1): EP tests (pull + sliced peer-memory combine, NCCL layouts), then
# BASELINE configs[4] (GLM-4.5 shape, decode T=64 and prefill T=8192) and the Air prefill at N ranks.
N=${1:-2}; TAG=${2:-r1d}
mkdir -p gpurun_out
run() {  # name, args...
  local name=$1; shift
  echo "== bench N=$N $name"
  timeout -k 10 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
     bench.py --gpus $N --no-cpu-baseline "$@" > gpurun_out/bench_${TAG}_n${N}_$name.json 2> gpurun_out/bench_${TAG}_n${N}_$name.err
  echo "rc=$?"; grep '^{' gpurun_out/bench_${TAG}_n${N}_$name.json | cut -c1-260; tail -2 gpurun_out/bench_${TAG}_n${N}_$name.err | cut -c1-300
}
if [ -z "$SKIP_TESTS" ]; then
echo "== EP tests world=$N"; GLM4MOE_TEST_WORLD=$N timeout -k 10 400 python -m pytest tests/test_gpu_ep.py -m gpu -q -x --timeout=380 > gpurun_out/pytest_ep${N}_$TAG.log 2>&1; echo "rc=$?"; tail -8 gpurun_out/pytest_ep${N}_$TAG.log
fi
run glm45_T8192_p2p_sliced --shape glm-4.5 --tokens 8192 --steps 20 --warmup 3 --ep-layout p2p --ep-sliced 1
run glm45_T8192_p2p_pull   --shape glm-4.5 --tokens 8192 --steps 20 --warmup 3 --ep-layout p2p --ep-sliced 0
run glm45_T8192_allreduce  --shape glm-4.5 --tokens 8192 --steps 20 --warmup 3 --ep-layout allreduce
run glm45_T8192_alltoall   --shape glm-4.5 --tokens 8192 --steps 20 --warmup 3 --ep-layout alltoall
run glm45_T64_p2p          --shape glm-4.5 --tokens 64 --steps 200 --warmup 20 --ep-layout p2p
run glm45_T64_p2p_sliced   --shape glm-4.5 --tokens 64 --steps 200 --warmup 20 --ep-layout p2p --ep-sliced 1
run glm45_T64_allreduce    --shape glm-4.5 --tokens 64 --steps 200 --warmup 20 --ep-layout allreduce
if [ -z "$SKIP_AIR" ]; then
run air_T4096_p2p          --tokens 4096 --steps 30 --warmup 3 --ep-layout p2p
run air_T4096_allreduce    --tokens 4096 --steps 30 --warmup 3 --ep-layout allreduce
run air_T32_p2p            --tokens 32 --steps 200 --warmup 20 --ep-layout p2p
fi
ls -la gpurun_out | tail -20
