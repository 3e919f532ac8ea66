#!/bin/bash
# N-GPU visit (N = $1): EP tests over N ranks, then bench at N with the peer-memory and the NCCL layout.
N=${1:-8}; LAYOUTS=${2:-"p2p allreduce"}
mkdir -p gpurun_out
if [ -z "$SKIP_TESTS" ]; then
echo "== EP tests world=$N"; GLM4MOE_TEST_WORLD=$N timeout -k 10 400 python -m pytest tests/test_gpu_ep.py -m gpu -q -x --timeout=380 > gpurun_out/pytest_ep$N.log 2>&1; echo "rc=$?"; tail -8 gpurun_out/pytest_ep$N.log
fi
for lay in $LAYOUTS; do
  echo "== bench N=$N $lay"
  timeout -k 10 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
     bench.py --gpus $N --steps 200 --warmup 20 --ep-layout $lay --no-cpu-baseline > gpurun_out/bench_n${N}_$lay.json 2> gpurun_out/bench_n${N}_$lay.err
  echo "rc=$?"; grep '^{' gpurun_out/bench_n${N}_$lay.json | cut -c1-230; tail -2 gpurun_out/bench_n${N}_$lay.err | cut -c1-300
done
