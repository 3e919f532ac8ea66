#!/bin/bash
# 2-GPU visit: world-of-one p2p test, EP tests (NCCL + peer-memory layouts), bench N=2 both layouts.
mkdir -p gpurun_out
echo "== world-1 p2p"; timeout -k 10 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout=150 -k "peer_memory" 2>&1 | tail -5
echo "== EP tests"; timeout -k 10 600 python -m pytest tests/test_gpu_ep.py -m gpu -q -x --timeout=500 > gpurun_out/pytest_ep.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/pytest_ep.log
for lay in p2p allreduce; do
  echo "== bench N=2 $lay"
  timeout -k 10 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
     bench.py --gpus 2 --steps 200 --warmup 20 --ep-layout $lay --no-cpu-baseline > gpurun_out/bench_n2_$lay.json 2> gpurun_out/bench_n2_$lay.err
  echo "rc=$?"; cut -c1-700 gpurun_out/bench_n2_$lay.json; tail -3 gpurun_out/bench_n2_$lay.err
done
