#!/bin/bash
# r1d single-GPU visit: regression (smoke, gpu tests), bench lines (Air T=32 default, GLM-4.5 shape
# decode 64 / prefill 8192 = BASELINE configs[4] on one GPU), stage timings incl. T=8192 and the
# 128-row tile experiment, bulk-copy / TMEM microbenchmarks.  Every step under its own timeout.
TAG=${1:-r1d}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
echo "== smoke"; timeout -k 10 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
echo "== pytest"; timeout -k 10 900 python -m pytest tests -m gpu -q -x --timeout=600 --durations=8 > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -22 gpurun_out/pytest_$TAG.log
echo "== bench"; timeout -k 10 600 python bench.py --sweep > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/bench_$TAG.json; tail -3 gpurun_out/bench_$TAG.err
echo "== bench reference"; timeout -k 10 400 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref_$TAG.json 2>> gpurun_out/bench_$TAG.err; echo "rc=$?"; cat gpurun_out/bench_ref_$TAG.json
echo "== bench prefill T=4096 (Air)"; timeout -k 10 400 python bench.py --tokens 4096 --steps 40 --warmup 5 --cpu-steps 1 > gpurun_out/bench_${TAG}_T4096.json 2> gpurun_out/bench_${TAG}_T4096.err; echo "rc=$?"; tail -c 2200 gpurun_out/bench_${TAG}_T4096.json; tail -2 gpurun_out/bench_${TAG}_T4096.err
echo "== stages bf16"; timeout -k 10 300 python scripts/stage_probe.py 1 32 256 4096 8192 > gpurun_out/stages_$TAG.jsonl 2> gpurun_out/stages_$TAG.err; cat gpurun_out/stages_$TAG.jsonl; tail -2 gpurun_out/stages_$TAG.err
echo "== stages bf16 BN=128"; GLM4MOE_FORCE_BN=128 timeout -k 10 300 python scripts/stage_probe.py 4096 8192 > gpurun_out/stages_${TAG}_bn128.jsonl 2> gpurun_out/stages_${TAG}_bn128.err; cat gpurun_out/stages_${TAG}_bn128.jsonl; tail -2 gpurun_out/stages_${TAG}_bn128.err
echo "== stages fp8"; timeout -k 10 300 python scripts/stage_probe.py 1 32 4096 8192 --fp8 > gpurun_out/stages_${TAG}_fp8.jsonl 2> gpurun_out/stages_${TAG}_fp8.err; cat gpurun_out/stages_${TAG}_fp8.jsonl; tail -2 gpurun_out/stages_${TAG}_fp8.err
echo "== GLM-4.5 shape stages"; GLM4MOE_SHAPE=glm-4.5 timeout -k 10 300 python scripts/stage_probe.py 1 32 64 256 8192 > gpurun_out/stages_${TAG}_glm45.jsonl 2> gpurun_out/stages_${TAG}_glm45.err; cat gpurun_out/stages_${TAG}_glm45.jsonl; tail -2 gpurun_out/stages_${TAG}_glm45.err
for T in 64 8192; do
  echo "== bench GLM-4.5 shape T=$T"; timeout -k 10 400 python bench.py --shape glm-4.5 --tokens $T --steps $([ $T -gt 256 ] && echo 30 || echo 200) --warmup 5 --cpu-steps 1 > gpurun_out/bench_${TAG}_glm45_T$T.json 2> gpurun_out/bench_${TAG}_glm45_T$T.err; echo "rc=$?"; tail -c 2200 gpurun_out/bench_${TAG}_glm45_T$T.json; tail -2 gpurun_out/bench_${TAG}_glm45_T$T.err
done
echo "== ubench"; timeout -k 10 120 scripts/ubench/ubench > gpurun_out/ubench_$TAG.jsonl 2>&1; echo "rc=$?"; cat gpurun_out/ubench_$TAG.jsonl
ls -la gpurun_out | tail -30
