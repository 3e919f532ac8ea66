#!/bin/bash
# Full GPU visit: smoke -> gpu tests -> bench (default args) -> stage timings (bf16 + fp8) ->
# ncu launch list of the bench command -> ncu full captures (decode T=1/32 fused kernel, prefill gate_up).
# Every step under its own timeout; ncu only after the same command exited 0 without it.
TAG=${1:-r1b}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
echo "== smoke"; timeout -k 10 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
if [ -z "$SKIP_PYTEST" ]; then
echo "== pytest"; timeout -k 10 1500 python -m pytest tests -m gpu -q -x --timeout=600 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest.log
fi
echo "== bench"; timeout -k 10 900 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; tail -c 2500 gpurun_out/bench_$TAG.json; tail -3 gpurun_out/bench_$TAG.err
echo "== bench reference"; timeout -k 10 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref_$TAG.json 2>> gpurun_out/bench_$TAG.err; echo "rc=$?"; cat gpurun_out/bench_ref_$TAG.json
for mode in "" "--fp8"; do
  echo "== stages $mode"; timeout -k 10 300 python scripts/stage_probe.py ${STAGE_TS:-1 2 4 8 16 32 64 128 256 1024 4096} $mode > gpurun_out/stages_${TAG}${mode}.jsonl 2> gpurun_out/stages_${TAG}${mode}.err; cat gpurun_out/stages_${TAG}${mode}.jsonl; tail -2 gpurun_out/stages_${TAG}${mode}.err
done
if [ -z "$SKIP_NCU" ]; then
echo "== ncu launch list (bench command)"
ARGS="--steps 2 --warmup 3 --no-cpu-baseline"
timeout -k 10 300 python bench.py $ARGS > gpurun_out/plain_$TAG.json 2> gpurun_out/plain_$TAG.err &&
timeout -k 10 600 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:glm4moe -c 600 --csv \
    --log-file gpurun_out/launches_$TAG.csv python bench.py $ARGS > gpurun_out/ncu_launch_$TAG.log 2>&1
echo "launch-list rc=$?"
bash scripts/ncu_full.sh ${TAG}_T32 decode_fused 3 32
bash scripts/ncu_full.sh ${TAG}_T1 decode_fused 3 1
bash scripts/ncu_full.sh ${TAG}_T4096_gu umma_grouped 4 4096
bash scripts/ncu_full.sh ${TAG}_T4096_fp8_gu umma_fp8_prefill 2 4096 4 --fp8
NCU_SKIP=0 NCU_COUNT=40 bash scripts/ncu_kernels.sh ${TAG}_T4096 4096 3
NCU_SKIP=0 NCU_COUNT=40 bash scripts/ncu_kernels.sh ${TAG}_T4096_fp8 4096 3 --fp8
NCU_SKIP=0 NCU_COUNT=12 bash scripts/ncu_kernels.sh ${TAG}_T1 1 6
fi
ls -la gpurun_out | tail -30
