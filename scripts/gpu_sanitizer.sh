#!/bin/bash
# compute-sanitizer over the mini-shape GPU tests (one tool per GPU visit: $1 = memcheck | racecheck | initcheck | synccheck)
TOOL=${1:-memcheck}
mkdir -p gpurun_out
SEL='(mini or tiny or small) and not second_device and not cuda_graph and not same_gpu'
# plain run first (the sanitizer run must repeat a command that already exited 0)
timeout -k 10 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider -k "$SEL" > gpurun_out/sanitizer_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/sanitizer_plain.log; exit 0; }
tail -2 gpurun_out/sanitizer_plain.log
EXTRA=""
[ "$TOOL" = memcheck ] && EXTRA="--leak-check no --report-api-errors no"
[ "$TOOL" = initcheck ] && EXTRA="--track-unused-memory no"
timeout -k 10 ${LIMIT:-1500} compute-sanitizer --tool $TOOL $EXTRA --error-exitcode 86 --print-limit 40 --log-file gpurun_out/sanitizer_$TOOL.log \
  python -m pytest tests/test_gpu_parity.py -m gpu -q -p no:cacheprovider -k "$SEL" > gpurun_out/sanitizer_${TOOL}_pytest.log 2>&1
echo "sanitizer $TOOL rc=$?"
tail -4 gpurun_out/sanitizer_${TOOL}_pytest.log
grep -c "=========" gpurun_out/sanitizer_$TOOL.log; tail -12 gpurun_out/sanitizer_$TOOL.log
