#!/bin/bash
# GPU visit with hard per-step limits (a hung kernel must not eat the box budget): gpu tests, then stage timings.
mkdir -p gpurun_out
echo "== pytest"; timeout -k 10 ${PYTEST_LIMIT:-240} python -m pytest tests -m gpu -q -x --timeout=120 ${PYTEST_ARGS} > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest.log
for mode in "" "--fp8"; do
  echo "== stages $mode"; timeout -k 10 120 python scripts/stage_probe.py ${STAGE_TS:-1 2 4 8 16 32} $mode 2>&1 | tail -8
done
