#!/bin/bash
# FP8 prefill kernel bring-up: fp8 tests under tight timeouts, then stage timings.
mkdir -p gpurun_out
echo "== pytest fp8"; timeout -k 10 ${PYTEST_LIMIT:-420} python -m pytest tests -m gpu -q -x --timeout=150 -k "${PYTEST_K:-fp8}" > gpurun_out/pytest_fp8.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_fp8.log
echo "== stages fp8"; timeout -k 10 200 python scripts/stage_probe.py ${STAGE_TS:-256 1024 4096 8192} --fp8 2>&1 | tail -6
echo "== stages bf16"; timeout -k 10 200 python scripts/stage_probe.py ${STAGE_TS_BF16:-64 256 1024 4096} 2>&1 | tail -6
