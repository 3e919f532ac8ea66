#!/bin/bash
# Quick GPU visit: gpu tests, then per-stage timings.
mkdir -p gpurun_out
echo "== pytest"; timeout 1500 python -m pytest tests -m gpu -q -x --timeout=600 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest.log
echo "== stages"; timeout 600 python scripts/stage_probe.py ${STAGE_TS:-1 4 32 256 1024 4096} > gpurun_out/stages_bf16.jsonl 2>gpurun_out/stages_bf16.err; cat gpurun_out/stages_bf16.jsonl; tail -3 gpurun_out/stages_bf16.err
