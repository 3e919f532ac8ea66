#!/bin/bash
# new tests (RMSNorm, GLM-4.5 shape) + layer-stack decode probe
mkdir -p gpurun_out
echo "== pytest"; timeout -k 10 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout=400 -k "rmsnorm or glm45" > gpurun_out/pytest_next.log 2>&1; echo "rc=$?"; tail -12 gpurun_out/pytest_next.log
echo "== stack fp8"; timeout -k 10 400 python scripts/stack_probe.py --fp8 1 4 32 > gpurun_out/stack_fp8.jsonl 2> gpurun_out/stack_fp8.err; cat gpurun_out/stack_fp8.jsonl; tail -3 gpurun_out/stack_fp8.err
echo "== stack bf16"; timeout -k 10 400 python scripts/stack_probe.py 1 4 32 > gpurun_out/stack_bf16.jsonl 2> gpurun_out/stack_bf16.err; cat gpurun_out/stack_bf16.jsonl; tail -3 gpurun_out/stack_bf16.err
