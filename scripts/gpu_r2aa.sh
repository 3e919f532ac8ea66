#!/bin/bash
# round 2, visit aa (1 GPU): the fused layer forward against the real decoder-layer forward
mkdir -p gpurun_out
timeout -k 10 300 python -m pytest tests/test_gpu_engine.py -m gpu -q -x --timeout=280 -p no:cacheprovider -k "fused_layer or real_decoder" > gpurun_out/r2aa_pytest.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/r2aa_pytest.log
