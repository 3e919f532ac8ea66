#!/bin/bash
# round 2, visit e: whole GPU suite (green run for the record) + one ncu --set full of the block-scaled gate_up kernel
mkdir -p gpurun_out
timeout -k 10 1200 python -m pytest tests -m gpu -q --timeout=600 -p no:cacheprovider > gpurun_out/r2e_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r2e_pytest_gpu.log
bash scripts/ncu_full.sh r2e_mx_gu "umma_fp8_mx_kernel.*true" 2 4096 4 --fp8-pow2
bash -c 'timeout -k 10 300 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"umma_fp8_mx_kernel.*false" -s 2 -c 1 -f -o gpurun_out/full_r2e_mx_down python scripts/one_forward.py 4096 4 --fp8-pow2 > gpurun_out/ncu_full_r2e_mx_down.log 2>&1; echo "ncu down rc=$?"'
