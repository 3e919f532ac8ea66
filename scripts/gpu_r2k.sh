#!/bin/bash
# round 2, visit k: ncu --set full of the CTA-pair block-scaled kernels (gate_up + down) at T=8192
mkdir -p gpurun_out
export GLM4MOE_FP8_MX=2
python scripts/ncu_traffic.py r2k_fp8mx2_T8192_gu "umma_fp8_mx2_kernel.*bool.1" --weights fp8_pow2 --tokens 8192 2>&1 | tail -1
python scripts/ncu_traffic.py r2k_fp8mx2_T8192_down "umma_fp8_mx2_kernel.*bool.0" --weights fp8_pow2 --tokens 8192 2>&1 | tail -1
