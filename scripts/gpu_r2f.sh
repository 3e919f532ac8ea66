#!/bin/bash
# round 2, visit f: ncu --set full captures of the dominant kernels on bench.py's own batch (traffic + pipe utilisation)
mkdir -p gpurun_out
python scripts/ncu_traffic.py r2f_bf16_T32 "decode_fused_kernel" --tokens 32 2>&1 | tail -2
python scripts/ncu_traffic.py r2f_fp8mx_T4096 "umma_fp8_mx_kernel.*bool.1" --weights fp8_pow2 --tokens 4096 2>&1 | tail -2
python scripts/ncu_traffic.py r2f_fp8mx_T8192 "umma_fp8_mx_kernel.*bool.1" --weights fp8_pow2 --tokens 8192 2>&1 | tail -2
python scripts/ncu_traffic.py r2f_bf16_T4096 "umma_grouped_kernel.*256, .int.1" --tokens 4096 2>&1 | tail -2
python scripts/ncu_traffic.py r2f_bf16_T1 "decode_fused_kernel" --tokens 1 2>&1 | tail -2
python scripts/ncu_traffic.py r2f_fp8_T32 "decode_fused_kernel" --weights fp8_pow2 --tokens 32 2>&1 | tail -2
ls -la gpurun_out/*.ncu-rep
