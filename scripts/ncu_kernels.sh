#!/bin/bash
# Per-kernel ncu metrics of our kernels only, at one T.  Usage: ncu_kernels.sh <tag> <T> [--fp8]
TAG=$1; shift
mkdir -p gpurun_out
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,lts__t_sector_hit_rate.pct,launch__grid_size
python scripts/one_forward.py "$@" > gpurun_out/one_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/one_$TAG.log; exit 1; }
ncu --metrics $M --clock-control none --kernel-name-base demangled -k regex:glm4moe -s ${NCU_SKIP:-6} -c ${NCU_COUNT:-12} --csv \
    --log-file gpurun_out/kern_$TAG.csv python scripts/one_forward.py "$@" > gpurun_out/ncu_kern_$TAG.log 2>&1
echo "ncu_kernels $TAG rc=$?"
