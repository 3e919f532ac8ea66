#!/bin/bash
# round 2, visit y (1 GPU): ncu launch list of the default bench command (this repo's kernels only)
mkdir -p gpurun_out
ARGS="--steps 2 --warmup 3 --no-cpu-baseline"
timeout -k 10 200 python bench.py $ARGS > gpurun_out/r2y_plain.json 2> gpurun_out/r2y_plain.err &&
timeout -k 10 400 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:glm4moe -c 600 --csv \
    --log-file gpurun_out/launches_r2_T32.csv python bench.py $ARGS > gpurun_out/r2y_ncu_launch.log 2>&1
echo "launch-list rc=$?"; grep -c glm4moe gpurun_out/launches_r2_T32.csv; tail -3 gpurun_out/launches_r2_T32.csv | cut -c1-260
