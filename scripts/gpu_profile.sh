#!/bin/bash
# ncu evidence for the bench command: launch list (per-launch device time) + one full capture of
# the dominant kernel.  Plain run first; ncu only if it exits 0.  Usage: gpu_profile.sh <tag> [bench args]
TAG=${1:-r1}; shift
ARGS="--steps 2 --warmup 3 --no-cpu-baseline $@"
mkdir -p gpurun_out
python bench.py $ARGS > gpurun_out/plain_$TAG.json 2> gpurun_out/plain_$TAG.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file gpurun_out/launches_$TAG.csv python bench.py $ARGS > gpurun_out/ncu_launch_$TAG.log 2>&1
echo "launch-list rc=$?"
python bench.py $ARGS > gpurun_out/plain2_$TAG.json 2>> gpurun_out/plain_$TAG.err &&
ncu --set full --clock-control none --import-source on -k regex:umma_grouped -s 6 -c 2 \
    -o gpurun_out/prof_$TAG python bench.py $ARGS > gpurun_out/ncu_full_$TAG.log 2>&1
echo "full-capture rc=$?"
ls -la gpurun_out | tail -12
