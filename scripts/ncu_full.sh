#!/bin/bash
# One `ncu --set full` capture of a kernel of one_forward.py.  Usage: ncu_full.sh <tag> <kernel-regex> <skip> <one_forward args...>
TAG=$1; KREGEX=$2; SKIP=$3; shift 3
mkdir -p gpurun_out
timeout -k 10 120 python scripts/one_forward.py "$@" > gpurun_out/one_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/one_$TAG.log; exit 1; }
timeout -k 10 300 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:$KREGEX -s $SKIP -c 1 \
    -f -o gpurun_out/full_$TAG python scripts/one_forward.py "$@" > gpurun_out/ncu_full_$TAG.log 2>&1
echo "ncu_full $TAG rc=$?"; ls -la gpurun_out/full_$TAG.ncu-rep
