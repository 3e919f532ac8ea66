#!/bin/bash
# round 2, visit p: is the T=4096 layer at the board power cap?
mkdir -p gpurun_out
timeout -k 10 400 python scripts/power_probe.py > gpurun_out/r2p_power.jsonl 2> gpurun_out/r2p_power.err; echo "rc=$?"; cat gpurun_out/r2p_power.jsonl; tail -5 gpurun_out/r2p_power.err
