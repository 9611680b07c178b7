#!/bin/bash
# usage: tools/gpu_prof.sh <kernel-regex> <config-substring> <outname>
# ncu --set full of one kernel on one config of tools/bench_configs.py (plain run first, same args).
set -u
mkdir -p gpurun_out
timeout 300 python tools/bench_configs.py 3 "$2" > gpurun_out/plain_$3.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$1 -s 6 -c 1 \
    -o gpurun_out/prof_$3 -f python tools/bench_configs.py 3 "$2" > gpurun_out/ncu_$3.log 2>&1
echo "ncu $3 exit $?"
