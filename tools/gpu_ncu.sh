#!/bin/bash
# usage: tools/gpu_ncu.sh <kernel-regex> <out-name>   (run under gpurun)
ARGS="--points 296 --steps 2 --warmup 1 --no-cpu-baseline"
python bench.py $ARGS > gpurun_out/plain_$2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$1 -s 1 -c 1 -o gpurun_out/prof_$2 python bench.py $ARGS > gpurun_out/ncu_$2.log 2>&1
echo "ncu exit=$?"; tail -2 gpurun_out/plain_$2.log | cut -c1-600
