#!/bin/bash
# bench + ncu launch list + one full capture of the dominant kernel (run under gpurun)
set -x
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench exit=$?"
tail -c 3000 gpurun_out/bench.json
ARGS="--points 296 --steps 2 --warmup 1 --no-cpu-baseline"
python bench.py $ARGS > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv python bench.py $ARGS > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches exit=$?"
python bench.py $ARGS > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:alpha -s 1 -c 1 -o gpurun_out/prof_alpha python bench.py $ARGS > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit=$?"
ls -la gpurun_out
