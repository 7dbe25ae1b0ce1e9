#!/bin/bash
# bench + ncu launch list + full captures of the three hot kernels (run under gpurun)
set -x
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench exit=$?"
tail -c 3000 gpurun_out/bench.json
ARGS="--points 296 --steps 2 --warmup 1 --no-cpu-baseline"
python bench.py $ARGS > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv python bench.py $ARGS > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches exit=$?"
for K in alpha_group integrate_diag; do
  ncu --set full --clock-control none --import-source on -k regex:"$K" -s 1 -c 1 -f -o "gpurun_out/prof_${K}" python bench.py $ARGS > gpurun_out/"ncu_full_${K}.log" 2>&1
  echo "ncu full $K exit=$?"
done
# both row kernels (Gamma, alphaTilde) of the second slice in one report
ncu --set full --clock-control none --import-source on -k regex:rows_group_kernel -s 2 -c 2 -f -o gpurun_out/prof_rows_group python bench.py $ARGS > gpurun_out/ncu_full_rows_group.log 2>&1
echo "ncu full rows_group exit=$?"
ls -la gpurun_out
