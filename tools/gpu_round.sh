#!/bin/bash
# usage (under gpurun): tools/gpu_round.sh <tag> [tests|bench|ncu ...]   -- everything lands in gpurun_out/<tag>_*
# tests: the whole -m gpu suite + smoke();  bench: headline + fine bench lines;  ncu: launch list + --set full captures of
# the step's kernels at the benchmarked launch shape (one 2048-point pipeline slice in scan order), each only after the
# same command has exited 0 without ncu.
TAG=${1:-rXX}; shift
WHAT=${*:-tests bench ncu}
mkdir -p gpurun_out
O=gpurun_out/$TAG
for W in $WHAT; do
case $W in
tests)
  timeout 2400 python -m pytest tests -m gpu -q -x --durations=8 > ${O}_pytest_gpu.log 2>&1; echo "pytest exit=$?"; tail -4 ${O}_pytest_gpu.log
  timeout 600 python -c 'import __graft_entry__ as g; g.smoke()' > ${O}_smoke.log 2>&1; echo "smoke exit=$?"; tail -1 ${O}_smoke.log | cut -c1-400
  ;;
bench)
  timeout 900 python bench.py --steps 3 --warmup 3 > ${O}_bench.json 2> ${O}_bench.err; echo "bench exit=$?"; cut -c1-1500 ${O}_bench.json
  timeout 900 python bench.py --workload fine --steps 2 --warmup 1 > ${O}_bench_fine.json 2> ${O}_bench_fine.err; echo "fine exit=$?"; cut -c1-900 ${O}_bench_fine.json
  ;;
ncu)
  ARGS="--points 2048 --steps 1 --warmup 1 --no-cpu-baseline"
  python bench.py $ARGS > ${O}_plain.log 2>&1 || { echo "plain run failed"; tail -5 ${O}_plain.log; continue; }
  ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file ${O}_launches.csv python bench.py $ARGS > ${O}_ncu_launch.log 2>&1
  echo "ncu launches exit=$?"
  # per slice: rows_shared<0>, rows_point<0>, rows_shared<1>, rows_point<1>, edges_group, alpha_group, integrate_split; skip the warm-up step's seven
  ncu --set full --clock-control none --import-source on -k regex:"rows_|edges_|alpha_group|integrate_split|integrate_diag" -s 7 -c 7 -f \
      -o ${O}_prof python bench.py $ARGS > ${O}_ncu_full.log 2>&1
  echo "ncu full exit=$?"
  ;;
esac
done
ls -la gpurun_out | head -40
