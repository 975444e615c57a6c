#!/bin/bash
# Run on the GPU box (under gpurun), ONE ncu pass per call:
#   tools/gpu_profile.sh list <tag> [workload]   bench line + ncu launch list of the same command
#   tools/gpu_profile.sh full <tag> [workload]   one --set full capture of the sweep kernel
set -u
MODE=${1:-list}
TAG=${2:-r01}
WL=${3:-weak125}
mkdir -p gpurun_out
CMD="python bench.py --workload $WL --steps 1 --warmup 3 --no-cpu --e2e-steps 2"
if [ "$MODE" = list ]; then
  python bench.py --workload $WL --all-stencils > gpurun_out/bench_${TAG}_${WL}.json 2> gpurun_out/bench_${TAG}_${WL}.err
  echo "bench rc=$?"; tail -c 2500 gpurun_out/bench_${TAG}_${WL}.json
  $CMD > gpurun_out/plain_${TAG}.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${TAG}_${WL}.csv $CMD > gpurun_out/ncu_list_${TAG}.log 2>&1
  echo "ncu list rc=$?"
else
  $CMD > gpurun_out/plain2_${TAG}.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:march_kernel -s 12 -c 2 -o gpurun_out/prof_${TAG}_${WL} -f $CMD > gpurun_out/ncu_full_${TAG}.log 2>&1
  echo "ncu full rc=$?"; tail -5 gpurun_out/ncu_full_${TAG}.log
fi
