#!/bin/bash
# Run on the GPU box (under gpurun): bench line, ncu launch list, one full capture of the sweep kernel.
# usage: tools/gpu_profile.sh <tag> [workload]
set -u
TAG=${1:-r01}
WL=${2:-weak125}
mkdir -p gpurun_out
python bench.py --workload $WL --all-stencils > gpurun_out/bench_${TAG}_${WL}.json 2> gpurun_out/bench_${TAG}_${WL}.err
echo "bench rc=$?"; tail -c 3000 gpurun_out/bench_${TAG}_${WL}.json
CMD="python bench.py --workload $WL --steps 1 --warmup 3 --no-cpu --e2e-steps 1"
$CMD > gpurun_out/plain_${TAG}.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${TAG}_${WL}.csv $CMD > gpurun_out/ncu_list_${TAG}.log 2>&1
echo "ncu list rc=$?"
$CMD > gpurun_out/plain2_${TAG}.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:march_kernel -s 12 -c 2 -o gpurun_out/prof_${TAG}_${WL} -f $CMD > gpurun_out/ncu_full_${TAG}.log 2>&1
echo "ncu full rc=$?"; tail -5 gpurun_out/ncu_full_${TAG}.log
