#!/bin/bash
# Run on an 8-GPU box: multi-GPU parity (exchange, weak step, strong, drivers), then weak scaling at
# N = 1, 2, 4, 8 with the early-exchange schedule, N = 8 without it, and strong125 at N = 1 and 8.
TAG=${1:-r01b}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_exchange.py tests/test_strong.py tests/test_drivers.py -m gpu -q -k "multi_gpu or strong or drivers" > gpurun_out/pytest_8gpu_${TAG}.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/pytest_8gpu_${TAG}.log
run() { # workload n extra-flags name
  local wl=$1 n=$2 extra=$3 name=$4
  if [ $n = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29800+n))"; fi
  timeout 300 $L bench.py --gpus $n --steps 50 --warmup 3 --no-cpu --e2e-steps 3 --workload $wl $extra > gpurun_out/${name}.json 2> gpurun_out/${name}.err
  echo "$name rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${name}.json').read().strip().splitlines()[-1])
    print('  value=%.1f GStencil/s  ms/step=%.3f  e2e=%.1f frac=%.3f clocks=%s'%(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['clocks']))
except Exception as e:
    print('  no line', e)
PY
}
for wl in weak125 weak7; do
  for n in 1 2 4 8; do run $wl $n "" scale_${TAG}_${wl}_$n; done
  run $wl 8 "--no-early" scale_${TAG}_${wl}_8_noearly
done
run strong125 1 "" strong_${TAG}_strong125_1
run strong125 8 "" strong_${TAG}_strong125_8
