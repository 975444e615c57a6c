#!/bin/bash
# Run on an 8-GPU box: multi-GPU strong parity tests, then the strong-scaling bench at N = 1, 2, 4, 8.
TAG=${1:-r01}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_strong.py -m gpu -x -q > gpurun_out/pytest_strong_8gpu_${TAG}.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/pytest_strong_8gpu_${TAG}.log
for wl in strong125 strong7; do
  for n in 1 2 4 8; do
    if [ $n = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700+n))"; fi
    timeout 400 $L bench.py --gpus $n --steps 20 --warmup 3 --no-cpu --e2e-steps 3 --workload $wl > gpurun_out/strong_${TAG}_${wl}_$n.json 2> gpurun_out/strong_${TAG}_${wl}_$n.err
    echo "$wl n=$n rc=$?"
    python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/strong_${TAG}_${wl}_$n.json').read().strip().splitlines()[-1])
    print('  value=%.1f GStencil/s  ms/step=%.3f  e2e=%.1f kernel frac=%.3f clocks=%s'%(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['clocks']))
except Exception as e:
    print('  no line', e)
PY
  done
done
