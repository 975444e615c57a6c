#!/bin/bash
# GPU box with N GPUs: the peer-memory exchange (bricks_step_p2p_*) in both modes -- push kernel and one
# copy-engine copy per region -- against NCCL, each command under a hard timeout.
# usage: tools/gpu_p2p_try.sh N [notest]
cd "$(dirname "$0")/.."
N=${1:-2}; O=gpurun_out/p2p3_n$N; mkdir -p $O
if [ "$2" != notest ]; then
  for mode in hybrid; do
    BRICKS_P2P_MODE=$mode timeout 120 python -m pytest tests/test_gpu_exchange.py -x -q -m gpu -k "copy_engine" 2>&1 | tail -3 > $O/pytest_$mode.log
    echo "pytest $mode: $(tail -1 $O/pytest_$mode.log)"
  done
fi
for v in "hybrid_default X=1" ${ALSO_CE:+"ce_default BRICKS_P2P_MODE=ce"}; do
  set -- $v; name=$1; envv=$2; shift 2
  env $envv timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
    bench.py --gpus $N --steps 40 --warmup 3 --no-strong --no-all-stencils --no-cpu --e2e-steps 2 --trace $O/trace_$name "$@" \
    > $O/$name.json 2> $O/$name.err
  echo "rc=$? $name"; python - "$O/$name.json" "$name" <<'PY'
import json,sys
try:
    l=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("%-22s value %8.1f GStencil/s  ms/step %.4f  parity %s  e2e %.1f" % (sys.argv[2], l["value"], l["ms_per_step"], l["parity"]["ok"], l["e2e"]["value"]))
except Exception as e:
    print(sys.argv[2], "FAILED", e)
PY
done 2>&1 | tee $O/SUMMARY.txt
