#!/bin/bash
# GPU box with N GPUs: the copy-engine exchange (bricks_step_p2p_*) -- parity tests, then bench with and
# without it, each under a hard timeout.   usage: tools/gpu_p2p_try.sh N
cd "$(dirname "$0")/.."
N=${1:-2}; O=gpurun_out/p2p_n$N; mkdir -p $O
timeout 150 python -m pytest tests/test_gpu_exchange.py -x -q -m gpu -k "copy_engine" 2>&1 | tail -8 > $O/pytest.log; cat $O/pytest.log
for v in "p2p_noearly --p2p --no-early" "p2p_early --p2p" "nccl_early --no-p2p" "p2p_shift_noearly --p2p --no-early --exchange shift"; do
  set -- $v; name=$1; shift
  timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
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
tail -5 $O/p2p_noearly.err
