#!/bin/bash
# GPU box with N GPUs: multi-GPU parity tests, then the weak 125-point step under every form of the
# exchange / schedule, each with a per-rank timeline of one step (bricks_step_trace).
# usage: tools/gpu_scale_r03.sh N [tag] [quick]
cd "$(dirname "$0")/.."
N=${1:-2}; TAG=${2:-r03}; QUICK=${3:-}
O=gpurun_out/scale_${TAG}_n$N; mkdir -p $O
python -m pytest tests -x -q -m gpu 2>&1 | tail -6 > $O/pytest.log
run() { # name, env..., -- bench args
  name=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
     --master-port $((29500 + RANDOM % 400)) bench.py --gpus $N --steps 40 --warmup 3 --no-strong --no-all-stencils --no-cpu \
     --e2e-steps 2 --trace $O/trace_$name "$@" > $O/$name.json 2> $O/$name.err
  python - "$O/$name.json" "$name" <<'PY'
import json,sys
try:
    l=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("%-28s value %8.1f GStencil/s  ms/step %.4f  parity %s" % (sys.argv[2], l["value"], l["ms_per_step"], l["parity"]["ok"] if l.get("parity") else None))
except Exception as e:
    print(sys.argv[2], "FAILED", e)
PY
}
{
run regions_early X=1 -- --no-unsplit
run regions_early_unsplit X=1 -- --unsplit
run packed_early_unsplit X=1 -- --exchange packed --unsplit
run shift_early_unsplit X=1 -- --exchange shift --unsplit
if [ -z "$QUICK" ]; then
run regions_noearly X=1 -- --no-early
run packed_early X=1 -- --exchange packed --no-unsplit
run regions_early_cta4 BRICKS_NCCL_MAX_CTAS=4 -- --no-unsplit
run packed_unsplit_cta8 BRICKS_NCCL_MAX_CTAS=8 -- --exchange packed --unsplit
run regions_serial X=1 -- --no-overlap
fi
} | tee $O/SUMMARY.txt
