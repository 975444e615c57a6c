#!/bin/bash
# GPU box with N GPUs: multi-GPU parity tests, then the weak 125-point step under every form of the
# exchange / schedule, each with a per-rank timeline of one step (bricks_step_trace).
# usage: tools/gpu_scale_r03.sh N [tag] [quick]
cd "$(dirname "$0")/.."
N=${1:-2}; TAG=${2:-r03}; QUICK=${3:-}
O=gpurun_out/scale_${TAG}_n$N; mkdir -p $O
if [ -n "$QUICK" ]; then  # the files with multi-GPU cases only
  python -m pytest tests/test_gpu_exchange.py tests/test_shift_exchange.py tests/test_strong.py tests/test_drivers.py \
    tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -6 > $O/pytest.log
else
  python -m pytest tests -x -q -m gpu 2>&1 | tail -6 > $O/pytest.log
fi
run() { # name, env..., -- bench args
  name=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
     --master-port $((29500 + RANDOM % 400)) bench.py --gpus $N --steps 40 --warmup 3 --no-strong --no-all-stencils --no-cpu \
     --e2e-steps ${E2E:-2} --trace $O/trace_$name "$@" > $O/$name.json 2> $O/$name.err
  python - "$O/$name.json" "$name" <<'PY'
import json,sys
try:
    l=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    hl = l["e2e"].get("host_link", {})
    print("%-28s value %8.1f GStencil/s  ms/step %.4f  parity %s | e2e %.1f GStencil/s %.1f ms/step, link alone h2d %.1f d2h %.1f both %.1f GB/s, pcie_frac %.2f (%s, node %s)" % (
        sys.argv[2], l["value"], l["ms_per_step"], l["parity"]["ok"] if l.get("parity") else None, l["e2e"]["value"],
        l["e2e"]["ms_per_step"], hl.get("h2d_alone_gbs", 0), hl.get("d2h_alone_gbs", 0), hl.get("both_directions_gbs_each", 0),
        hl.get("pcie_frac", 0), hl.get("pinned_alloc"), hl.get("numa_node_of_pinned_buffers")))
except Exception as e:
    print(sys.argv[2], "FAILED", e)
PY
}
{
E2E=4 run regions_early_unsplit X=1 -- --unsplit
E2E=4 run regions_unsplit_plainalloc X=1 -- --unsplit --e2e-plain-alloc
run regions_early X=1 -- --no-unsplit
run packed_early_unsplit X=1 -- --exchange packed --unsplit
run shift_early_unsplit X=1 -- --exchange shift --unsplit
if [ "$QUICK" = ce ]; then  # NCCL P2P through the copy engines instead of an SM-resident kernel
run regions_noearly X=1 -- --no-early
run ce_regions_early NCCL_P2P_USE_CUDA_MEMCPY=1 --
run ce_regions_noearly NCCL_P2P_USE_CUDA_MEMCPY=1 -- --no-early
run ce_packed_noearly NCCL_P2P_USE_CUDA_MEMCPY=1 -- --exchange packed --no-early
fi
if [ -z "$QUICK" ]; then
run regions_noearly X=1 -- --no-early
run packed_early X=1 -- --exchange packed --no-unsplit
run regions_early_cta4 BRICKS_NCCL_MAX_CTAS=4 -- --no-unsplit
run packed_unsplit_cta8 BRICKS_NCCL_MAX_CTAS=8 -- --exchange packed --unsplit
run regions_serial X=1 -- --no-overlap
fi
} | tee $O/SUMMARY.txt
