#!/bin/bash
# one box: the weak 125-point step at N = 1 and N = 2 back to back (same GPUs, same thermal state), the
# N = 2 run being the full default bench line (parity, all stencils, strong block, e2e)
cd "$(dirname "$0")/.."
O=gpurun_out/same_box; mkdir -p $O
timeout 120 python bench.py --steps 40 --warmup 3 --no-cpu --no-strong --no-all-stencils --e2e-steps 2 --trace $O/trace_n1 > $O/n1.json 2> $O/n1.err; echo "rc=$? n1"
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus 2 --steps 40 --warmup 3 --trace $O/trace_n2 > $O/n2.json 2> $O/n2.err; echo "rc=$? n2"
python - <<'PY'
import json
for n in ("n1","n2"):
    try:
        l=json.loads(open("gpurun_out/same_box/%s.json"%n).read().strip().splitlines()[-1])
        print(n, "value %.1f ms/step %.4f parity %s e2e %.1f strong %s" % (l["value"], l["ms_per_step"], l["parity"]["ok"], l["e2e"]["value"],
              {k: l["strong"].get(k) for k in ("value","ms_per_step","error")} if l.get("strong") else None))
    except Exception as e:
        print(n, "FAILED", e)
PY
