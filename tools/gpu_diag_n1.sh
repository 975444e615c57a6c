#!/bin/bash
# one GPU: does the harness itself cost step time?  the same N = 1 bench with / without --trace and the clock sampler
cd "$(dirname "$0")/.."
O=gpurun_out/diag_n1; mkdir -p $O
B="python bench.py --steps 40 --warmup 3 --no-cpu --no-strong --no-all-stencils --e2e-steps 2 --no-parity"
timeout 60 $B > $O/plain.json 2>$O/plain.err
timeout 60 $B --trace $O/t > $O/trace.json 2>$O/trace.err
BENCH_NO_CLOCKS=1 timeout 60 $B > $O/noclocks.json 2>$O/noclocks.err
timeout 60 $B --steps 200 > $O/steps200.json 2>$O/steps200.err
timeout 60 $B --no-overlap > $O/serial.json 2>$O/serial.err
timeout 60 $B > $O/plain2.json 2>$O/plain2.err
python - <<'PY'
import json
for n in ("plain","trace","noclocks","steps200","serial","plain2"):
    try:
        l=json.loads(open("gpurun_out/diag_n1/%s.json"%n).read().strip().splitlines()[-1])
        print("%-9s ms/step %.4f kernel_ms %.4f clocks %s" % (n, l["ms_per_step"], l["roofline"]["kernel_ms"], l["clocks"]))
    except Exception as e:
        print(n, "FAILED", e)
PY
