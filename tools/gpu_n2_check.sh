#!/bin/bash
# two GPUs: the multi-GPU parity tests and the default bench line as the driver launches it
cd "$(dirname "$0")/.."
O=gpurun_out/n2_check; mkdir -p $O
timeout 200 python -m pytest tests/test_gpu_exchange.py tests/test_shift_exchange.py tests/test_strong.py tests/test_drivers.py -x -q -m gpu \
  -k "copy_engine or multi_gpu or kinds" 2>&1 | tail -5 > $O/pytest.log; cat $O/pytest.log
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
  bench.py --gpus 2 --steps 20 --warmup 3 > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench rc=$?"
python - <<'PY'
import json
l=json.loads(open("gpurun_out/n2_check/bench_n2.json").read().strip().splitlines()[-1])
print("value %.1f ms/step %.4f parity %s p2p %s note %s exchange: %s" % (l["value"], l["ms_per_step"], l["parity"]["ok"], l["peer_memory_exchange"], l["peer_memory_exchange_note"], l["exchange"]))
print("e2e %.1f strong %s cpu %s" % (l["e2e"]["value"], l["strong"].get("value"), l["cpu_baseline"]))
PY
tail -3 $O/bench_n2.err
