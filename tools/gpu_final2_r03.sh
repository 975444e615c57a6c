#!/bin/bash
# one GPU: the whole GPU suite with the final library, then one --set full capture each of the 125- and 25-point sweep
cd "$(dirname "$0")/.."
O=gpurun_out/final2_r03; mkdir -p $O
timeout 300 python -m pytest tests -x -q -m gpu 2>&1 | tail -6 > $O/pytest_gpu_1gpu.log; cat $O/pytest_gpu_1gpu.log
timeout 100 ncu --set full --clock-control none --import-source on -k regex:march_kernel -s 2 -c 1 -o $O/march125 python tools/variant_check.py 125 > $O/ncu125.log 2>&1; echo "ncu125 rc=$?"
timeout 100 ncu --set full --clock-control none --import-source on -k regex:march_kernel -s 2 -c 1 -o $O/march25 python tools/variant_check.py 25 > $O/ncu25.log 2>&1; echo "ncu25 rc=$?"
grep VCHECK $O/ncu125.log $O/ncu25.log
