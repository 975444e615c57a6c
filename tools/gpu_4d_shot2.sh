#!/bin/bash
# Second short GPU call for the 4-D path: its GPU tests through pytest, then one ncu --set full capture of
# line_kernel from the C++ driver.
mkdir -p gpurun_out/d4
timeout 33 python -m pytest tests/test_4d.py -m gpu -x -q -s > gpurun_out/d4/pytest_gpu_4d.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/d4/rc2.log
tail -n 6 gpurun_out/d4/pytest_gpu_4d.log
timeout 20 ncu --set full --clock-control none --import-source on -k regex:line_kernel -s 20 -c 1 -f -o gpurun_out/d4/line64 \
  drivers/_build/weak4d -s 64 -I 1 > gpurun_out/d4/ncu_line64.log 2>&1; echo "ncu rc=$?" | tee -a gpurun_out/d4/rc2.log
tail -n 3 gpurun_out/d4/ncu_line64.log
