#!/bin/bash
# One short GPU call for the 4-D path (plain C++ driver, no Python start-up): validation against the host
# array twin at 32^4, then timing at 64^4 with the marching kernel and with the twin kernel.
mkdir -p gpurun_out/d4
D=drivers/_build/weak4d
timeout 25 $D -s 32 -I 5 -v > gpurun_out/d4/validate32.log 2>&1; echo "validate32 rc=$?" | tee -a gpurun_out/d4/rc.log
timeout 25 $D -s 16,24,16,20 -I 2 -v -a > gpurun_out/d4/validate_aligned.log 2>&1; echo "validate_aligned rc=$?" | tee -a gpurun_out/d4/rc.log
timeout 30 $D -s 64 -I 10 -v > gpurun_out/d4/time64.log 2>&1; echo "time64 rc=$?" | tee -a gpurun_out/d4/rc.log
timeout 20 $D -s 64 -I 5 -t > gpurun_out/d4/time64_twin.log 2>&1; echo "time64_twin rc=$?" | tee -a gpurun_out/d4/rc.log
tail -n 4 gpurun_out/d4/validate32.log gpurun_out/d4/validate_aligned.log
tail -n 12 gpurun_out/d4/time64.log
tail -n 3 gpurun_out/d4/time64_twin.log
