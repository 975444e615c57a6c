#!/bin/bash
# One GPU, end of round 3: column-length variants of the 25-point kernel, the whole GPU suite, the default
# bench line, the reference arm, the ncu launch list of the bench command and one --set full capture each
# of the 125- and the 25-point sweep.  Every command under its own timeout.
cd "$(dirname "$0")/.."
O=gpurun_out/final_r03; mkdir -p $O
L=$PWD/bricklib_b200/lib
for mc in 8 13 17; do
  VARIANT=maxcol BRICKS_LIB=$L/libbricks_b200_maxcol.so BRICKS_MAXCOL=$mc timeout 60 python tools/variant_check.py 25 2>&1 | grep VCHECK
done | tee $O/variants_l_maxcol25.log
VARIANT=product timeout 60 python tools/variant_check.py 125,25,13,7 2>&1 | grep VCHECK | tee -a $O/variants_l_maxcol25.log
timeout 240 python -m pytest tests -x -q -m gpu 2>&1 | tail -6 > $O/pytest_gpu_1gpu.log; cat $O/pytest_gpu_1gpu.log
timeout 200 python bench.py --steps 20 --warmup 3 > $O/bench_weak125.json 2> $O/bench_weak125.err; echo "bench rc=$?"
timeout 100 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref_weak125.json 2>/dev/null; echo "ref125 rc=$?"
timeout 60 python bench.py --impl reference --workload weak7 --steps 2 --warmup 1 > $O/bench_ref_weak7.json 2>/dev/null; echo "ref7 rc=$?"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-strong --no-all-stencils --no-parity --e2e-steps 2"
timeout 120 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file $O/ncu_launches_weak125.csv $CMD > $O/ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 120 ncu --set full --clock-control none --import-source on -k regex:march_kernel -s 2 -c 1 -o $O/march125 python tools/variant_check.py 125 > $O/ncu125.log 2>&1; echo "ncu125 rc=$?"
timeout 120 ncu --set full --clock-control none --import-source on -k regex:march_kernel -s 2 -c 1 -o $O/march25 python tools/variant_check.py 25 > $O/ncu25.log 2>&1; echo "ncu25 rc=$?"
ls -la $O | tail -20
