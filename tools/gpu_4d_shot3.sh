#!/bin/bash
# Third short GPU call for the 4-D path: load width and CTA shape variants of line_kernel on one box.
mkdir -p gpurun_out/d4
D=drivers/_build/weak4d
L=gpurun_out/d4/variants.log
: > $L
BRICKS_4D_VARIANT=221 timeout 5 $D -s 32 -I 3 -v 2>&1 | grep -E "result|rank" | sed 's/^/221 validate: /' >> $L
for v in 221 220 411 211 111; do
  BRICKS_4D_VARIANT=$v timeout 5 $D -s 64 -I 10 2>&1 | grep -E "perf|full-grid|rank" | tr '\n' ' ' | sed "s/^/$v: /" >> $L
  echo >> $L
done
cat $L
