#!/bin/bash
# time the kernel variants built by BRICKS_EXP_NAME (see bricklib_b200/build.py); results are throw-away
cd "$(dirname "$0")/.."
L=bricklib_b200/lib
for v in "" $@; do
  for P in 4 8; do
    if [ -z "$v" ]; then lib=$L/libbricks_b200.so; else lib=$L/libbricks_b200_$v.so; fi
    echo "== variant=${v:-product} P=$P"
    BRICKS_LIB=$PWD/$lib BRICKS_PLANES=$P timeout 120 python tools/kernel_timing.py 7,13,25,125 2>&1 | grep KTIME | sed 's/env=.*ms=/ms=/'
  done
done
