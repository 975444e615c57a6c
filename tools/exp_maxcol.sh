#!/bin/bash
cd "$(dirname "$0")/.."
for m in ${@:-6 7 8 9 10 11 12 14 17 22}; do
  echo "== maxcol=$m"
  BRICKS_MAXCOL=$m timeout 120 python tools/kernel_timing.py 7,13,25,125 2>&1 | grep KTIME | sed 's/env=.*ms=/ms=/'
done
