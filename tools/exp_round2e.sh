#!/bin/bash
# GPU box: product-kernel experiments (pinned adjacency prefetch, corner pieces by cp.async) and the
# four-stage unpadded tile kernel
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=$PWD/bricklib_b200/lib
OUT=gpurun_out/variants_e.log
: > $OUT
run() { # variant stencils [maxcol]
  local lib=$L/libbricks_b200_$1.so; [ "$1" = product ] && lib=$L/libbricks_b200.so
  VARIANT=$1 BRICKS_LIB=$lib BRICKS_MAXCOL=${3:-} timeout 90 python tools/variant_check.py $2 2>&1 | grep -E "VCHECK|Error|error|Traceback" >> $OUT
  echo "rc=$? $1" >> $OUT
}
run product 7,13,25,125
run pinadj 7,13,25,125
run corner 125
run cornerpin 125
run tile2s4np 125
run tile2s4np 125 11
run tile2s4np 125 8
run product 125 11
run product 125 22
cat $OUT
