#!/bin/bash
# GPU box: parity (CRC + max rel err vs the naive twin) and timing of experiment builds.
# usage: tools/run_variants.sh <out.log> <variant[:stencils[:maxcol]]> ...   ("product" = the product library)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=$PWD/bricklib_b200/lib
OUT=gpurun_out/$1; shift
: > $OUT
for spec in "$@"; do
  IFS=: read -r v st mc <<< "$spec"
  lib=$L/libbricks_b200_$v.so; [ "$v" = product ] && lib=$L/libbricks_b200.so
  VARIANT=$v BRICKS_LIB=$lib BRICKS_MAXCOL=${mc:-} timeout 90 python tools/variant_check.py ${st:-125} 2>&1 | grep -E "VCHECK|Error|error|Traceback" >> $OUT
  echo "rc=$? $v" >> $OUT
done
cat $OUT
