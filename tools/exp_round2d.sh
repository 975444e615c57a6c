#!/bin/bash
# GPU box: parity + timing of the tile-form cube kernel builds (branch-free stores, copies after the wait)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=$PWD/bricklib_b200/lib
OUT=gpurun_out/variants_d.log
: > $OUT
run() { # variant stencils [maxcol]
  local lib=$L/libbricks_b200_$1.so; [ "$1" = product ] && lib=$L/libbricks_b200.so
  VARIANT=$1 BRICKS_LIB=$lib BRICKS_MAXCOL=${3:-} timeout 90 python tools/variant_check.py $2 2>&1 | grep -E "VCHECK|Error|error|Traceback" >> $OUT
  echo "rc=$? $1" >> $OUT
}
run st256nb 7,13,25,125
for v in tile2 tile2s2 tile2s4 tile4s2 tile4s2np; do run $v 125; done
for mc in 8 11 13; do run tile2 125 $mc; run tile4s2np 125 $mc; done
cat $OUT
V=${NCU_VARIANT:-tile4s2np}
VARIANT=$V BRICKS_LIB=$L/libbricks_b200_$V.so timeout 200 ncu --set full --clock-control none --import-source on -k regex:cube_tile -s 3 -c 1 -o gpurun_out/prof_r02d_$V -f python tools/variant_check.py 125 > gpurun_out/ncu_r02d_$V.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_r02d_$V.log
