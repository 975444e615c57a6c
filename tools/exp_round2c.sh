#!/bin/bash
# GPU box: parity + timing of the tile-form cube kernel builds (memory-order cp.async, 256-bit stores)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=$PWD/bricklib_b200/lib
OUT=gpurun_out/variants_c.log
: > $OUT
run() { # variant stencils [maxcol]
  local lib=$L/libbricks_b200_$1.so; [ "$1" = product ] && lib=$L/libbricks_b200.so
  VARIANT=$1 BRICKS_LIB=$lib BRICKS_MAXCOL=${3:-} timeout 90 python tools/variant_check.py $2 2>&1 | grep -E "VCHECK|Error|error|Traceback" >> $OUT
  echo "rc=$? $1" >> $OUT
}
run st256 7,13,25,125
for v in tile2 tile2nost tile2s2 tile2s4 tile2r168 tile4s2; do run $v 125; done
for mc in 6 8 11 22; do run tile2 125 $mc; run tile4s2 125 $mc; done
cat $OUT
V=${NCU_VARIANT:-tile2}
VARIANT=$V BRICKS_LIB=$L/libbricks_b200_$V.so timeout 200 ncu --set full --clock-control none --import-source on -k regex:cube_tile -s 3 -c 1 -o gpurun_out/prof_r02c_$V -f python tools/variant_check.py 125 > gpurun_out/ncu_r02c_$V.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_r02c_$V.log
