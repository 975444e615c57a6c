#!/bin/bash
# GPU box: parity + timing of the tile-form cube kernel builds, the GPU stencil tests on the tile
# build, then one ncu --set full capture.  Output: gpurun_out/variants_b.log
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=$PWD/bricklib_b200/lib
OUT=gpurun_out/variants_b.log
: > $OUT
run() { # variant stencils [maxcol]
  local lib=$L/libbricks_b200_$1.so; [ "$1" = product ] && lib=$L/libbricks_b200.so
  VARIANT=$1 BRICKS_LIB=$lib BRICKS_MAXCOL=${3:-} timeout 90 python tools/variant_check.py $2 2>&1 | grep -E "VCHECK|Error|error|Traceback" >> $OUT
  echo "rc=$? $1" >> $OUT
}
for v in tile2 tile2s4 tile2r168 tile4s2; do for mc in "" 6 8 11 22; do run $v 125 $mc; done; done
cat $OUT
V=${NCU_VARIANT:-tile2}
BRICKS_LIB=$L/libbricks_b200_$V.so timeout 120 python -m pytest tests/test_gpu_stencil.py tests/test_gpu_exchange.py -m gpu -x -q 2>&1 | tail -4
VARIANT=$V BRICKS_LIB=$L/libbricks_b200_$V.so timeout 200 ncu --set full --clock-control none --import-source on -k regex:cube_tile -s 3 -c 1 -o gpurun_out/prof_r02_$V -f python tools/variant_check.py 125 > gpurun_out/ncu_r02_$V.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_r02_$V.log
