#!/bin/bash
# GPU box: parity + timing of the cube-sweep experiment builds (tools/build_variants.sh), then one
# ncu --set full capture of the window kernel.  Output: gpurun_out/variants.log
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=$PWD/bricklib_b200/lib
OUT=gpurun_out/variants.log
: > $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader >> $OUT
run() { # variant stencils [maxcol]
  local lib=$L/libbricks_b200_$1.so; [ "$1" = product ] && lib=$L/libbricks_b200.so
  VARIANT=$1 BRICKS_LIB=$lib BRICKS_MAXCOL=${3:-} timeout 90 python tools/variant_check.py $2 2>&1 | grep -E "VCHECK|Error|error|Traceback" >> $OUT
  echo "rc=$? $1" >> $OUT
}
run product 7,13,25,125
run ldgsts 7,13,25,125
for v in win4 win4r128 win2r128 win2s3r128 win2s4r128 win4s3; do run $v 125; done
for mc in 11 22 31; do run win4 125 $mc; run win2s3r128 125 $mc; done
for mc in 5 11; do run ldgsts 7,13 $mc; done
cat $OUT
V=${NCU_VARIANT:-win4}
VARIANT=$V BRICKS_LIB=$L/libbricks_b200_$V.so timeout 200 ncu --set full --clock-control none --import-source on -k regex:cube_window -s 3 -c 1 -o gpurun_out/prof_r02_$V -f python tools/variant_check.py 125 > gpurun_out/ncu_r02_$V.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_r02_$V.log
