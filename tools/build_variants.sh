#!/bin/bash
# build the kernel-experiment libraries (lib/libbricks_b200_<name>.so) next to the product library
cd "$(dirname "$0")/.."
build() { n=$1; shift; BRICKS_EXP_NAME=$n BRICKS_EXP_FLAGS="$*" python bricklib_b200/build.py >/dev/null && echo "built $n"; }
build ldgsts -DMARCH_HALO_LDGSTS &
build win4 -DMARCH_CUBE_WINDOW=4 &
build win4r128 -DMARCH_CUBE_WINDOW=4 -DMARCH_WINDOW_REGS=128 &
build win2r128 -DMARCH_CUBE_WINDOW=2 -DMARCH_WINDOW_REGS=128 &
wait
build win2s3r128 -DMARCH_CUBE_WINDOW=2 -DMARCH_WINDOW_REGS=128 -DMARCH_WINDOW_STAGES=3 &
build win2s4r128 -DMARCH_CUBE_WINDOW=2 -DMARCH_WINDOW_REGS=128 -DMARCH_WINDOW_STAGES=4 &
build win4s3 -DMARCH_CUBE_WINDOW=4 -DMARCH_WINDOW_STAGES=3 &
wait
