#!/bin/bash
# build the kernel-experiment libraries (lib/libbricks_b200_<name>.so) next to the product library
cd "$(dirname "$0")/.."
build() { n=$1; shift; BRICKS_EXP_NAME=$n BRICKS_EXP_FLAGS="$*" python bricklib_b200/build.py >/dev/null && echo "built $n"; }
build st256 -DMARCH_ST256 &
build st256nb -DMARCH_ST256 -DMARCH_STORE_NB &
build tile2 -DMARCH_CUBE_TILE=2 -DMARCH_ST256 &
build tile2s2 -DMARCH_CUBE_TILE=2 -DMARCH_TILE_STAGES=2 -DMARCH_ST256 &
wait
build tile4s2 -DMARCH_CUBE_TILE=4 -DMARCH_TILE_STAGES=2 -DMARCH_ST256 &
build tile4s2np -DMARCH_CUBE_TILE=4 -DMARCH_TILE_STAGES=2 -DMARCH_ST256 -DMARCH_TILE_NOPAD &
build tile2s4 -DMARCH_CUBE_TILE=2 -DMARCH_TILE_STAGES=4 -DMARCH_ST256 &
wait
