#!/bin/bash
# build the kernel-experiment libraries (lib/libbricks_b200_<name>.so) next to the product library
cd "$(dirname "$0")/.."
build() { n=$1; shift; BRICKS_EXP_NAME=$n BRICKS_EXP_FLAGS="$*" python bricklib_b200/build.py >/dev/null && echo "built $n"; }
build pinadj -DMARCH_PIN_ADJ &
build corner -DMARCH_CORNER_LDGSTS &
build cornerpin -DMARCH_CORNER_LDGSTS -DMARCH_PIN_ADJ &
build tile2s4np -DMARCH_CUBE_TILE=2 -DMARCH_TILE_STAGES=4 -DMARCH_TILE_NOPAD &
wait
