#!/bin/bash
# Build kernel-experiment libraries (lib/libbricks_b200_<name>.so) next to the product library.
# usage: tools/build_variants.sh [name ...]   (default: all).  Time/verify them on a B200 with
#        VARIANT=<name> BRICKS_LIB=$PWD/bricklib_b200/lib/libbricks_b200_<name>.so python tools/variant_check.py 125
cd "$(dirname "$0")/.."
declare -A FLAGS=(
  [uni]="-DMARCH_UNI_ISSUE"
  [nbahead]="-DMARCH_NB_AHEAD"
  [nbtab]="-DMARCH_NB_TABLE"
  [nbtabahead]="-DMARCH_NB_TABLE -DMARCH_NB_AHEAD"
  [persist]="-DMARCH_NB_TABLE -DMARCH_PERSIST"
  [jlswz]="-DMARCH_JL_SWIZZLE"
  [edgehi]="-DMARCH_EDGE_HI"
  [jlswzedge]="-DMARCH_JL_SWIZZLE -DMARCH_EDGE_HI"
  [jlswzpersist]="-DMARCH_JL_SWIZZLE -DMARCH_NB_TABLE -DMARCH_PERSIST"
  [win4]="-DMARCH_CUBE_WINDOW=4"
  [win2r128]="-DMARCH_CUBE_WINDOW=2 -DMARCH_WINDOW_REGS=128"
  [win2s3r128]="-DMARCH_CUBE_WINDOW=2 -DMARCH_WINDOW_REGS=128 -DMARCH_WINDOW_STAGES=3"
  [tile2]="-DMARCH_CUBE_TILE=2"
  [tile2s4np]="-DMARCH_CUBE_TILE=2 -DMARCH_TILE_STAGES=4 -DMARCH_TILE_NOPAD"
  [tile2r168]="-DMARCH_CUBE_TILE=2 -DMARCH_TILE_REGS=168 -DMARCH_TILE_STAGES=2"
  [tile4s2]="-DMARCH_CUBE_TILE=4 -DMARCH_TILE_STAGES=2"
  [tile4s2np]="-DMARCH_CUBE_TILE=4 -DMARCH_TILE_STAGES=2 -DMARCH_TILE_NOPAD"
)
names=("$@"); [ ${#names[@]} -eq 0 ] && names=("${!FLAGS[@]}")
for n in "${names[@]}"; do
  [ -z "${FLAGS[$n]:-}" ] && { echo "unknown variant $n"; exit 1; }
  # WAIT_TRAP=0 builds without the hang guard (final timing of a variant that already ran clean)
  guard="-DMARCH_WAIT_TRAP -DBRICKS_EXPERIMENT"; [ "${WAIT_TRAP:-1}" = 0 ] && guard=""
  ( BRICKS_EXP_NAME=$n BRICKS_EXP_FLAGS="${FLAGS[$n]} $guard" python bricklib_b200/build.py >/dev/null && echo "built $n" ) &
done
wait
