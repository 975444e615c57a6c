#!/bin/bash
# Build kernel-experiment libraries (lib/libbricks_b200_<name>.so) next to the product library.
# usage: tools/build_variants.sh [name ...]   (default: all).  Time/verify them on a B200 with
#        VARIANT=<name> BRICKS_LIB=$PWD/bricklib_b200/lib/libbricks_b200_<name>.so python tools/variant_check.py 125
cd "$(dirname "$0")/.."
declare -A FLAGS=(
  # add an entry per experiment: [name]="-DSOME_FLAG ..." (the flag must exist in csrc/); experiment
  # builds also get -DMARCH_WAIT_TRAP (a stage that never arrives traps instead of hanging) and
  # -DBRICKS_EXPERIMENT (BRICKS_MAXCOL overrides the column length)
  [maxcol]="-DBRICKS_MAXCOL_BUILD"
  [nomath]="-DMARCH_EXP_NOMATH"
  [cube160]="-DMARCH_CUBE_REGS=160"
  [star4r200]="-DMARCH_STAR4_REGS=200"
)
names=("$@"); [ ${#names[@]} -eq 0 ] && names=("${!FLAGS[@]}")
for n in "${names[@]}"; do
  [ -z "${FLAGS[$n]:-}" ] && { echo "unknown variant $n"; exit 1; }
  # WAIT_TRAP=0 builds without the hang guard (final timing of a variant that already ran clean)
  guard="-DMARCH_WAIT_TRAP -DBRICKS_EXPERIMENT"; [ "${WAIT_TRAP:-1}" = 0 ] && guard=""
  ( BRICKS_EXP_NAME=$n BRICKS_EXP_FLAGS="${FLAGS[$n]} $guard" python bricklib_b200/build.py >/dev/null && echo "built $n" ) &
done
wait
