#!/bin/bash
# The prepared sweep experiments of DESIGN.md 4.1 / 9 in one go.
#   here (no GPU):   tools/exp_next.sh build
#   on the GPU box:  gpurun --timeout 240 -- 'tools/exp_next.sh run'      (~5 s per entry)
# Every entry is checked bit for bit against the naive twin and by CRC against the product
# (tools/variant_check.py); experiment builds trap instead of hanging (-DMARCH_WAIT_TRAP).
cd "$(dirname "$0")/.."
V="jlswz edgehi jlswzedge jlswzpersist nbtab nbahead nbtabahead persist"
case "$1" in
  build) tools/build_variants.sh $V ;;
  run)   timeout 220 tools/run_variants.sh variants_next.log product:125,25,13,7 jlswz:125,25,13 edgehi:125,25,13,7 \
           jlswzedge:125,25 jlswzpersist:125 nbtab:125,7 nbahead:125 nbtabahead:125 persist:125:13 persist:7 ;;
  *) echo "usage: $0 build|run"; exit 1 ;;
esac
