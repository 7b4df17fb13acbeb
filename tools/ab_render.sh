#!/bin/bash
# A/B of render throughput on one GPU box: the default library and every variant in csrc/variants (tools/build_variant.sh).
# usage: tools/ab_render.sh [spp] [reps]   (SCENES="random_ned cornell_smoke_ned" by default; each line: Mpaths/s + per-iteration kernel times)
spp=${1:-256}; reps=${2:-2}
V=ray-tracing-one-weekend_b200/csrc/variants
for rep in $(seq $reps); do
for lib in default $(ls $V/*.so 2>/dev/null); do
  [ $lib = default ] && unset RTOW_B200_LIB || export RTOW_B200_LIB=$PWD/$lib
  for sc in ${SCENES:-random_ned cornell_smoke_ned}; do
    echo -n "$(basename $lib) "; python tools/prof_render.py --scene $sc --spp $spp --reps 2 2>&1 | tr '\n' ' '; echo
  done
done
done
unset RTOW_B200_LIB
