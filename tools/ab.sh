#!/bin/bash
# A/B on one GPU box: the default library and every variant in csrc/variants (tools/build_variant.sh).
# usage: tools/ab.sh [spp]   -> config-5 closest-hit Mrays/s and config-2 / Cornell render Mpaths/s per library
spp=${1:-128}
V=ray-tracing-one-weekend_b200/csrc/variants
for lib in default $(ls $V/*.so 2>/dev/null); do
  echo "== $lib"
  [ $lib = default ] && unset RTOW_B200_LIB || export RTOW_B200_LIB=$PWD/$lib
  python tools/prof_trace.py --rays 22 --reps 5
  for sc in ${SCENES:-random_ned cornell_smoke_ned}; do
  python bench.py --steps 2 --warmup 3 --spp $spp --no-cpu --no-bvh --scene $sc 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$sc', round(d['value'],1),'Mpaths/s', 'extend_ms',round(r['kernel_ms'],4),'shade_ms',round(r.get('shade_kernel_ms',0),4))"
  done
done
