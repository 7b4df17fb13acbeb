#!/bin/bash
# ncu --set full capture of steady-state launches of a wavefront kernel.  usage: tools/r2_ncu.sh <tag> <kernel regex> <scene>
tag=$1; kern=$2; scene=${3:-random_ned}
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$kern -s ${SKIP:-60} -c ${COUNT:-1} -o gpurun_out/${tag} -f \
  python tools/prof_render.py --scene $scene --spp ${SPP:-128} --reps 1 > gpurun_out/${tag}.log 2>&1; echo "ncu $tag rc=$?"
