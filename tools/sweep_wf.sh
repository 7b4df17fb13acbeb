# wavefront knobs on the render scenes (CUDA-event timings printed by tools/prof_render.py)
for s in random_ned cornell_smoke_ned; do
  for cfg in "1048576 5" "524288 5" "262144 5" "1048576 8" "524288 8"; do set -- $cfg
    echo "POOL=$1 MINB=$2"; RTOW_WF_POOL=$1 RTOW_WF_MINB=$2 python tools/prof_render.py --scene $s --spp 64 2>&1 | grep -v Trace
  done
done
