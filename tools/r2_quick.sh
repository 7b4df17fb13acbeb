#!/bin/bash
# One GPU call: parity tests, then render throughput of the three scene classes.   usage: tools/r2_quick.sh <tag> [pytest args]
tag=${1:-x}; shift
python -m pytest tests -m gpu -x -q "$@" > gpurun_out/${tag}_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/${tag}_tests.log
for r in ${REFILLS:-default}; do
  [ $r = default ] && unset RTOW_WF_REFILL || export RTOW_WF_REFILL=$r
  echo "== RTOW_WF_REFILL=$r"
  python tools/prof_render.py --scene random_ned --spp ${SPP:-256} --reps 2 2>&1 | tee -a gpurun_out/${tag}_render.log
done
unset RTOW_WF_REFILL
for sc in cornell_ned cornell_smoke_ned; do
  python tools/prof_render.py --scene $sc --spp 128 --reps 2 2>&1 | tee -a gpurun_out/${tag}_render.log
done
python tools/prof_trace.py --rays 22 --reps 5 2>&1 | tee gpurun_out/${tag}_trace.log
