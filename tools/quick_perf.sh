#!/bin/bash
# One GPU call: parity tests, config-5 closest-hit throughput, short render benches of the three scenes.
# usage: tools/quick_perf.sh <tag>
tag=${1:-x}
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${tag}_tests.log
python tools/prof_trace.py --rays 22 --reps 5 2>&1 | tee gpurun_out/${tag}_trace.log
for sc in random_ned cornell_ned cornell_smoke_ned; do
  python bench.py --steps 2 --warmup 3 --spp 128 --no-cpu --no-bvh --scene $sc 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$sc', round(d['value'],1),'Mpaths/s', 'extend_ms',round(r['kernel_ms'],4),'shade_ms',round(r.get('shade_kernel_ms',0),4))" 2>&1 | tee -a gpurun_out/${tag}_render.log
done
