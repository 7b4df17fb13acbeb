#!/bin/bash
# refill threshold sweep, then the bench line and ncu captures with the job-size-dependent stream length (1024 spp -> 2^25 entries)
for rep in 1 2; do for r in 2 4 6 8; do echo -n "refill $r: "; RTOW_WF_REFILL=$r python tools/prof_render.py --scene random_ned --spp 256 --reps 3 | head -1; done; done
python bench.py > gpurun_out/r02l_bench_n1.json 2> gpurun_out/r02l_bench_n1.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/r02l_bench_n1.json
SPP=1024 SKIP=60 tools/r2_ncu.sh r02l_extend 'wf_extend_kernel' random_ned
SPP=1024 SKIP=60 tools/r2_ncu.sh r02l_shade wf_shade random_ned
