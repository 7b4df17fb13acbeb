#!/bin/bash
# final checks with refill-below-2 and the job-size-dependent stream length: refill 1 / 2 / 3, tests, bench line, ncu captures at the bench's spp
for r in 1 2 3; do echo -n "refill $r: "; RTOW_WF_REFILL=$r python tools/prof_render.py --scene random_ned --spp 256 --reps 3 | head -1; done
python -m pytest tests -m gpu -x -q > gpurun_out/r02m_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/r02m_tests.log
python bench.py > gpurun_out/r02m_bench_n1.json 2> gpurun_out/r02m_bench_n1.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/r02m_bench_n1.json
SPP=1024 SKIP=60 tools/r2_ncu.sh r02m_extend 'wf_extend_kernel' random_ned
SPP=1024 SKIP=60 tools/r2_ncu.sh r02m_shade wf_shade random_ned
