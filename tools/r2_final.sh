#!/bin/bash
# Round-end measurements on one B200: launch list of the bench command + full ncu captures of the four hot kernels at 128 spp
# (full 64-sample blocks, full pool), each only after the plain command has run.   usage: tools/r2_final.sh <tag>
tag=${1:-r02c}
python bench.py --steps 1 --warmup 3 --spp 128 --no-cpu --no-bvh --no-extra > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${tag}_launches_bench.csv python bench.py --steps 1 --warmup 3 --spp 128 --no-cpu --no-bvh --no-extra > gpurun_out/${tag}_ncu_launch.log 2>&1; echo "launch list rc=$?"
SKIP=60 tools/r2_ncu.sh ${tag}_extend 'wf_extend_kernel' random_ned
SKIP=60 tools/r2_ncu.sh ${tag}_shade wf_shade random_ned
SKIP=60 tools/r2_ncu.sh ${tag}_list 'wf_extend_list' cornell_smoke_ned
SKIP=60 tools/r2_ncu.sh ${tag}_shade_smoke wf_shade cornell_smoke_ned
timeout 600 ncu --set full --clock-control none --import-source on -k regex:trace_persistent_kernel -c 4 -o gpurun_out/${tag}_trace -f python tools/prof_trace.py --rays 22 > gpurun_out/${tag}_ncu_trace.log 2>&1; echo "trace capture rc=$?"
