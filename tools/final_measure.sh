#!/bin/bash
# Round-end measurements on one B200: tests, both bench arms, ncu launch list of the bench command, full ncu captures.
tag=${1:-r01f}
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2> gpurun_out/${tag}_bench_reference.err; echo "ref rc=$?"
python bench.py > gpurun_out/${tag}_bench_n1.json 2> gpurun_out/${tag}_bench_n1.err; echo "bench rc=$?"
cut -c1-400 gpurun_out/${tag}_bench_n1.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_bench.csv python bench.py --steps 1 --warmup 3 --spp 16 --no-cpu --no-bvh > gpurun_out/${tag}_ncu_launch.log 2>&1; echo "launch list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wf_ -s 40 -c 4 -o gpurun_out/${tag}_wf -f python bench.py --steps 1 --warmup 1 --spp 16 --no-cpu --no-bvh > gpurun_out/${tag}_ncu_wf.log 2>&1; echo "wf capture rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wf_extend -s 20 -c 1 -o gpurun_out/${tag}_wfext -f python bench.py --steps 1 --warmup 1 --spp 16 --no-cpu --no-bvh > gpurun_out/${tag}_ncu_wfext.log 2>&1; echo "extend capture rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:trace_persistent_kernel -c 5 -o gpurun_out/${tag}_trace -f python tools/prof_trace.py --rays 22 > gpurun_out/${tag}_ncu_trace.log 2>&1; echo "trace capture rc=$?"
