#!/bin/bash
# config 5: refill threshold / leaf vote threshold of trace_persistent_kernel (env knobs are read once per process)
for r in 8 16 24 32; do echo "== RTOW_TRACE_REFILL=$r"; RTOW_TRACE_REFILL=$r python tools/prof_trace.py --rays 22 --reps 3; done
for l in 4 16; do echo "== RTOW_TRACE_LEAFMIN=$l"; RTOW_TRACE_LEAFMIN=$l python tools/prof_trace.py --rays 22 --reps 3; done
