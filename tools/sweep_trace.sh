# config-5 tuning sweep of the persistent closest-hit kernel: leaf vote threshold x refill threshold x blocks per SM
for cfg in "8 28 5" "6 28 5" "10 28 5" "8 24 5" "8 30 5" "8 28 4" "8 28 6"; do set -- $cfg; echo "LEAFMIN=$1 REFILL=$2 MINB=$3"; RTOW_TRACE_LEAFMIN=$1 RTOW_TRACE_REFILL=$2 RTOW_TRACE_MINB=$3 python tools/prof_trace.py --rays 22 --reps 5; done
