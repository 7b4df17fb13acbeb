python -m pytest tests/test_gpu_render.py -q -m gpu -k "medium_boundary or render_matches_oracle" 2>&1 | tail -2
SCENES="random_ned" tools/ab_render.sh 256 2
V=$PWD/ray-tracing-one-weekend_b200/csrc/variants/librtow_b200_amax.so
echo "== config5 default"; python tools/prof_trace.py --rays 22 --reps 3
echo "== config5 amax"; RTOW_B200_LIB=$V python tools/prof_trace.py --rays 22 --reps 3
for r in 4 12 16; do echo "== refill $r"; RTOW_WF_REFILL=$r python tools/prof_render.py --scene random_ned --spp 256 --reps 2 | head -1; done
for l in 8 16; do echo "== leafmin $l"; RTOW_TRACE_LEAFMIN=$l python tools/prof_render.py --scene random_ned --spp 256 --reps 2 | head -1; done
for p in 8388608 16777216 33554432; do echo "== pool $p"; RTOW_WF_POOL=$p python tools/prof_render.py --scene random_ned --spp 1024 --reps 2 | head -1; done
