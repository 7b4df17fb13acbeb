#!/bin/bash
# Builds an alternative librtow_b200 with extra nvcc flags, for A/B kernel experiments on one GPU box:
#   tools/build_variant.sh noprefetch -DRTOW_NO_PREFETCH
#   RTOW_B200_LIB=$PWD/ray-tracing-one-weekend_b200/csrc/variants/librtow_b200_noprefetch.so python tools/prof_trace.py
set -e
name=$1; shift
cd "$(dirname "$0")/../ray-tracing-one-weekend_b200/csrc"
mkdir -p variants/obj_$name
A="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -ffp-contract=off"
for u in kernels.cu wavefront.cu api.cu multi.cu; do nvcc $A -fmad=false "$@" -c $u -o variants/obj_$name/$u.o & done
for u in fast_bvh.cpp host_scenes.cpp errors.cpp; do nvcc $A "$@" -c $u -o variants/obj_$name/$u.o & done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o variants/librtow_b200_$name.so variants/obj_$name/*.o -lcudart
rm -rf variants/obj_$name
echo variants/librtow_b200_$name.so
