#!/bin/bash
# tools/build_variant.sh <tag> <extra nvcc flags...>: builds learn_comms_b200/liblearncomms_<tag>.so for A/B runs (LEARN_COMMS_LIB)
set -e
tag=$1; shift
cd "$(dirname "$0")/../learn_comms_b200/csrc"
F="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -prec-div=true -prec-sqrt=true -ftz=false -Xcompiler -fPIC,-ffp-contract=off,-fno-fast-math,-O2"
mkdir -p /tmp/lcv_$tag
for s in crowd.cu capi.cu batch.cu walls.cu peak.cu splat.cu host.cpp planner.cpp; do nvcc $F "$@" -x cu -c $s -o /tmp/lcv_$tag/${s%.*}.o & done; wait
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o ../liblearncomms_$tag.so /tmp/lcv_$tag/*.o
