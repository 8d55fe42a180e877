#!/bin/bash
# An experimental build of the CUDA library next to the product one:  tools/build_variant.sh NAME [-DMACRO=..]...
# -> exp/libmbgfe_NAME.so (git-ignored, travels to the GPU box); run it with tools/profile_step.py --library exp/libmbgfe_NAME.so
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p exp
[ -f verkko_b200/hostpack.o ] || g++ -O3 -std=c++17 -fPIC -pthread -c -o verkko_b200/hostpack.o verkko_b200/csrc/hostpack.cpp
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -ccbin /usr/bin/g++ --shared "$@" \
    -o exp/libmbgfe_$name.so verkko_b200/csrc/mbgfe.cu verkko_b200/hostpack.o -lz
echo exp/libmbgfe_$name.so
