#!/bin/bash
# Development aid: liblocust.so with map_kernel.cu compiled under extra -D flags, written to variants/<name>/liblocust.so
# (git-ignored, travels to the GPU box); run anything against it with LOCUST_LIB=variants/<name>/liblocust.so.
#   tools/build_variant.sh noseltab -DLOCUST_SELTAB=0
set -e
cd "$(dirname "$0")/../locust_b200/csrc"
name=$1; shift
mkdir -p ../../variants/$name
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xptxas -v "$@" \
    -c map_kernel.cu -o ../../variants/$name/map_kernel.o 2> ../../variants/$name/ptxas.log
objs=$(ls build/*.o | grep -v map_kernel.o)
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../variants/$name/liblocust.so ../../variants/$name/map_kernel.o $objs
grep -A2 'tokenize_combine_kernelILb1' ../../variants/$name/ptxas.log | grep -E 'spill|registers' | tr '\n' ' '; echo " <- $name"
