#!/bin/bash
# A/B builds of librrt_b200.so with extra -D flags: tools/ab_build.sh name "-DRRT_X=1 ..."  ->  build_ab/lib_<name>.so
# (select at run time with RRT_B200_LIB=build_ab/lib_<name>.so)
set -e
cd "$(dirname "$0")/../relativistic-ray-tracer_b200/csrc"
name=$1; shift
NV="/usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -Xcompiler -fPIC,-O2 $*"
mkdir -p ../../build_ab
$NV -Xptxas -v -c rrt_api.cu -o ../../build_ab/api_$name.o 2> ../../build_ab/ptxas_$name.log
[ -f rrt_bvh.o ] || $NV -c rrt_bvh.cpp -o rrt_bvh.o
/usr/local/cuda/bin/nvcc -shared -gencode arch=compute_100a,code=sm_100a -o ../../build_ab/lib_$name.so ../../build_ab/api_$name.o rrt_bvh.o -lcudart_static -lpthread -ldl -lrt
rm -f ../../build_ab/api_$name.o
echo built build_ab/lib_$name.so
