#!/bin/bash
# scratch/build_variant.sh NAME "-DLSG_JT_THREADS=128 ..." : the library with other compile-time constants of join.cu,
# as scratch/variants/libv_NAME.so (select it with LLMSLO_GPU_LIB=...); the other objects come from the normal build.
set -e
cd "$(dirname "$0")/../llm-slo-ebpf-toolkit_b200/csrc"
make -s
mkdir -p ../../scratch/variants build
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-Wall -Xptxas -v --fmad=false $2 -c join.cu -o build/join_v_$1.o 2> build/join_v_$1.ptxas.log
objs=$(ls build/*.o | grep -v "join.o" | grep -v "join_v_")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../scratch/variants/libv_$1.so $objs build/join_v_$1.o -ldl
grep -A2 "k_join_tilesILi3ELb1ELb1E" build/join_v_$1.ptxas.log | grep -i "spill\|Used" 
