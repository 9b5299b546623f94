#!/bin/bash
# Experimental build of the library with extra compiler switches:  tools/build_exp.sh <name> [-DHUT_...=...]
# -> hutoken_b200/exp_<name>.so (git-ignored; picked up by tools/ab_libs.sh through HUTOKEN_B200_LIB)
set -e
cd "$(dirname "$0")/../hutoken_b200/csrc"
name=$1; shift
mkdir -p _build
make -s _build/hut_model.o
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xptxas -v "$@" \
    -c hut_engine.cu -o _build/exp_$name.o 2> _build/exp_$name.ptxas
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o ../exp_$name.so _build/exp_$name.o _build/hut_model.o
grep -A1 "tile_kernelILb0" _build/exp_$name.ptxas | grep -o "Used [0-9]* registers\|[0-9]* bytes spill stores\|[0-9]* bytes smem" | tr '\n' ' '; echo " <- $name tile_kernel<false>"
