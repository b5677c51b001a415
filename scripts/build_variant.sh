#!/bin/bash
# build_variant.sh <name> <source.cu> <nvcc -D flags...> : experimental build of one kernel file -> csrc/build/variants/lib_<name>.so
# (all other objects come from the regular build; select the library with SVR2_LIB=<path>)
set -e
name=$1; src=$2; shift 2
C=$(dirname $0)/../comfyui-seedvr2_b200/csrc
mkdir -p $C/build/variants
nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC --expt-relaxed-constexpr -Xptxas -v -I$C "$@" -c $C/$src -o $C/build/variants/$name.o 2> $C/build/variants/$name.log
objs=$(ls $C/build/*.o | grep -v "/${src%.cu}.o")
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o $C/build/variants/lib_$name.so $objs $C/build/variants/$name.o -lcudart
echo "$name: $(grep -A1 'ILb0' $C/build/variants/$name.log | grep spill)"
