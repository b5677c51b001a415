#!/bin/bash
# build_variant.sh <name> <nvcc -D flags...> : experimental build of the attention kernel -> csrc/build/variants/lib_<name>.so
set -e
C=/root/repo/comfyui-seedvr2_b200/csrc
name=$1; shift
mkdir -p $C/build/variants
nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC --expt-relaxed-constexpr -Xptxas -v -I$C "$@" -c $C/window_attn_v2.cu -o $C/build/variants/$name.o 2> $C/build/variants/$name.log
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o $C/build/variants/lib_$name.so $C/build/api.o $C/build/window_index.o $C/build/adaln.o $C/build/qkv_pack.o $C/build/window_attn_mma.o $C/build/window_attn.o $C/build/window_attn_v3.o $C/build/window_attn_v4.o $C/build/window_attn_v5.o $C/build/variants/$name.o $C/build/halo_blend.o -lcudart
echo "$name: $(grep -A1 'ILb0' $C/build/variants/$name.log | grep spill)"
