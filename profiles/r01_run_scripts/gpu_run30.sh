#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "modnorm or gate_residual" 2>&1 | tail -2
for s in c1 c2; do timeout 120 python scripts/prof_kernels.py $s 2>&1 | tail -5; done
echo "--- qkv_pack with __launch_bounds__(256, 4)"
for s in c1 c2; do SVR2_LIB=comfyui-seedvr2_b200/csrc/build/variants/lib_pack_minb4.so timeout 120 python scripts/prof_kernels.py $s 2>&1 | grep qkv_pack; done
