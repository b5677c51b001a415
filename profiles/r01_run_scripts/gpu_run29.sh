#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "modnorm or gate_residual or qkv_pack or generations_bit" 2>&1 | tail -3
timeout 600 python -m pytest tests/test_gpu_block.py tests/test_gpu_stack.py -m gpu -q -x 2>&1 | tail -2
for s in c1 c2; do timeout 120 python scripts/prof_kernels.py $s 2>&1 | tail -5; done
