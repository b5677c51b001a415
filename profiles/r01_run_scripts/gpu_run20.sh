#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_block.py -m gpu -q -x -k "not window_attn" 2>&1 | tail -3
timeout 120 python scripts/prof_kernels.py 2>&1 | tail -4
