#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "qkv_pack" 2>&1 | tail -3
echo "--- staged"; timeout 120 python scripts/prof_kernels.py 2>&1 | tail -1
echo "--- v1"; SVR2_QKV_PACK=v1 timeout 120 python scripts/prof_kernels.py 2>&1 | tail -1
