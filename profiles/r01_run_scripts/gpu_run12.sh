#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_block.py -m gpu -q -x -k "window_attn or block" > gpurun_out/t_attn.log 2>&1; echo "attn+block tests rc=$?"; tail -3 gpurun_out/t_attn.log
timeout 120 python scripts/attn_ab.py gen5 2>&1 | tail -1
SVR2_ATTN_IMPL=v2 timeout 120 python scripts/attn_ab.py gen4 2>&1 | tail -1
timeout 120 python scripts/attn_ab.py gen5-again 2>&1 | tail -1
timeout 120 python scripts/attn_trace.py c1 2>&1 | grep -E "TFLOP|item 2|item 3"
