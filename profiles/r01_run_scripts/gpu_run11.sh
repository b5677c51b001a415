#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_block.py -m gpu -q -x -k "window_attn or block" > gpurun_out/t_attn.log 2>&1; echo "attn+block tests rc=$?"; tail -3 gpurun_out/t_attn.log
bash scripts/gpu_ab.sh 2>&1 | tee gpurun_out/ab.txt
python scripts/attn_trace.py c1 2>&1 | grep -E "TFLOP|item 2|item 3"
