#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "window_attn" > gpurun_out/t_attn.log 2>&1; echo "attn tests rc=$?"; tail -3 gpurun_out/t_attn.log
SVR2_ATTN_IMPL=v4 timeout 300 python -m pytest tests/test_gpu_block.py -m gpu -q -x 2>&1 | tail -2
SVR2_ATTN_IMPL=v4 timeout 120 python scripts/attn_ab.py gen6 2>&1 | tail -1
timeout 120 python scripts/attn_ab.py gen4 2>&1 | tail -1
SVR2_ATTN_IMPL=v4 timeout 120 python scripts/attn_ab.py gen6-again 2>&1 | tail -1
for cfg in "c2 swin" "c1 swin" "c3 swin"; do SVR2_ATTN_IMPL=v4 timeout 200 python scripts/attn_stress.py $cfg 300 2>&1 | grep -E "svr2:|ok|differs|rror" | head -3; done
