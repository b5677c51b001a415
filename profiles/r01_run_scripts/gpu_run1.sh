mkdir -p gpurun_out
nvidia-smi > gpurun_out/smi.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "not window_attn" > gpurun_out/t1_kernels.log 2>&1; echo "t1 rc=$?"
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "window_attn_vs_fp32 and mma" > gpurun_out/t2_mma.log 2>&1; echo "t2 rc=$?"
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "window_attn_vs_fp32 and tcgen05" > gpurun_out/t3_tc.log 2>&1; echo "t3 rc=$?"
SVR2_ATTN_IMPL=mma timeout 300 python -m pytest tests/test_gpu_block.py -m gpu -q -s > gpurun_out/t4_block_mma.log 2>&1; echo "t4 rc=$?"
timeout 300 python -m pytest tests/test_gpu_block.py -m gpu -q -s > gpurun_out/t5_block_tc.log 2>&1; echo "t5 rc=$?"
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "full_size" > gpurun_out/t6_full.log 2>&1; echo "t6 rc=$?"
tail -5 gpurun_out/t*.log
