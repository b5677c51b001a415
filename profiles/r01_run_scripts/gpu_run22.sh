#!/bin/bash
# Run 22: joined vid+txt rows in shared-weight layers, affine modnorm for the output norm; suite + bench C1 (+ C2 short)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x -s > gpurun_out/t_gpu.log 2>&1; echo "gpu tests rc=$?"; tail -3 gpurun_out/t_gpu.log; grep -E "PSNR|block vid" gpurun_out/t_gpu.log | head -20
timeout 600 python bench.py > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "bench c1 rc=$?"; tail -2 gpurun_out/bench_c1.err
cat gpurun_out/bench_c1.json
timeout 600 python bench.py --steps 3 --warmup 3 --workload c2 --no-cpu-baseline > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench c2 rc=$?"; tail -2 gpurun_out/bench_c2.err
cat gpurun_out/bench_c2.json
