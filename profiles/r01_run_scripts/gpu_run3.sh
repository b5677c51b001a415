#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/t_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -15 gpurun_out/t_gpu.log
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "bench c1 rc=$?"; tail -c 2200 gpurun_out/bench_c1.json; tail -5 gpurun_out/bench_c1.err
