#!/bin/bash
# Run 64 (final state of the round): whole GPU suite, smoke, default bench (C1 incl. cpu_baseline), reference arm
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x -s > gpurun_out/t_gpu.log 2>&1; echo "gpu tests rc=$?"; tail -1 gpurun_out/t_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "bench c1 rc=$?"; tail -2 gpurun_out/bench_c1.err
cat gpurun_out/bench_c1.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "reference arm rc=$?"; cut -c1-200 gpurun_out/bench_ref.json
