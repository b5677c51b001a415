#!/bin/bash
# Run 21: whole GPU suite (incl. the new whole-NaDiT parity tests), default bench at C1, launch list of the bench command.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x -s > gpurun_out/t_gpu.log 2>&1; echo "gpu tests rc=$?"; tail -3 gpurun_out/t_gpu.log; grep -E "PSNR|block vid" gpurun_out/t_gpu.log | head -20
timeout 600 python bench.py > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "bench c1 rc=$?"; tail -2 gpurun_out/bench_c1.err
cat gpurun_out/bench_c1.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 1200 --launch-count 1400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu launch list rc=$?"
wc -l gpurun_out/launches.csv
