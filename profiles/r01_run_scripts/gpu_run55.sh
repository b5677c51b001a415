#!/bin/bash
# Run 55 (round-end style, after the C-ABI change for the 4-tile work list): whole GPU suite, smoke, default bench (C1, incl. cpu_baseline), reference arm, C2 bench, ncu launch list
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x -s > gpurun_out/t_gpu.log 2>&1; echo "gpu tests rc=$?"; tail -1 gpurun_out/t_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "bench c1 rc=$?"; tail -2 gpurun_out/bench_c1.err
cat gpurun_out/bench_c1.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "reference arm rc=$?"; cat gpurun_out/bench_ref.json
timeout 600 python bench.py --steps 3 --warmup 3 --workload c2 --no-cpu-baseline > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench c2 rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 1000 --launch-count 1200 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu launch list rc=$?"
wc -l gpurun_out/launches.csv
