#!/bin/bash
# plain bench run, then the ncu launch list of the same command (model set-up launches skipped)
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err || { echo "plain bench failed"; tail -3 gpurun_out/bench_short.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip ${1:-1200} --launch-count 1400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu launch list rc=$?"
wc -l gpurun_out/launches.csv
timeout 600 python bench.py --steps 3 --warmup 3 --workload c2 --no-cpu-baseline > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench c2 rc=$?"; tail -2 gpurun_out/bench_c2.err
