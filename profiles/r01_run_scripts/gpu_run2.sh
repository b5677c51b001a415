#!/bin/bash
# round 1, GPU run 2: first bench line + ncu launch list + ncu full capture of the attention kernel
mkdir -p gpurun_out
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "bench c1 rc=$?"; tail -c 3000 gpurun_out/bench_c1.json; tail -5 gpurun_out/bench_c1.err
python bench.py --steps 2 --warmup 3 > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 2500 -c 1200 --csv --log-file gpurun_out/launches_c1.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
python bench.py --steps 2 --warmup 3 > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:window_attn_tc -s 6 -c 2 -o gpurun_out/attn_c1 python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
python bench.py --workload c2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench c2 rc=$?"; tail -c 2500 gpurun_out/bench_c2.json; tail -5 gpurun_out/bench_c2.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_c1.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref_c1.json
nproc; free -g | head -2
