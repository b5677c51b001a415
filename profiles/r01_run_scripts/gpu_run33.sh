#!/bin/bash
# Run 33 (2 GPUs): NCCL halo exchange parity + C3 stitch timing, bench at N=2; then C3 / C4 bench lines on one GPU
mkdir -p gpurun_out
bash scripts/gpu_multi.sh 2
for w in c3 c4; do
  timeout 900 python bench.py --steps 2 --warmup 3 --workload $w --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"; tail -1 gpurun_out/bench_$w.err
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/bench_$w.json'))
    print('$w', {k:round(d[k],3) for k in ('value','ms_per_step')}, 'e2e', round(d['e2e']['value'],2), 'attn', round(d['roofline']['achieved'],1), round(d['roofline']['frac'],3), d['clocks']['sm_mhz'])
except Exception as e: print('$w parse failed', e)
PY
done
