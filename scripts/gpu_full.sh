#!/bin/bash
# Round-end style run: the GPU test suite, the bench at C1 and C2, then the ncu launch list of the bench command.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/t_gpu.log 2>&1; echo "gpu tests rc=$?"; tail -3 gpurun_out/t_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "bench c1 rc=$?"
timeout 600 python bench.py --steps 3 --warmup 3 --workload c2 --no-cpu-baseline > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench c2 rc=$?"
python - <<'PY'
import json
for n in ('c1','c2'):
    try:
        d=json.load(open(f'gpurun_out/bench_{n}.json'))
        print(n, {k:d[k] for k in ('value','ms_per_step','gpu_launches')}, 'roofline', d['roofline'], 'e2e', d['e2e']['value'], 'cpu', d.get('cpu_baseline'), d['clocks'])
    except Exception as e: print(n,'bench parse failed', e)
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu launch list rc=$?"
