#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout 120 python scripts/prof_kernels.py c1 2>&1 | tail -5
timeout 600 python bench.py > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "bench c1 rc=$?"; tail -2 gpurun_out/bench_c1.err
python - <<PY
import json
d=json.load(open('gpurun_out/bench_c1.json'))
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, 'e2e', d['e2e']['value'], 'attn', round(d['roofline']['achieved'],1), round(d['roofline']['frac'],3), d['clocks'])
for k,v in d['kernels'].items(): print(k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items()})
PY
