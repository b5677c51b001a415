#!/bin/bash
# multi-GPU check: NCCL halo exchange + the bench at N ranks
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L | head -8
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scripts/nccl_segments_check.py 2>&1 | grep -vE "^W[0-9]|OMP_NUM|\*\*\*" | tail -8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c1_n$N.json 2> gpurun_out/bench_c1_n$N.err; echo "bench N=$N rc=$?"; tail -3 gpurun_out/bench_c1_n$N.err
python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/bench_c1_n$N.json') if l.startswith('{')][-1])
    print({k:d[k] for k in ('value','n_gpus','ms_per_step','scaling','gpu_launches')}, d['e2e']['value'], d['clocks'])
except Exception as e: print('parse failed', e)
PY
