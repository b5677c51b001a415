#!/bin/bash
# Run 51 (NP GPUs, default 2): temporal-segment stitch over NCCL and as one peer-memory kernel per rank
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node ${NP:-2} --master-addr 127.0.0.1 --master-port 29511 scripts/nccl_segments_check.py 2>&1 | grep -v "^W\|^\*\*\*" | tail -40
