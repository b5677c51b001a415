#!/bin/bash
# Run 43: stress of generation 8 (repeated launches, all BASELINE grids incl. shifted), then the whole GPU suite with the new dispatch
mkdir -p gpurun_out
export SVR2_ATTN_IMPL=v6
for a in "c1 win 400" "c1 swin 400" "c2 swin 200" "c3 swin 60" "c3 win 60"; do timeout 300 python scripts/attn_stress.py $a 2>&1 | tail -2; done
unset SVR2_ATTN_IMPL
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -2
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
