#!/bin/bash
for s in c1 c2; do SVR2_ATTN_IMPL=v6 timeout 120 python scripts/attn_v5_prof.py $s 2>&1 | tail -6; done
