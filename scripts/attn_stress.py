"""Repeated launches of the window-attention kernel (timing-dependent protocol bugs): python scripts/attn_stress.py c2 swin 300"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import seedvr2_b200 as svr2
grids = {"c1": ((1, 128, 128), 20), "c2": ((9, 90, 160), 20), "c3": ((31, 68, 120), 24)}
name = sys.argv[1]; shifted = sys.argv[2] == "swin"; n = int(sys.argv[3])
grid, heads = grids[name]
dev = torch.device("cuda:0")
plan = svr2.get_plan(grid, (4, 3, 3), shifted, 58)
g = torch.Generator(device="cuda").manual_seed(0)
q, k, v = (torch.randn(plan.packed_rows, heads, 128, generator=g, device=dev).bfloat16() for _ in range(3))
ref = None
for it in range(n):
    out = svr2.ops.window_attn(plan, q, k, v, heads)
    if it % 25 == 0:
        torch.cuda.synchronize()
        o = out[0] if isinstance(out, (tuple, list)) else out
        if ref is None: ref = o.clone()
        elif not torch.equal(ref, o): print(f"launch {it}: output differs from launch 0 (max abs {(ref.float() - o.float()).abs().max().item():.4g})")
torch.cuda.synchronize()
print(f"{name} shifted={shifted}: {n} launches ok")
