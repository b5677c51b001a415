"""Three launches of the window-attention kernel at a BASELINE grid (for ncu): python scripts/prof_attn.py [c1|c2] [swin]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import seedvr2_b200 as svr2
grids = {"v1080": ((9, 68, 120), 20), "c1": ((1, 128, 128), 20), "c2": ((9, 90, 160), 20), "c3": ((31, 68, 120), 24)}
name = sys.argv[1] if len(sys.argv) > 1 else "c1"
shifted = len(sys.argv) > 2 and sys.argv[2] == "swin"
grid, heads = grids[name]
dev = torch.device("cuda:0")
plan = svr2.get_plan(grid, (4, 3, 3), shifted, 58)
g = torch.Generator(device="cuda").manual_seed(0)
q, k, v = (torch.randn(plan.packed_rows, heads, 128, generator=g, device=dev).bfloat16() for _ in range(3))
timer = svr2.ops.KernelTimer(only={"svr2_window_attn_fwd"})
svr2.ops.set_timer(timer)
for _ in range(3):
    svr2.ops.window_attn(plan, q, k, v, heads)
torch.cuda.synchronize()
n, ms = timer.summary()["svr2_window_attn_fwd"]
print(f"{name}: {ms / n * 1e3:.1f} us per launch, {plan.attention_flops(heads) * n / ms / 1e9:.0f} TFLOP/s")
