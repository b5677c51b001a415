"""Median launch time of the window-attention kernel at the BASELINE grids; SVR2_LIB selects an experimental build.
python scripts/attn_ab.py [label]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import seedvr2_b200 as svr2
label = sys.argv[1] if len(sys.argv) > 1 else "default"
impl = int(sys.argv[2]) if len(sys.argv) > 2 else svr2.ops.ATTN_AUTO      # SVR2_ATTN_* code: 4, 8, 10 or 0 = what the library picks
dev = torch.device("cuda:0")
out = []
for name, grid, heads, shifted in (("c1", (1, 128, 128), 20, False), ("c1s", (1, 128, 128), 20, True), ("c2", (9, 90, 160), 20, False),
                                   ("v1080", (9, 68, 120), 20, False), ("v1080s", (9, 68, 120), 20, True), ("c3", (31, 68, 120), 24, False)):
    plan = svr2.get_plan(grid, (4, 3, 3), shifted, 58)
    g = torch.Generator(device="cuda").manual_seed(0)
    q, k, v = (torch.randn(plan.packed_rows, heads, 128, generator=g, device=dev).bfloat16() for _ in range(3))
    for _ in range(3):
        svr2.ops.window_attn(plan, q, k, v, heads, impl=impl)
    timer = svr2.ops.KernelTimer(only={"svr2_window_attn_fwd"})
    svr2.ops.set_timer(timer)
    for _ in range(15):
        svr2.ops.window_attn(plan, q, k, v, heads, impl=impl)
    torch.cuda.synchronize()
    svr2.ops.set_timer(None)
    n, ms = timer.summary()["svr2_window_attn_fwd"]
    us = ms / n * 1e3
    out.append(f"{name} {us:7.1f} us {plan.attention_flops(heads) / us / 1e6:5.0f} TF/s")
print(f"{label:28s} " + " | ".join(out))
