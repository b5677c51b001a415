"""Per-CTA timeline of the persistent window-attention kernel (debug hook svr2_debug_attn_trace):
python scripts/attn_trace.py [c1|c2|c3] [swin]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import seedvr2_b200 as svr2

grids = {"c1": ((1, 128, 128), 20), "c2": ((9, 90, 160), 20), "c3": ((31, 68, 120), 24)}
name = sys.argv[1] if len(sys.argv) > 1 else "c1"
shifted = len(sys.argv) > 2 and sys.argv[2] == "swin"
grid, heads = grids[name]
dev = torch.device("cuda:0")
plan = svr2.get_plan(grid, (4, 3, 3), shifted, 58)
P = plan.packed_rows
g = torch.Generator(device="cuda").manual_seed(0)
q, k, v = (torch.randn(P, heads, 128, generator=g, device=dev).bfloat16() for _ in range(3))
for _ in range(20):      # let the clocks settle
    svr2.ops.window_attn(plan, q, k, v, heads)
torch.cuda.synchronize()
n_ctas = 148
L = svr2._lib.lib()
slots = L.svr2_debug_attn_trace(None, 0)
buf = torch.zeros(n_ctas, slots, dtype=torch.int64, device=dev)
L.svr2_debug_attn_trace(buf.data_ptr(), n_ctas)
timer = svr2.ops.KernelTimer(only={"svr2_window_attn_fwd"})
svr2.ops.set_timer(timer)
svr2.ops.window_attn(plan, q, k, v, heads)
torch.cuda.synchronize()
svr2.ops.set_timer(None)
L.svr2_debug_attn_trace(None, 0)
ms = timer.summary()["svr2_window_attn_fwd"][1]
fl = plan.attention_flops(heads)
print(f"{name} shifted={shifted}: {plan.work.shape[0] * heads} items, kernel {ms*1e3:.1f} us, {fl/ms/1e9:.0f} TFLOP/s (with trace stores)")
t = buf.cpu().numpy().astype(np.float64)
ent = t[:, 0]
print(f"setup done {np.median(t[:,1]-ent):.0f} | CTA end {np.median(t[:,5]-ent):.0f} cycles | items per CTA median {np.median(t[:,7]):.0f}")
names = ["MMA start", "S0(0) seen", "P0(0) pub", "S0(last) seen", "P0(last) pub", "O0 final seen", "tile0 stored", "tile1 stored", "TMA last KV issued"]
for n in range(4):
    base = 8 + 9 * n
    st = t[:, base]
    ok = st > 0
    if not ok.any():
        continue
    line = f"item {n}: start @ {np.median(st[ok]-ent[ok]):.0f}"
    if n > 0:
        prev = t[:, base - 9]
        line += f" (period {np.median((st-prev)[ok]):.0f})"
    for w in range(1, 9):
        x = t[:, base + w]
        m = ok & (x > 0)
        if m.any():
            line += f" | {names[w]} {np.median((x-st)[m]):+.0f}"
    print(line)


f = t[:, 44:60]
ok = (f[:, :4] > 0).all(axis=1)
if ok.any():
    ref = t[ok, 8 + 9 * 2]           # item 2: MMA start
    for i in range(2):
        parts = []
        for j in range(4):
            a, b = t[ok, 44 + 8 * i + 2 * j], t[ok, 45 + 8 * i + 2 * j]
            m = (a > 0) & (b > 0)
            if m.any():
                parts.append(f"j={j}: S seen {np.median((a - ref)[m]):+.0f}, P published {np.median((b - ref)[m]):+.0f} (softmax {np.median((b - a)[m]):.0f})")
        print(f"item 2, tile {i}: " + " | ".join(parts))
