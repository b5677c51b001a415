"""Per-CTA timeline of the window-attention kernel (debug): python scripts/attn_trace.py [c1|c2] [swin]"""
import os, sys, ctypes
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
for _ in range(3):
    svr2.ops.window_attn(plan, q, k, v, heads)
torch.cuda.synchronize()
n_ctas = plan.work.shape[0] * heads
L = svr2._lib.lib()
slots = L.svr2_debug_attn_trace(None, 0)
buf = torch.zeros(n_ctas, slots, dtype=torch.int64, device=dev)
L.svr2_debug_attn_trace(buf.data_ptr(), n_ctas)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
svr2.ops.window_attn(plan, q, k, v, heads)
e1.record()
torch.cuda.synchronize()
L.svr2_debug_attn_trace(None, 0)
t = buf.cpu().numpy()
print(f"{name} shifted={shifted}: {n_ctas} CTAs, kernel+pool {e0.elapsed_time(e1)*1e3:.1f} us, flops {plan.attention_flops(heads)/1e9:.1f} G")
nkv = t[:, 7]
for nk in sorted(set(nkv.tolist())):
    m = nkv == nk
    tt = t[m]
    ent = tt[:, 0]
    def col(s):
        return (tt[:, s] - ent)
    print(f"-- n_kv={nk}: {m.sum()} CTAs;  median cycles since entry:")
    print(f"   setup done {np.median(col(1)):.0f} | Q+K0 landed {np.median(col(2)):.0f} | O final {np.median(col(3)):.0f} | stores issued {np.median(col(4)):.0f} | end {np.median(col(5)):.0f}")
    s_ready = [np.median(col(8 + j)) for j in range(min(nk, 16))]
    p_pub = [np.median(col(24 + j)) for j in range(min(nk, 16))]
    s1 = [np.median(col(40 + j)) for j in range(min(nk, 16))]
    print("   S0(j) ready:", [int(x) for x in s_ready])
    print("   P0(j) done :", [int(x) for x in p_pub])
    print("   S1(j) ready:", [int(x) for x in s1])
# per-SM occupancy: gaps between consecutive CTAs on one SM
sm = t[:, 6]
gaps = []
for s_ in np.unique(sm):
    rows = t[sm == s_]
    rows = rows[np.argsort(rows[:, 0])]
    gaps += list(rows[1:, 0] - rows[:-1, 5])
print(f"gap between a CTA's end and the next CTA's entry on the same SM: median {np.median(gaps):.0f} cycles, mean {np.mean(gaps):.0f}")
dur = t[:, 5] - t[:, 0]
print(f"CTA lifetime cycles: median {np.median(dur):.0f} mean {dur.mean():.0f}; CTAs per SM ~{n_ctas/148:.1f}")
