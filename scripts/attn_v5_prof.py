"""Where generation 7 (SVR2_ATTN_IMPL=v5) spends its cycles: per-role wait counters of CTA 0 (kProf instantiation, armed
through svr2_debug_attn_trace).  python scripts/attn_v5_prof.py [c1|c2|c3] [swin]"""
import os, sys
os.environ.setdefault("SVR2_ATTN_IMPL", "v5")      # v5 (generation 7) or v6 (generation 8: wg rows of CTA 1 follow)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import seedvr2_b200 as svr2

grids = {"c1": ((1, 128, 128), 20), "c2": ((9, 90, 160), 20), "c3": ((31, 68, 120), 24)}
name = sys.argv[1] if len(sys.argv) > 1 else "c1"
shifted = len(sys.argv) > 2 and sys.argv[2] == "swin"
grid, heads = grids[name]
dev = torch.device("cuda:0")
plan = svr2.get_plan(grid, (4, 3, 3), shifted, 58)
g = torch.Generator(device="cuda").manual_seed(0)
q, k, v = (torch.randn(plan.packed_rows, heads, 128, generator=g, device=dev).bfloat16() for _ in range(3))
for _ in range(10):
    svr2.ops.window_attn(plan, q, k, v, heads)
torch.cuda.synchronize()
L = svr2._lib.lib()
buf = torch.zeros(64, dtype=torch.int64, device=dev)
L.svr2_debug_attn_trace(buf.data_ptr(), 1)
timer = svr2.ops.KernelTimer(only={"svr2_window_attn_fwd"})
svr2.ops.set_timer(timer)
svr2.ops.window_attn(plan, q, k, v, heads)
torch.cuda.synchronize()
svr2.ops.set_timer(None)
L.svr2_debug_attn_trace(None, 0)
ms = timer.summary()["svr2_window_attn_fwd"][1]
t = buf.cpu().tolist()
print(f"{name} shifted={shifted}: kernel {ms * 1e3:.1f} us (profiling instantiation), {plan.attention_flops(heads) / ms / 1e9:.0f} TFLOP/s")
for wg in (0, 1):
    a = t[8 * wg: 8 * wg + 8]
    steps, tot = max(a[0], 1), max(a[1], 1)
    busy = tot - sum(a[2:7])
    print(f"softmax warpgroup {wg}: {a[0]} steps, {tot} cycles = {tot / steps:.0f} per step | waits per step: S full {a[2] / steps:.0f}, "
          f"m(g-1) {a[3] / steps:.0f}, P buffer free {a[4] / steps:.0f}, l(g-1) {a[5] / steps:.0f}, O drained {a[6] / steps:.0f} | "
          f"working {busy / steps:.0f} per step ({100 * busy / tot:.0f} %) | rescales {a[7]}")
if os.environ["SVR2_ATTN_IMPL"] == "v6":
    for wg in (0, 1):
        a = t[24 + 8 * wg: 32 + 8 * wg]
        steps, tot = max(a[0], 1), max(a[1], 1)
        busy = tot - sum(a[2:7])
        print(f"CTA 1 softmax warpgroup {wg}: {a[0]} steps, {tot / steps:.0f} per step | waits: S full {a[2] / steps:.0f}, m {a[3] / steps:.0f}, "
              f"P buffer free {a[4] / steps:.0f}, l {a[5] / steps:.0f}, O drained {a[6] / steps:.0f} | working {busy / steps:.0f}")
if os.environ["SVR2_ATTN_IMPL"] == "v6":
    for wg in (0, 1):
        ph, steps = t[40 + 4 * wg: 44 + 4 * wg], max(t[8 * wg], 1)
        print(f"CTA 0 warpgroup {wg} phases per step: ld S {ph[0] / steps:.0f} | mask + row max {ph[1] / steps:.0f} | exp2 + pack + st issue {ph[2] / steps:.0f} | wait::st {ph[3] / steps:.0f}")
        x = t[48 + 8 * wg: 54 + 8 * wg]
        print(f"CTA 0 warpgroup {wg} glue per step: S-loaded fence + remote arrive {x[1] / steps:.0f} | m hand-over {x[2] / steps:.0f} | "
              f"P-ready fence + remote arrive {x[3] / steps:.0f} | l hand-over {x[4] / steps:.0f} | cursor to next step {x[5] / steps:.0f}")
a = t[16:24]
steps, tot = max(a[0], 1), max(a[1], 1)
print(f"MMA warp: {a[0]} steps, {tot} cycles = {tot / steps:.0f} per step | waits per step: S in registers {a[2] / steps:.0f}, K tile {a[3] / steps:.0f}, "
      f"V tile {a[4] / steps:.0f}, P ready {a[5] / steps:.0f}, O drained {a[6] / steps:.0f}, Q tile {a[7] / steps:.0f} | "
      f"issuing {(tot - sum(a[2:8])) / steps:.0f} per step")
