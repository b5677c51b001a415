"""Standalone launches of the HBM-bound kernels at C1 sizes (for ncu --set full): python scripts/prof_kernels.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import seedvr2_b200 as svr2
ops = svr2.ops
dev = torch.device("cuda:0")
SIZE = sys.argv[1] if len(sys.argv) > 1 else "c1"
GRID = {"c1": (1, 128, 128), "c2": (9, 90, 160)}[SIZE]
L, D, heads, F = GRID[0] * GRID[1] * GRID[2], 2560, 20, 6912
g = torch.Generator(device="cuda").manual_seed(0)
x = torch.randn(L, D, generator=g, device=dev).bfloat16()
r = torch.randn(L, D, generator=g, device=dev).bfloat16()
emb = torch.randn(6 * D, generator=g, device=dev).bfloat16()
sh, sc, gt = (torch.randn(D, generator=g, device=dev).bfloat16() for _ in range(3))
plan = svr2.get_plan(GRID, (4, 3, 3), False, 58)
ga = torch.randn(L, F, generator=g, device=dev).bfloat16()
up = torch.randn(L, F, generator=g, device=dev).bfloat16()
qkv_v = torch.randn(L, 3 * heads * 128, generator=g, device=dev).bfloat16()
qkv_t = torch.randn(58, 3 * heads * 128, generator=g, device=dev).bfloat16()
w = [torch.ones(128, device=dev).bfloat16() for _ in range(4)]
freqs = svr2.block._Rope(ops.ROPE_3B_MM, 128).rope.freqs.to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
timer = ops.KernelTimer()
ops.set_timer(timer)
for it in range(4):
    flush.zero_()
    ops.modnorm(x, emb, sh, sc, 0, 1e-5)
    flush.zero_()
    ops.gate_residual_modnorm(x, r, emb, gt, sh, sc, 1e-5)
    flush.zero_()
    ops.gate_residual(x, r, emb, gt, 1)
    flush.zero_()
    ops.qkv_pack(plan, qkv_v, qkv_t, *w, freqs, ops.ROPE_3B_MM, heads, 1e-5)
    flush.zero_()
    ops.swiglu(ga, up)
torch.cuda.synchronize()
P = plan.packed_rows
bytes_ = {"svr2_modnorm_fwd": 4 * L * D, "svr2_gate_residual_modnorm_fwd": 8 * L * D, "svr2_gate_residual_fwd": 6 * L * D,
          "svr2_qkv_pack_fwd": (L + 58 + P) * 3 * heads * 256, "svr2_swiglu_fwd": 6 * L * F}
for k, recs in timer.records.items():
    ms = [a.elapsed_time(b) for a, b in recs][1:]
    best = min(ms)
    print(f"[{SIZE}] {k}: best {best*1e3:.1f} us  -> {bytes_[k]/best/1e6:.0f} GB/s (L2 flushed, cold)")
