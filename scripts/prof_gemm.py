"""Two launches of each tcgen05 GEMM form at the size of one v1080 forward (73,440 video + 58 text rows, 3B widths), for
`ncu --set full -k regex:linear_tc_kernel`:  python scripts/prof_gemm.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import seedvr2_b200 as svr2
ops = svr2.ops
dev = torch.device("cuda:0")
grid, heads, D, Fh = (9, 68, 120), 20, 2560, 6912
plan = svr2.get_plan(grid, (4, 3, 3), False, 58)
M = plan.n_tokens + 58
g = torch.Generator(device="cuda").manual_seed(0)
r = lambda *s: torch.randn(*s, generator=g, device=dev).bfloat16()
x, wqkv, wo, bo = r(M, D), r(3 * D, D) / D ** 0.5, r(D, D) / D ** 0.5, r(D)
wg, wu, w2, h = r(Fh, D) / D ** 0.5, r(Fh, D) / D ** 0.5, r(D, Fh) / Fh ** 0.5, r(M, Fh)
emb, gb, resid = r(6 * D), r(D), r(M, D)
nw = [(1.0 + 0.1 * torch.randn(128, generator=g, device=dev)).bfloat16() for _ in range(4)]
f = (1.0 / (10000 ** (torch.arange(0, 42, 2)[:21].float() / 42))).to(dev)
fkey = tuple(float(v) for v in f.tolist())
bufs = None
timer = ops.KernelTimer()
ops.set_timer(timer)
for _ in range(2):
    bufs = ops.linear_qkv_pack(plan, x, wqkv, 0, *nw, f, fkey, 1, heads, 1e-5, out=bufs)
    ops.linear_swiglu(x, wg, wu)
    ops.linear_gate_residual(h, w2, None, emb, gb, 1, resid)
    ops.linear_gate_residual(x, wo, bo, emb, gb, 0, resid)
torch.cuda.synchronize()
fl = {"svr2_linear_qkv_pack_fwd": 2 * M * D * 3 * D, "svr2_linear_swiglu_fwd": 4 * M * D * Fh,
      "svr2_linear_gate_residual_fwd": 2 * M * D * (Fh + D)}
for k, (n, ms) in timer.summary().items():
    per = 2 if k == "svr2_linear_gate_residual_fwd" else 1
    print(f"{k}: {ms / n * 1e3:.1f} us per launch, {fl[k] / per * n / ms / 1e9:.0f} TFLOP/s")
