"""Time the tcgen05 linear kernels against the library GEMM (cuBLAS through torch) on the block's real shapes.
usage: python scripts/gemm_bench.py [c1|c2|c3r]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

import seedvr2_b200 as svr2

ops = svr2.ops
dev = torch.device("cuda:0")
which = sys.argv[1] if len(sys.argv) > 1 else "c1"
M = {"c1": 16384 + 58, "c2": 129600 + 58, "c3r": 48960, "v1080": 9 * 68 * 120 + 58}[which]
D, Fh = (3072, 12288) if which == "c3r" else (2560, 6912)
g = torch.Generator(device="cuda").manual_seed(0)
r = lambda *s: torch.randn(*s, generator=g, device=dev).bfloat16()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def report(name, flops, ours, lib):
    print(f"{name:44s} ours {ours * 1e3:8.1f} us {flops / ours / 1e9:7.1f} TF/s | library path {lib * 1e3:8.1f} us {flops / lib / 1e9:7.1f} TF/s | x{lib / ours:.3f}")


x = r(M, D)
wqkv, wo, bo = r(3 * D, D) / D ** 0.5, r(D, D) / D ** 0.5, r(D)
emb, gb, resid = r(6 * D), r(D), r(M, D)
report(f"proj_qkv [{M},{D}]x[{3 * D},{D}]", 2 * M * D * 3 * D, timeit(lambda: ops.linear(x, wqkv)), timeit(lambda: F.linear(x, wqkv)))
if which in ("v1080", "c1", "c3r"):
    # the fused projection (GEMM + q/k RMSNorm + RoPE + window scatter) against GEMM + svr2_qkv_pack_fwd
    import math
    grid = {"v1080": (9, 68, 120), "c1": (1, 128, 128), "c3r": (6, 68, 120)}[which]
    heads, kind = D // 128, (2 if which == "c3r" else 1)
    plan = svr2.get_plan(grid, (4, 3, 3), False, 58)
    Mx = plan.n_tokens + 58
    xq = r(Mx, D)
    nw = [(1.0 + 0.1 * torch.randn(128, generator=g, device=dev)).bfloat16() for _ in range(4)]
    f = (1.0 / (10000 ** (torch.arange(0, 42, 2)[:21].float() / 42)) if kind == 1 else torch.linspace(1.0, 128.0, 10) * math.pi).to(dev)
    fkey = tuple(float(v) for v in f.tolist())
    bufs = ops.linear_qkv_pack(plan, xq, wqkv, 0, *nw, f, fkey, kind, heads, 1e-5)

    def fused():
        ops.linear_qkv_pack(plan, xq, wqkv, 0, *nw, f, fkey, kind, heads, 1e-5, out=bufs)
        ops.replicate_text(plan, bufs[3], bufs[0], bufs[1], bufs[2], heads)

    def unfused():
        y = ops.linear(xq, wqkv)
        ops.qkv_pack(plan, y[:plan.n_tokens], y[plan.n_tokens:], *nw, f, kind, heads, 1e-5)

    report("proj_qkv + norm + rope + partition (fused)", 2 * Mx * D * 3 * D, timeit(fused), timeit(unfused))
report("proj_out + gate + residual", 2 * M * D * D, timeit(lambda: ops.linear_gate_residual(x, wo, bo, emb, gb, 0, resid)),
       timeit(lambda: ops.gate_residual(F.linear(x, wo, bo), resid, emb, gb, 0)))
if which == "c3r":
    w1, b1, w2, b2 = r(Fh, D) / D ** 0.5, r(Fh), r(D, Fh) / Fh ** 0.5, r(D)
    report("mlp proj_in + gelu", 2 * M * D * Fh, timeit(lambda: ops.linear(x, w1, b1, ops.ACT_GELU_TANH)),
           timeit(lambda: torch._addmm_activation(b1, x, w1.t(), use_gelu=True)))
    h = r(M, Fh)
    report("mlp proj_out + gate + residual", 2 * M * D * Fh, timeit(lambda: ops.linear_gate_residual(h, w2, b2, emb, gb, 1, resid)),
           timeit(lambda: ops.gate_residual(F.linear(h, w2, b2), resid, emb, gb, 1)))
else:
    wg, wu, w2 = r(Fh, D) / D ** 0.5, r(Fh, D) / D ** 0.5, r(D, Fh) / Fh ** 0.5
    report("mlp swiglu in-projections", 2 * 2 * M * D * Fh, timeit(lambda: ops.linear_swiglu(x, wg, wu)),
           timeit(lambda: ops.swiglu(F.linear(x, wg), F.linear(x, wu))))
    h = r(M, Fh)
    report("mlp proj_out + gate + residual", 2 * M * D * Fh, timeit(lambda: ops.linear_gate_residual(h, w2, None, emb, gb, 1, resid)),
           timeit(lambda: ops.gate_residual(F.linear(h, w2), resid, emb, gb, 1)))
