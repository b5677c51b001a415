"""Where do the vendor libraries land at NaSwin's sequence lengths?  Same FLOPs as one window-attention launch of the v1080 / C2 / C1
grids, as a dense batch [windows, heads, seq, 128] (equal-length windows only: the libraries get the EASIER problem — no ragged
sequences, no text interleave, no window reverse), through torch's cuDNN and flash SDPA backends and flash_attn_func, timed with
CUDA events after warm-up, against svr2_window_attn_fwd on the real packed layout.  python scripts/attn_vs_library.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from torch.nn.attention import SDPBackend, sdpa_kernel
import seedvr2_b200 as svr2

dev = torch.device("cuda:0")


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


for name, grid, heads in (("v1080", (9, 68, 120), 20), ("c2", (9, 90, 160), 20), ("c1", (1, 128, 128), 20), ("c3", (31, 68, 120), 24)):
    plan = svr2.get_plan(grid, (4, 3, 3), False, 58)
    g = torch.Generator(device="cuda").manual_seed(0)
    q, k, v = (torch.randn(plan.packed_rows, heads, 128, generator=g, device=dev).bfloat16() for _ in range(3))
    ours_ms = timeit(lambda: svr2.ops.window_attn(plan, q, k, v, heads))
    flops = plan.attention_flops(heads)
    line = [f"{name:6s} ours (packed, ragged, reverse fused) {ours_ms * 1e3:8.1f} us {flops / ours_ms / 1e9:6.0f} TF/s"]
    # dense stand-in: every window at the MAXIMUM length (>= the real work), so the library's TF/s are computed on ITS flops
    seq = int(plan.window_lens.max()) + 58
    nw = plan.n_windows
    qd, kd, vd = (torch.randn(nw, heads, seq, 128, generator=g, device=dev).bfloat16() for _ in range(3))
    dflops = nw * heads * 4.0 * 128 * seq * seq
    for label, backend in (("cudnn sdpa", SDPBackend.CUDNN_ATTENTION), ("flash sdpa", SDPBackend.FLASH_ATTENTION)):
        try:
            with sdpa_kernel(backend):
                ms = timeit(lambda: F.scaled_dot_product_attention(qd, kd, vd))
            line.append(f"{label} {ms * 1e3:8.1f} us {dflops / ms / 1e9:6.0f} TF/s")
        except Exception as e:
            line.append(f"{label} unavailable ({type(e).__name__})")
    try:
        from flash_attn import flash_attn_func
        qf, kf, vf = (t.transpose(1, 2).contiguous() for t in (qd, kd, vd))
        ms = timeit(lambda: flash_attn_func(qf, kf, vf))
        line.append(f"flash_attn {ms * 1e3:8.1f} us {dflops / ms / 1e9:6.0f} TF/s")
    except Exception as e:
        line.append(f"flash_attn unavailable ({type(e).__name__})")
    print(f"{line[0]} | dense [{nw},{heads},{seq},128]: " + " | ".join(line[1:]), flush=True)
    del qd, kd, vd
