"""svr2_causal_conv3d_fwd against cuDNN (torch.nn.functional.conv3d, channels_last_3d and contiguous) at the VAE decoder's shapes.
Usage: python scripts/conv_bench.py [fp16|bf16]   — prints one line per shape: TFLOP/s of each and the ratio; L2 is flushed
between timed launches (a 256-MiB write)."""
import json
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, ".")
import seedvr2_b200 as svr2      # noqa: E402

dtype = torch.bfloat16 if (len(sys.argv) > 1 and sys.argv[1] == "bf16") else torch.float16
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
peaks = {}
try:
    peaks = json.load(open("MEASURED_PEAKS.json"))
except Exception:      # noqa: BLE001
    pass
peak = float(peaks.get("bf16_tflops_sustained", 1400.8))


def timed(fn, n=9):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


# (c_in, c_out, kt, ks, T, H, W): decoder levels of a 1080p clip (latent 135 x 240), one temporal slice of 4 latent frames
SHAPES = [
    (128, 128, 3, 3, 5, 1080, 1920),
    (256, 128, 3, 3, 5, 1080, 1920),
    (256, 256, 3, 3, 5, 540, 960),
    (512, 256, 3, 3, 5, 540, 960),
    (512, 512, 3, 3, 3, 270, 480),
    (512, 512, 3, 3, 2, 135, 240),
    (256, 128, 1, 1, 5, 1080, 1920),
]
quick = "quick" in sys.argv[1:]      # ours only, three shapes: variant A/B runs
if quick:
    SHAPES = [SHAPES[0], SHAPES[2], SHAPES[6]]
out = []
for ci, co, kt, ks, T, H, W in SHAPES:
    x = torch.randn(1, ci, T, H, W, device=dev).to(dtype).contiguous(memory_format=torch.channels_last_3d)
    w = (torch.randn(co, ci, kt, ks, ks, device=dev) * (ci * kt * ks * ks) ** -0.5).to(dtype)
    b = torch.randn(co, device=dev).to(dtype)
    conv = svr2.vae_conv.CausalConv3dB200(w, b)
    xt = x[0].permute(1, 2, 3, 0)
    y = torch.empty(T, H, W, co, dtype=dtype, device=dev)
    t_ours = timed(lambda: svr2.vae_conv.causal_conv3d(xt, conv.weight_packed, conv.bias, out=y))
    flops = 2.0 * T * H * W * kt * ks * ks * ci * co
    # the reference: extend_head (a concatenation) + cuDNN
    w_cl = w.contiguous(memory_format=torch.channels_last_3d)

    def ref_cl():
        xx = torch.cat((x[:, :, :1].expand(-1, -1, kt - 1, -1, -1), x), 2) if kt > 1 else x
        return F.conv3d(xx, w_cl, b, padding=(0, ks // 2, ks // 2))

    xc = x.contiguous()

    def ref_nc():
        xx = torch.cat((xc[:, :, :1].expand(-1, -1, kt - 1, -1, -1), xc), 2) if kt > 1 else xc
        return F.conv3d(xx, w, b, padding=(0, ks // 2, ks // 2))

    res = {}
    if quick:
        print(json.dumps({"shape": [ci, co, kt, ks, T, H, W], "ms": round(t_ours, 3), "tflops": round(flops / t_ours / 1e9, 1)}), flush=True)
        continue
    for name, fn in (("cudnn_channels_last", ref_cl), ("cudnn_contiguous", ref_nc)):
        try:
            res[name] = timed(fn, 3)
        except RuntimeError as e:      # noqa: PERF203 — out of memory / unsupported
            res[name] = None
            torch.cuda.empty_cache()
    chk = ref_cl()
    err = float((chk[0].permute(1, 2, 3, 0).float() - y.float()).abs().max())
    del chk
    line = {"shape": [ci, co, kt, ks, T, H, W], "dtype": str(dtype).split(".")[-1], "ms": round(t_ours, 3),
            "tflops": round(flops / t_ours / 1e9, 1), "frac_of_sustained_peak": round(flops / t_ours / 1e9 / peak, 3),
            "cudnn_ms": {k: (None if v is None else round(v, 3)) for k, v in res.items()},
            "cudnn_tflops": {k: (None if v is None else round(flops / v / 1e9, 1)) for k, v in res.items()},
            "max_abs_diff_vs_cudnn": err}
    out.append(line)
    print(json.dumps(line), flush=True)
    del x, xc, w, y, conv
    torch.cuda.empty_cache()
