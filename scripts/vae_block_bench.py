"""The VAE decoder's residual blocks at 1080p: comfyui-seedvr2_b200/vae_block.py against the reference's block composed from its real
parts (InflatedCausalConv3d through cuDNN, causal_norm_wrapper, nn.SiLU; staged copy), fp16.  One line per decoder level."""
import json
import sys

import torch
from torch import nn

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import seedvr2_b200 as svr2      # noqa: E402
from oracle import refstage      # noqa: E402

refstage.activate()
from src.models.video_vae_v3.modules.causal_inflation_lib import causal_norm_wrapper, init_causal_conv3d      # noqa: E402
from src.models.video_vae_v3.modules.types import MemoryState      # noqa: E402
from vae_standins import ResnetBlock3D      # noqa: E402

dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fn, n=5):
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


# (c_in, c_out, T, H, W): one temporal slice (4 latent frames -> up to 13 + frames at the top; 5 keeps the reference inside memory)
for ci, co, T, H, W in ((128, 128, 5, 1080, 1920), (256, 128, 5, 1080, 1920), (256, 256, 5, 540, 960), (512, 512, 3, 270, 480)):
    torch.manual_seed(0)
    ref = ResnetBlock3D(ci, co, True).eval().to(dev).half()
    mine = svr2.vae_block.ResnetBlock3DB200.from_reference(ref)
    x = torch.randn(1, ci, T, H, W, device=dev).half().contiguous(memory_format=torch.channels_last_3d)
    xc = x.contiguous()
    with torch.no_grad():
        t_mine = timed(lambda: mine(x, None, memory_state=MemoryState.DISABLED))
        res = {}
        for name, inp in (("reference_contiguous", xc), ("reference_channels_last", x)):
            try:
                res[name] = timed(lambda: ref(inp, memory_state=MemoryState.DISABLED), 3)
            except RuntimeError:
                res[name] = None
                torch.cuda.empty_cache()
        # per-kernel split of one call (CUDA events around every C-ABI launch)
        kt = svr2.ops.KernelTimer()
        svr2.ops.set_timer(kt)
        mine(x, None, memory_state=MemoryState.DISABLED)
        torch.cuda.synchronize()
        svr2.ops.set_timer(None)
        split = {k: [round(v, 3) for v in vs] for k, vs in kt.durations().items()}
        gn_bytes = 3.0 * T * H * W * 2       # per channel: statistics read + apply read + write
        a = mine(x, None, memory_state=MemoryState.DISABLED).float()
        b = ref(xc, memory_state=MemoryState.DISABLED).float()
        diff = float((a - b).abs().max())
        del a, b
    flops = 2.0 * T * H * W * 27 * (ci * co + co * co) + (2.0 * T * H * W * ci * co if ci != co else 0)
    print(json.dumps({"block": [ci, co, T, H, W], "ms": round(t_mine, 3), "tflops": round(flops / t_mine / 1e9, 1),
                      "reference_ms": {k: (None if v is None else round(v, 3)) for k, v in res.items()},
                      "speedup": {k: (None if v is None else round(v / t_mine, 2)) for k, v in res.items()},
                      "max_abs_diff": diff, "kernels_ms": split,
                      "groupnorm_GBps": [round(gn_bytes * c / (ms * 1e6), 0) for c, ms in zip((ci, co), split.get("svr2_groupnorm_silu_fwd", []))]}),
          flush=True)
    del ref, mine, x, xc
    torch.cuda.empty_cache()
