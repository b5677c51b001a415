"""The VAE mid block's attention layer at a 1080p latent frame (135 x 240 = 32,400 positions, one 512-channel head): AttentionB200
against the stand-in of diffusers' layer (torch SDPA, memory-efficient backend), fp16."""
import json
import sys

import torch
from torch import nn

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import seedvr2_b200 as svr2      # noqa: E402
from vae_standins import Attention      # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)
ref = Attention(512).eval().to(dev).half()
mine = svr2.vae_attn.AttentionB200.from_reference(ref)
x = torch.randn(2, 135, 240, 512, device=dev).half().permute(0, 3, 1, 2)      # two frames, NHWC-strided


def timed(fn, n=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


with torch.no_grad():
    t_mine = timed(lambda: mine(x))
    t_ref = timed(lambda: ref(x.contiguous()))
    kt = svr2.ops.KernelTimer()
    svr2.ops.set_timer(kt)
    a = mine(x)
    torch.cuda.synchronize()
    svr2.ops.set_timer(None)
    b = ref(x.contiguous())
print(json.dumps({"vae_mid_attention_1080p_latent": {"frames": 2, "ms_per_frame": round(t_mine / 2, 3), "reference_ms_per_frame": round(t_ref / 2, 3),
                                                      "speedup": round(t_ref / t_mine, 2),
                                                      "kernel_ms": {k: round(v[1], 3) for k, v in kt.summary().items()},
                                                      "max_abs_diff": float((a.float() - b.float()).abs().max()),
                                                      "out_rms": float(b.float().pow(2).mean().sqrt())}}))
