"""The VAE encoder at 1080p, built from the reference's real parts in the order of Encoder3D.forward (tests/vae_standins.py; the mid block's
attention layer is left out of both arms): the reference's modules through cuDNN against the same modules after install_vae_blocks(), fp16,
streamed (first slice of 4 n + 1 frames INITIALIZING, then slices of 4 n with ACTIVE).  Usage: python scripts/vae_encoder_bench.py [n] [slices] [--no-ref]"""
import copy
import json
import sys

import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import seedvr2_b200 as svr2      # noqa: E402
from vae_standins import Encoder3D, reference_parts      # noqa: E402

_, _, _, MemoryState = reference_parts()
dev = torch.device("cuda:0")
args = [a for a in sys.argv[1:] if not a.startswith("--")]
n = int(args[0]) if len(args) > 0 else 1
n_slices = int(args[1]) if len(args) > 1 else 2
torch.manual_seed(0)
ATTN = "--attention" in sys.argv
ref = Encoder3D(attention=ATTN).eval().to(dev).half()
mine = copy.deepcopy(ref)
installed = svr2.vae_block.install_vae_blocks(mine)
mine.use_rebound_norm = True
xs = [torch.randn(1, 3, 4 * n + (1 if i == 0 else 0), 1080, 1920, device=dev).half() for i in range(n_slices)]


def run(model):
    return [model(x, memory_state=MemoryState.INITIALIZING if i == 0 else MemoryState.ACTIVE) for i, x in enumerate(xs)]


def timed(model, reps):
    run(model)
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        outs = run(model)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2], outs


with torch.no_grad():
    frames = sum(x.shape[2] for x in xs)
    t_mine, outs_m = timed(mine, 3)
    kt = svr2.ops.KernelTimer()
    svr2.ops.set_timer(kt)
    run(mine)
    torch.cuda.synchronize()
    svr2.ops.set_timer(None)
    res = {"encoder_1080p": {"mid_attention": ATTN, "input_frames": frames, "slices": [x.shape[2] for x in xs], "latent_frames": sum(o.shape[2] for o in outs_m),
                             "installed": list(installed), "ms": round(t_mine, 2), "ms_per_frame": round(t_mine / frames, 2),
                             "frames_per_s": round(frames / t_mine * 1e3, 2), "kernel_ms_ours": {k: round(v[1], 2) for k, v in kt.summary().items()}}}
    if "--no-ref" not in sys.argv:
        a = torch.cat([o.float() for o in outs_m], 2)
        del outs_m
        torch.cuda.empty_cache()
        try:
            t_ref, outs_r = timed(ref, 1)
            b = torch.cat([o.float() for o in outs_r], 2)
            res["encoder_1080p"].update({"reference_ms": round(t_ref, 2), "reference_ms_per_frame": round(t_ref / frames, 2),
                                         "speedup": round(t_ref / t_mine, 2), "max_abs_diff": float((a - b).abs().max()),
                                         "rms_diff": float((a - b).pow(2).mean().sqrt()), "out_rms": float(b.pow(2).mean().sqrt())})
        except RuntimeError as e:
            res["encoder_1080p"]["reference_error"] = str(e)[:200]
print(json.dumps(res))
