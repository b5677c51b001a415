"""The upper half of the VAE decoder at 1080p (up block 2 at 540 x 960 with its spatial up-sampler, up block 3 at 1080 x 1920, output
norm + convolution: 6 residual blocks, 8 further convolutions, ~33 of the decoder's ~48 TFLOP per frame) built from the reference's real
parts (tests/vae_standins.py), fp16: the reference's modules through cuDNN against the same modules after install_vae_blocks().
Usage: python scripts/vae_decoder_top_bench.py [frames]"""
import copy
import json
import sys

import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import seedvr2_b200 as svr2      # noqa: E402
from vae_standins import DecoderTop, reference_parts      # noqa: E402

_, _, _, MemoryState = reference_parts()
dev = torch.device("cuda:0")
T = int(sys.argv[1]) if len(sys.argv) > 1 else 4
torch.manual_seed(0)
ref = DecoderTop(512, 256, 128, slicing=False).eval().to(dev).half()
mine = copy.deepcopy(ref)
n_blocks, n_convs = svr2.vae_block.install_vae_blocks(mine)
mine.use_rebound_norm = True
x = (torch.randn(1, 512, T, 540, 960, device=dev) * 0.5).half()


def timed(fn, n=3):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


with torch.no_grad():
    t_mine = timed(lambda: mine(x, memory_state=MemoryState.DISABLED))
    kt = svr2.ops.KernelTimer()
    svr2.ops.set_timer(kt)
    a = mine(x, memory_state=MemoryState.DISABLED)
    torch.cuda.synchronize()
    svr2.ops.set_timer(None)
    ours_ms = {k: round(v[1], 2) for k, v in kt.summary().items()}
    a = a.float()
    torch.cuda.empty_cache()
    try:
        t_ref = timed(lambda: ref(x, memory_state=MemoryState.DISABLED), 2)
        b = ref(x, memory_state=MemoryState.DISABLED).float()
        diff, scale = float((a - b).abs().max()), float(b.abs().max())
    except RuntimeError as e:
        t_ref, diff, scale = None, None, None
        print("reference failed:", str(e)[:200])
hw = 540 * 960
conv = lambda ci, co, pos: 2.0 * pos * 27 * ci * co      # noqa: E731
flops = T * (conv(512, 256, hw) + 5 * conv(256, 256, hw) + 2.0 * hw * 512 * 256          # up block 2 (+ shortcut)
             + 2.0 * hw * 256 * 1024 + conv(256, 256, 4 * hw)                             # expansion + up-sampler convolution
             + conv(256, 128, 4 * hw) + 5 * conv(128, 128, 4 * hw) + 2.0 * 4 * hw * 256 * 128   # up block 3 (+ shortcut)
             + conv(128, 3, 4 * hw))
print(json.dumps({"decoder_top_half": {"frames": T, "installed": [n_blocks, n_convs], "ms": round(t_mine, 2), "ms_per_frame": round(t_mine / T, 2),
                                        "tflops_algorithmic": round(flops / t_mine / 1e9, 1),
                                        "reference_ms": None if t_ref is None else round(t_ref, 2),
                                        "speedup": None if t_ref is None else round(t_ref / t_mine, 2),
                                        "kernel_ms_ours": ours_ms, "torch_side_ms": round(max(t_mine - sum(ours_ms.values()), 0.0), 2),
                                        "max_abs_diff": diff, "out_scale": scale}}))
