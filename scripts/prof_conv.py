"""One launch of svr2_causal_conv3d_fwd per decoder shape class, for ncu (scripts/gpu.sh ncu:conv3d_tc:scripts/prof_conv.py)."""
import sys

import torch

sys.path.insert(0, ".")
import seedvr2_b200 as svr2      # noqa: E402

dev = torch.device("cuda:0")
for ci, co, T, H, W in ((128, 128, 3, 1080, 1920), (256, 256, 3, 540, 960)):
    x = torch.randn(T, H, W, ci, device=dev).half()
    w = (torch.randn(co, ci, 3, 3, 3, device=dev) * (ci * 27) ** -0.5).half()
    b = torch.randn(co, device=dev).half()
    wp = svr2.vae_conv.pack_conv_weight(w)
    for _ in range(2):
        y = svr2.vae_conv.causal_conv3d(x, wp, b)
    torch.cuda.synchronize()
    print(ci, co, float(y.float().abs().mean()))
