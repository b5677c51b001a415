"""Two launches of svr2_groupnorm_silu_fwd at the decoder's top level, for ncu (scripts/gpu.sh ncu:gn:scripts/prof_gn.py)."""
import sys

import torch

sys.path.insert(0, ".")
import seedvr2_b200 as svr2      # noqa: E402

dev = torch.device("cuda:0")
x = torch.randn(5, 1080, 1920, 128, device=dev).half()
w = torch.ones(128, device=dev).half()
b = torch.zeros(128, device=dev).half()
for _ in range(2):
    y = svr2.vae_block.groupnorm_silu(x, w, b, 32, 1e-6)
torch.cuda.synchronize()
print(float(y.float().abs().mean()))
