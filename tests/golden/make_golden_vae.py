"""Golden vectors of the video VAE's causal convolutions: tests/golden/vae_conv.pt, produced by the REAL
InflatedCausalConv3d (disk02/ComfyUI-SeedVR2, /root/reference) in fp32 on the CPU of the build container.
Run once:  python tests/golden/make_golden_vae.py        (never runs on the GPU box)

Per case: weight / bias / input (fp16-representable values, stored as fp16), the outputs of the module over streamed temporal slices
(MemoryState INITIALIZING -> ACTIVE ...) and the memory it held after every slice.  The oracle (oracle/vae_conv.py) and the CUDA
kernels are both checked against these, so the pin does not depend on the reference being importable where the tests run."""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[0:0] = [os.path.join(ROOT, "oracle", "shim"), "/root/reference", ROOT]

import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

from src.models.video_vae_v3.modules.causal_inflation_lib import init_causal_conv3d  # noqa: E402
from src.models.video_vae_v3.modules.types import MemoryState  # noqa: E402

CASES = {
    # name: (c_in, c_out, kernel, stride, padding of the module, pad (0,1,0,1) in front (Downsample3D), slices)
    "conv333": (64, 128, (3, 3, 3), (1, 1, 1), (1, 1, 1), False, ((0, 3), (3, 4), (4, 7))),
    "conv133": (64, 128, (1, 3, 3), (1, 1, 1), (0, 1, 1), False, ((0, 4), (4, 7))),
    "down_spatial": (64, 128, (1, 3, 3), (1, 2, 2), (0, 0, 0), True, ((0, 4), (4, 7))),
    "down_temporal": (64, 128, (3, 3, 3), (2, 2, 2), (1, 0, 0), True, ((0, 5), (5, 9), (9, 11))),
}


def main():
    out = {}
    for name, (ci, co, k, s, p, pad, slices) in CASES.items():
        g = torch.Generator().manual_seed(sum(map(ord, name)))
        conv = init_causal_conv3d(ci, co, kernel_size=k, stride=s, padding=p, inflation_mode="tail").eval()
        w = (torch.randn(co, ci, *k, generator=g) * (ci * k[0] * k[1] * k[2]) ** -0.5).half()
        b = (torch.randn(co, generator=g) * 0.3).half()
        T = slices[-1][1]
        x = torch.randn(1, ci, T, 9, 12, generator=g).half()
        with torch.no_grad():
            conv.weight.copy_(w.float())
            conv.bias.copy_(b.float())
            ys, mems = [], []
            for i, (a, e) in enumerate(slices):
                xi = x[:, :, a:e].float()
                if pad:
                    xi = F.pad(xi, (0, 1, 0, 1))
                ys.append(conv(xi, memory_state=MemoryState.INITIALIZING if i == 0 else MemoryState.ACTIVE).clone())
                mems.append(None if conv.memory is None else conv.memory.clone())
        out[name] = {"weight": w, "bias": b, "x": x, "kernel": k, "stride": s, "pad_hw": int(pad), "slices": slices, "y": ys, "memory": mems}
        print(name, [tuple(y.shape) for y in ys])
    torch.save(out, os.path.join(HERE, "vae_conv.pt"))


if __name__ == "__main__":
    main()
