"""Restatement of the pipeline's default colour correction and output conversion.  TEST INFRASTRUCTURE ONLY.

Reference: src/utils/color_fix.py:122-246 (wavelet_blur, wavelet_decomposition, wavelet_reconstruction),
src/core/generation_phases.py:1296-1337 (applied per batch, then ``clamp_(-1, 1).mul_(0.5).add_(0.5)`` on the [T, H, W, C] video),
inference_cli.py:691-695,734-738 (``(frames * 255.0).astype(uint8)``, RGB -> BGR for OpenCV).  Plain torch ops in the reference's
op order, device-agnostic; pinned against the real module by tests/test_oracle_golden.py (build container) and
tests/test_gpu_postprocess.py (GPU box, staged copy)."""
import numpy as np
import torch
import torch.nn.functional as F


def wavelet_blur(image: torch.Tensor, radius: int) -> torch.Tensor:
    radius = min(radius, max(1, min(image.shape[-2:]) // 8))
    c = image.shape[1]
    k = torch.tensor([[0.0625, 0.125, 0.0625], [0.125, 0.25, 0.125], [0.0625, 0.125, 0.0625]], dtype=image.dtype, device=image.device)
    k = k[None, None].repeat(c, 1, 1, 1)
    image = F.pad(image, (radius, radius, radius, radius), mode="replicate")
    return F.conv2d(image, k, groups=c, dilation=radius)


def wavelet_decomposition(image: torch.Tensor, levels: int = 5):
    high = torch.zeros_like(image)
    low = image
    for i in range(levels):
        low = wavelet_blur(image, 2 ** i)
        high.add_(image).sub_(low)
        image = low
    return high, low


def wavelet_color_fix(sample: torch.Tensor, style: torch.Tensor) -> torch.Tensor:
    """sample / style [T, C, H, W] in [-1, 1] -> [T, H, W, C] in [0, 1] (same dtype)."""
    high, _ = wavelet_decomposition(sample)
    _, low = wavelet_decomposition(style)
    out = high.add_(low).clamp_(-1.0, 1.0)
    out = out.permute(0, 2, 3, 1).contiguous()
    return out.clamp_(-1, 1).mul_(0.5).add_(0.5)


def to_uint8(frames: torch.Tensor, bgr: bool = False) -> np.ndarray:
    a = (frames.float().cpu().numpy() * 255.0).astype(np.uint8)
    if bgr and a.shape[-1] >= 3:
        a = np.concatenate([a[..., 2::-1], a[..., 3:]], axis=-1)
    return a
