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


def _rgb_to_lab(rgb: torch.Tensor) -> torch.Tensor:
    """color_fix.py:368-411 (_rgb_to_lab_batch), D65."""
    m = torch.tensor([[0.4124564, 0.3575761, 0.1804375], [0.2126729, 0.7151522, 0.0721750], [0.0193339, 0.1191920, 0.9503041]],
                     dtype=torch.float32, device=rgb.device)
    eps, kappa = 6.0 / 29.0, (29.0 / 3.0) ** 3
    lin = torch.where(rgb > 0.04045, torch.pow((rgb + 0.055) / 1.055, 2.4), rgb / 12.92)
    B, C, H, W = lin.shape
    xyz = torch.matmul(lin.permute(0, 2, 3, 1).reshape(-1, 3), m.T).reshape(B, H, W, 3).permute(0, 3, 1, 2)
    xyz[:, 0].div_(0.95047)
    xyz[:, 2].div_(1.08883)
    f = torch.where(xyz > eps ** 3, torch.pow(xyz, 1.0 / 3.0), xyz.mul(kappa).add_(16.0).div_(116.0))
    return torch.stack([f[:, 1].mul(116.0).sub_(16.0), (f[:, 0] - f[:, 1]).mul_(500.0), (f[:, 1] - f[:, 2]).mul_(200.0)], dim=1)


def _lab_to_rgb(lab: torch.Tensor) -> torch.Tensor:
    """color_fix.py:414-470 (_lab_to_rgb_batch)."""
    mi = torch.tensor([[3.2404542, -1.5371385, -0.4985314], [-0.9692660, 1.8760108, 0.0415560], [0.0556434, -0.2040259, 1.0572252]],
                      dtype=torch.float32, device=lab.device)
    eps, kappa = 6.0 / 29.0, (29.0 / 3.0) ** 3
    L, a, b = lab[:, 0], lab[:, 1], lab[:, 2]
    fy = (L + 16.0) / 116.0
    fx = a.div(500.0).add_(fy)
    fz = fy - b / 200.0
    inv = lambda f: torch.where(f > eps, torch.pow(f, 3.0), f.mul(116.0).sub_(16.0).div_(kappa))
    x, y, z = inv(fx).mul_(0.95047), inv(fy), inv(fz).mul_(1.08883)
    xyz = torch.stack([x, y, z], dim=1)
    B, C, H, W = xyz.shape
    lin = torch.matmul(xyz.permute(0, 2, 3, 1).reshape(-1, 3), mi.T).reshape(B, H, W, 3).permute(0, 3, 1, 2)
    rgb = torch.where(lin > 0.0031308, torch.pow(torch.clamp(lin, min=0.0), 1.0 / 2.4).mul_(1.055).sub_(0.055), lin * 12.92)
    return torch.clamp(rgb, 0.0, 1.0)


def _hist_match(source: torch.Tensor, reference: torch.Tensor) -> torch.Tensor:
    """color_fix.py:473-513 for equal element counts."""
    flat = source.flatten()
    _, idx = torch.sort(flat)
    ref_sorted, _ = torch.sort(reference.flatten())
    out = torch.empty_like(flat)
    out.scatter_(0, idx, ref_sorted)
    return out.reshape(source.shape)


def lab_color_fix(sample: torch.Tensor, style: torch.Tensor, luminance_weight: float = 0.8) -> torch.Tensor:
    """lab_color_transfer (color_fix.py:249-367) + the batch normalisation: sample / style [T, 3, H, W] in [-1, 1] -> [T, H, W, 3]
    in [0, 1] (same dtype)."""
    dt = sample.dtype
    high, _ = wavelet_decomposition(sample)
    _, low = wavelet_decomposition(style)
    content = high.add_(low).clamp_(-1.0, 1.0).float()
    st = style.float().clone()
    content = content.add_(1.0).mul_(0.5).clamp_(0.0, 1.0)
    st = st.add_(1.0).mul_(0.5).clamp_(0.0, 1.0)
    cl, sl = _rgb_to_lab(content), _rgb_to_lab(st)
    a, b = _hist_match(cl[:, 1], sl[:, 1]), _hist_match(cl[:, 2], sl[:, 2])
    if luminance_weight < 1.0:
        mL = _hist_match(cl[:, 0], sl[:, 0])
        rL = cl[:, 0].mul(luminance_weight).add_(mL.mul(1.0 - luminance_weight))
    else:
        rL = cl[:, 0]
    out = _lab_to_rgb(torch.stack([rL, a, b], dim=1)).mul_(2.0).sub_(1.0).to(dt)
    out = out.permute(0, 2, 3, 1).contiguous()
    return out.clamp_(-1, 1).mul_(0.5).add_(0.5)


def to_uint8(frames: torch.Tensor, bgr: bool = False) -> np.ndarray:
    a = (frames.float().cpu().numpy() * 255.0).astype(np.uint8)
    if bgr and a.shape[-1] >= 3:
        a = np.concatenate([a[..., 2::-1], a[..., 3:]], axis=-1)
    return a
