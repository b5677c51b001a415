"""TEST INFRASTRUCTURE ONLY — restatement of the video VAE's residual block.

Follows ResnetBlock3D.forward (src/models/video_vae_v3/modules/attn_video_vae.py:311-361) as the VAE configures it
(temb_channels=None, no up / down sampler inside the block, dropout 0, non_linearity "silu", time_embedding_norm "default"):

    h = silu(group_norm_per_frame(x, norm1)); h = conv1(h)           :316-336   (causal_norm_wrapper, causal_inflation_lib.py:354-409)
    h = silu(group_norm_per_frame(h, norm2)); h = conv2(h)           :346-356
    x = conv_shortcut(x) if in_channels != out_channels              :358-359
    out = (x + h) / output_scale_factor                              :361

diffusers (the base class ResnetBlock2D) is not in this image, so the real block cannot be instantiated; its parts that do import
— InflatedCausalConv3d and causal_norm_wrapper — are the real ones in tests/test_gpu_vae_block.py, composed in this order.
Parity unpinned for the block as a whole (no fixture of the real ResnetBlock3D); pinned for its convolution (oracle/vae_conv.py)."""
from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F

from .vae_conv import causal_conv3d


def group_norm_per_frame(x: torch.Tensor, groups: int, weight, bias, eps: float) -> torch.Tensor:
    """causal_norm_wrapper for nn.GroupNorm: "b c t h w -> (b t) c h w", normalise, and back (causal_inflation_lib.py:366-409)."""
    b, c, t, h, w = x.shape
    y = F.group_norm(x.permute(0, 2, 1, 3, 4).reshape(b * t, c, h, w), groups, weight, bias, eps)
    return y.reshape(b, t, c, h, w).permute(0, 2, 1, 3, 4)


def resnet_block3d(x: torch.Tensor, p: Dict[str, torch.Tensor], groups: int = 32, eps: float = 1e-6, output_scale_factor: float = 1.0,
                   memory: Optional[Tuple] = None) -> Tuple[torch.Tensor, Tuple]:
    """x [B, C, T, H, W]; p: norm1.weight/.bias, conv1.weight/.bias, norm2.*, conv2.*, optional conv_shortcut.*.
    memory = (conv1 memory, conv2 memory) of the previous temporal slice.  Returns (out, new memory)."""
    m1, m2 = memory if memory is not None else (None, None)
    h = F.silu(group_norm_per_frame(x, groups, p["norm1.weight"], p["norm1.bias"], eps))
    h, n1 = causal_conv3d(h, p["conv1.weight"], p["conv1.bias"], m1)
    h = F.silu(group_norm_per_frame(h, groups, p["norm2.weight"], p["norm2.bias"], eps))
    h, n2 = causal_conv3d(h, p["conv2.weight"], p["conv2.bias"], m2)
    if "conv_shortcut.weight" in p:
        x, _ = causal_conv3d(x, p["conv_shortcut.weight"], p.get("conv_shortcut.bias"))
    return (x + h) / output_scale_factor, (n1, n2)
