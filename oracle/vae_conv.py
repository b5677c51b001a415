"""TEST INFRASTRUCTURE ONLY — restatement of the video VAE's causal 3-D convolution (first slice).

Follows InflatedCausalConv3d (src/models/video_vae_v3/modules/causal_inflation_lib.py:61-272), stride 1:
  * the temporal padding is dropped from the module's own padding (:74) and replaced by a causal head in front of the input —
    2 * temporal_padding copies of the first frame (extend_head :422-437 / cache_send_recv context_parallel_lib.py:55-67) on the
    first temporal slice, the module's ``memory`` (the previous slice's last kernel_t - stride_t frames, :226-240 / :273-283) after;
  * spatially the convolution zero-pads (Conv3d padding, :255 ``padding`` tuple handed to memory_limit_conv);
  * memory_limit_conv (:117-211) only splits the same convolution along H / W with exact halos: the function computed is unchanged.
Pinned: tests/test_oracle_golden.py runs this against the real module (staged copy) over several memory states, bit-equal on CPU.
"""
from typing import Optional, Tuple

import torch
import torch.nn.functional as F


def causal_conv3d(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], memory: Optional[torch.Tensor] = None
                  ) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """x [B, C, T, H, W], weight [Co, Ci, kt, kh, kw] (odd kernel, stride 1, "same" spatial padding).  Returns (y, new memory)."""
    kt, kh, kw = weight.shape[2:]
    if kt > 1:
        head = memory.to(x) if memory is not None else x[:, :, :1].repeat(1, 1, kt - 1, 1, 1)
        xx = torch.cat((head, x), dim=2)
    else:
        xx = x
    y = F.conv3d(xx, weight, bias, stride=1, padding=(0, kh // 2, kw // 2))
    new_memory = xx[:, :, -(kt - 1):].detach() if kt > 1 else None
    return y, new_memory


def causal_conv3d_down(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], memory: Optional[torch.Tensor] = None,
                       pad_hw: int = 1) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """Downsample3D.forward (attn_video_vae.py:228-246) as the encoder builds it: zero row / column after the last one
    (`pad (0, 1, 0, 1)`, :239-241), then InflatedCausalConv3d with kernel (kt, 3, 3), stride (2 if kt == 3 else 1, 2, 2), no spatial
    padding; the causal head is two copies of the first frame (times = temporal_padding * 2, causal_inflation_lib.py:265) or the
    memory, which is the last kernel_t - stride_t = 1 input frame (:262, :273-283).  Returns (y, new memory)."""
    kt = weight.shape[2]
    st = 2 if kt == 3 else 1
    if kt > 1:
        head = memory.to(x) if memory is not None else x[:, :, :1].repeat(1, 1, 2, 1, 1)
        xx = torch.cat((head, x), dim=2)
    else:
        xx = x
    y = F.conv3d(F.pad(xx, (0, pad_hw, 0, pad_hw)), weight, bias, stride=(st, 2, 2))
    return y, (xx[:, :, -1:].detach() if kt > 1 else None)
