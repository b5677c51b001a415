"""The attention layer of the VAE's mid block on the B200.

UNetMidBlock3D (src/models/video_vae_v3/modules/attn_video_vae.py:563-668) puts one diffusers ``Attention`` between its two residual
blocks (``heads = in_channels // attention_head_dim = 1``, GroupNorm(32), biased q / k / v / out projections, residual connection) and
calls it frame by frame: ``rearrange("b c f h w -> (b f) c h w")``, ``attn(hidden_states, temb=temb)``, and back (:660-665).  The layer
itself is diffusers code (``AttnProcessor2_0``: GroupNorm -> to_q / to_k / to_v -> torch's fused SDPA call -> to_out[0] ->
+ residual -> / rescale_output_factor); diffusers is not in this image, so this mirror restates that sequence — **parity unpinned** —
and is checked against a stand-in with the same attribute names (tests/vae_standins.py).

One head of 512 channels over the 32,400 positions of a 1080p latent frame is two GEMMs around a row softmax: S = (Q / sqrt(d)) K^T
and O = P V run on the convolution kernel as 1 x 1 x 1 GEMMs (K and V^T are the "weights"), svr2_row_softmax_fwd sits between them
(S is materialised: 2.1 GB per frame).  torch's SDPA takes this head dimension only on its memory-efficient backend: 12.2 ms per
frame on a B200.  CUDA only — there is no other path."""
import math
from typing import Optional

import torch
from torch import nn

from . import _lib, ops
from .vae_block import groupnorm_silu
from .vae_conv import _DTYPE_CODE, CausalConv3dB200, causal_conv3d, transpose_2b

MAX_TOKENS = 32768      # S is materialised per frame: 2 x MAX_TOKENS^2 bytes (2.1 GB at this limit)


def row_softmax_(s: torch.Tensor, n_valid: int) -> torch.Tensor:
    """In-place softmax over the first ``n_valid`` columns of every row of the 2-D matrix ``s``; the padding columns become 0."""
    ops._need(s, "s", s.dtype)
    ops._call("svr2_row_softmax_fwd", 1, s.device, _lib.lib().svr2_row_softmax_fwd, s.data_ptr(), s.shape[0], int(n_valid), s.shape[1],
              _DTYPE_CODE[s.dtype], ops._stream(s))
    return s


class AttentionB200(nn.Module):
    """Drop-in for the mid block's ``Attention`` (one head, GroupNorm, residual connection); see the module docstring."""

    def __init__(self, group_norm: Optional[nn.GroupNorm], to_q: nn.Linear, to_k: nn.Linear, to_v: nn.Linear, to_out: nn.Linear,
                 residual_connection: bool = True, rescale_output_factor: float = 1.0):
        super().__init__()
        c = to_q.in_features
        if not (to_q.out_features == to_k.out_features == to_v.out_features == to_out.in_features == c == to_out.out_features):
            raise NotImplementedError("seedvr2_b200.vae_attn: one head as wide as the layer (inner_dim == channels) only")
        if c % 128:
            raise NotImplementedError("seedvr2_b200.vae_attn: channel count outside the kernel's granularity")
        self.channels = c
        self.norm = None if group_norm is None else (group_norm.num_groups, group_norm.eps)
        self.register_buffer("norm_weight", None if group_norm is None or group_norm.weight is None else group_norm.weight.detach().clone(),
                             persistent=False)
        self.register_buffer("norm_bias", None if group_norm is None or group_norm.bias is None else group_norm.bias.detach().clone(),
                             persistent=False)
        scale = 1.0 / math.sqrt(c)                      # SDPA's default scale, folded into the query projection

        def proj(lin: nn.Linear, k: float = 1.0) -> CausalConv3dB200:
            w = (lin.weight.detach().float() * k).to(lin.weight.dtype)[:, :, None, None, None]
            b = None if lin.bias is None else (lin.bias.detach().float() * k).to(lin.bias.dtype)
            return CausalConv3dB200(w, b)
        self.to_q, self.to_k, self.to_v, self.to_out = proj(to_q, scale), proj(to_k), proj(to_v), proj(to_out)
        self.residual_connection, self.rescale_output_factor = bool(residual_connection), float(rescale_output_factor)

    @classmethod
    def from_reference(cls, attn: nn.Module) -> "AttentionB200":
        if getattr(attn, "heads", 1) != 1 or getattr(attn, "norm_cross", None) is not None or getattr(attn, "spatial_norm", None) is not None \
                or getattr(attn, "added_kv_proj_dim", None) is not None:
            raise NotImplementedError("seedvr2_b200.vae_attn: single-head self-attention without cross / spatial norms only")
        if attn.to_q.weight.dtype not in _DTYPE_CODE:
            raise NotImplementedError("seedvr2_b200.vae_attn: fp16 / bf16 layers only")
        return cls(getattr(attn, "group_norm", None), attn.to_q, attn.to_k, attn.to_v, attn.to_out[0],
                   getattr(attn, "residual_connection", False), getattr(attn, "rescale_output_factor", 1.0))

    def forward(self, hidden_states: torch.Tensor, encoder_hidden_states=None, attention_mask=None, temb=None, **kwargs) -> torch.Tensor:
        if encoder_hidden_states is not None or attention_mask is not None:
            raise NotImplementedError("seedvr2_b200.vae_attn: self-attention without a mask (what the VAE's mid block calls)")
        if hidden_states.dim() != 4:
            raise NotImplementedError("seedvr2_b200.vae_attn: [(b f), C, H, W] input, as UNetMidBlock3D hands it over")
        x = hidden_states.permute(0, 2, 3, 1)                        # [F, H, W, C]: a view for the channels-last tensors of this library
        if not x.is_contiguous():
            x = x.contiguous()
        cd = self.to_q._compute_dtype()
        if x.dtype != cd:
            x = x.to(cd)
        F_, H, W, C = x.shape
        n = H * W
        if n > MAX_TOKENS:
            raise NotImplementedError(f"seedvr2_b200.vae_attn: {n} positions per frame exceed {MAX_TOKENS} (S is materialised per frame)")
        npad = -(-n // 128) * 128
        dev = x.device
        out = torch.empty_like(x)
        h = x
        if self.norm is not None:
            nw = None if self.norm_weight is None else self.norm_weight.to(cd)
            nb = None if self.norm_bias is None else self.norm_bias.to(cd)
            h = groupnorm_silu(x, nw, nb, self.norm[0], self.norm[1], silu=False)
        wq, bq = self.to_q._weights_for(cd)
        wk, bk = self.to_k._weights_for(cd)
        wv, bv = self.to_v._weights_for(cd)
        wo, bo = self.to_out._weights_for(cd)
        kpad = torch.zeros(npad, 1, 1, 1, C, dtype=cd, device=dev)   # K as the "weight" of the S GEMM, rows padded to the N tile
        vt = torch.zeros(C, 1, 1, 1, npad, dtype=cd, device=dev)     # V^T as the "weight" of the O GEMM, K padded alike
        s = torch.empty(1, H, W, npad, dtype=cd, device=dev)
        for f in range(F_):
            hf = h[f:f + 1]
            q = causal_conv3d(hf, wq, bq)
            causal_conv3d(hf, wk, bk, out=kpad.view(npad, C)[:n].view(1, H, W, C))
            v = causal_conv3d(hf, wv, bv)
            transpose_2b(v, vt, n, C, C, npad, 1, 0, 0)              # [n, C] -> [C, npad] (the padding columns stay zero)
            causal_conv3d(q, kpad, None, out=s)                      # S = Q K^T  [n, npad]
            row_softmax_(s.view(n, npad), n)                         # P, zeros in the padding
            o = causal_conv3d(s, vt, None)                           # O = P V   [n, C]
            causal_conv3d(o, wo, bo, None, x[f:f + 1] if self.residual_connection else None, out=out[f:f + 1])
        if self.rescale_output_factor != 1.0:
            out = out / self.rescale_output_factor
        return out.permute(0, 3, 1, 2)
