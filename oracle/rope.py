"""RoPE angle tables for the packed window layout.  TEST INFRASTRUCTURE ONLY.

3B  ``mmrope3d``: reference src/models/dit_3b/rope.py:88-176 — lang frequencies (buffer ``attn.rope.rope.freqs`` [21]),
     video token at window-local (t,h,w) uses positions (Lt+t, h, w) (rope.py:172), text token j uses (j,j,j) (:173).
7B  pixel rope : reference src/models/dit_7b/rope.py:73-111 — pixel frequencies (buffer [10]), per-axis positions
     linspace(-1,1,n) over the WINDOW's own extent, video only, angles cast to the activation dtype (rope.py:85).

Angle layout follows the third-party ``get_axial_freqs`` (restated in oracle/shim): per axis the vector pos*freqs with
every entry repeated twice (interleaved), axes concatenated (t,h,w).
"""
from typing import Sequence, Tuple

import torch


def _axis_angles(pos: torch.Tensor, freqs: torch.Tensor) -> torch.Tensor:
    return torch.repeat_interleave(pos.to(freqs.dtype)[:, None] * freqs[None, :], 2, dim=-1)


def angles_3b(window_shape: Sequence[Tuple[int, int, int]], txt_len: int, freqs: torch.Tensor):
    """-> (vid_angles [sum Lw, 6*len(freqs)], txt_angles [txt_len, 6*len(freqs)]) fp32, built on the CPU (the caller moves
    them): the tables are then bit-identical whatever device the activations live on."""
    freqs = freqs.float().cpu()
    out = []
    for (f, h, w) in window_shape:
        at = _axis_angles(torch.arange(txt_len, txt_len + f), freqs)[:, None, None, :].expand(f, h, w, -1)
        ah = _axis_angles(torch.arange(h), freqs)[None, :, None, :].expand(f, h, w, -1)
        aw = _axis_angles(torch.arange(w), freqs)[None, None, :, :].expand(f, h, w, -1)
        out.append(torch.cat([at, ah, aw], dim=-1).reshape(f * h * w, -1))
    txt = _axis_angles(torch.arange(txt_len), freqs).repeat(1, 3)
    return torch.cat(out, dim=0), txt


def angles_7b(window_shape: Sequence[Tuple[int, int, int]], freqs: torch.Tensor) -> torch.Tensor:
    """-> vid_angles [sum Lw, 6*len(freqs)] fp32 on the CPU (caller moves it and casts to the activation dtype as the
    reference does)."""
    freqs = freqs.float().cpu()
    out = []
    for (f, h, w) in window_shape:
        at = _axis_angles(torch.linspace(-1, 1, steps=f), freqs)[:, None, None, :].expand(f, h, w, -1)
        ah = _axis_angles(torch.linspace(-1, 1, steps=h), freqs)[None, :, None, :].expand(f, h, w, -1)
        aw = _axis_angles(torch.linspace(-1, 1, steps=w), freqs)[None, None, :, :].expand(f, h, w, -1)
        out.append(torch.cat([at, ah, aw], dim=-1).reshape(f * h * w, -1))
    return torch.cat(out, dim=0)


def rotate(x: torch.Tensor, angles: torch.Tensor) -> torch.Tensor:
    """x [L, heads, d] (any float dtype), angles [L, rot] -> rotated, same dtype as x.

    Interleaved-pair rotation in fp32 over the first ``rot`` dims, remaining dims pass through, result cast back
    (reference rope.py:118-119 ``apply_rotary_emb(freqs, q.float()).to(q.dtype)`` + third-party apply_rotary_emb).
    cos/sin are evaluated in the dtype of ``angles`` (bf16 for 7B, fp32 for 3B) exactly like ``freqs.cos()``.
    """
    rot = angles.shape[-1]
    xf = x.float()
    mid = xf[..., :rot]
    pair = mid.reshape(*mid.shape[:-1], rot // 2, 2)
    half = torch.stack((-pair[..., 1], pair[..., 0]), dim=-1).reshape(mid.shape)
    cos, sin = angles.cos()[:, None, :], angles.sin()[:, None, :]
    out = torch.cat([mid * cos + half * sin, xf[..., rot:]], dim=-1)
    return out.to(x.dtype)
