"""Restatement of the slice of lucidrains' ``rotary_embedding_torch`` the reference calls.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  The real package is a third-party
dependency pinned only as ``rotary_embedding_torch>=0.5.3`` (reference requirements.txt:11,
pyproject.toml:38) and is NOT installed in this image, so the reference cannot be imported
without it.  This file restates the published v0.5.3 algorithm for exactly the call sites the
reference uses:

* ``RotaryEmbedding(dim, freqs_for="lang", theta=10000)``   src/models/dit_3b/rope.py:77-81
* ``RotaryEmbedding(dim, freqs_for="pixel", max_freq=256)`` src/models/dit_7b/rope.py:28-32
* ``.get_axial_freqs(*dims)``                               dit_3b/rope.py:46, dit_7b/rope.py:46
* ``apply_rotary_emb(freqs, t)``                            dit_3b/rope.py:118-126, dit_7b/rope.py:88-89

PARITY UNPINNED at this boundary: the reference has no tests or golden vectors for it.
"""
from math import pi

import torch
from torch import nn


def rotate_half(x: torch.Tensor) -> torch.Tensor:
    # interleaved pairs (x0, x1) -> (-x1, x0)
    x = x.reshape(*x.shape[:-1], x.shape[-1] // 2, 2)
    x1, x2 = x.unbind(dim=-1)
    x = torch.stack((-x2, x1), dim=-1)
    return x.reshape(*x.shape[:-2], -1)


def apply_rotary_emb(freqs, t, start_index=0, scale=1.0, seq_dim=-2, freqs_seq_dim=None):
    dtype = t.dtype
    if freqs_seq_dim is None and (freqs.ndim == 2 or t.ndim == 3):
        freqs_seq_dim = 0
    if t.ndim == 3 or freqs_seq_dim is not None:
        seq_len = t.shape[seq_dim]
        freqs = freqs.narrow(freqs_seq_dim, freqs.shape[freqs_seq_dim] - seq_len, seq_len)
    rot_dim = freqs.shape[-1]
    end_index = start_index + rot_dim
    assert rot_dim <= t.shape[-1]
    t_left = t[..., :start_index]
    t_middle = t[..., start_index:end_index]
    t_right = t[..., end_index:]
    t_transformed = (t_middle * freqs.cos() * scale) + (rotate_half(t_middle) * freqs.sin() * scale)
    out = torch.cat((t_left, t_transformed, t_right), dim=-1)
    return out.type(dtype)


class RotaryEmbedding(nn.Module):
    def __init__(self, dim, custom_freqs=None, freqs_for="lang", theta=10000, max_freq=10,
                 num_freqs=1, learned_freq=False, **kwargs):
        super().__init__()
        if custom_freqs is not None:
            freqs = custom_freqs
        elif freqs_for == "lang":
            freqs = 1.0 / (theta ** (torch.arange(0, dim, 2)[: (dim // 2)].float() / dim))
        elif freqs_for == "pixel":
            freqs = torch.linspace(1.0, max_freq / 2, dim // 2) * pi
        elif freqs_for == "constant":
            freqs = torch.ones(num_freqs).float()
        else:
            raise ValueError(freqs_for)
        self.freqs_for = freqs_for
        self.freqs = nn.Parameter(freqs, requires_grad=learned_freq)

    def forward(self, t: torch.Tensor, seq_len=None, offset=0):
        freqs = self.freqs
        freqs = torch.einsum("..., f -> ... f", t.type(freqs.dtype), freqs)
        freqs = torch.repeat_interleave(freqs, 2, dim=-1)
        return freqs

    def get_axial_freqs(self, *dims):
        all_freqs = []
        for ind, dim in enumerate(dims):
            if self.freqs_for == "pixel":
                pos = torch.linspace(-1, 1, steps=dim, device=self.freqs.device)
            else:
                pos = torch.arange(dim, device=self.freqs.device)
            freqs = self.forward(pos, seq_len=dim)
            all_axis = [None] * len(dims)
            all_axis[ind] = slice(None)
            new_axis_slice = (Ellipsis, *all_axis, slice(None))
            all_freqs.append(freqs[new_axis_slice])
        all_freqs = torch.broadcast_tensors(*all_freqs)
        return torch.cat(all_freqs, dim=-1)
