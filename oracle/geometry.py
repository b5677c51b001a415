"""Window geometry and partition indices (integer, bit-exact).  TEST INFRASTRUCTURE ONLY.

Follows reference src/models/dit_3b/window.py:28-83 (identical copy in dit_7b/window.py) for the bounds and
src/models/dit_3b/na.py:583-641 (``window`` / ``window_idx``) for the index tensors.  Python is used on purpose:
the reference relies on Python's banker's ``round``, ``math.ceil`` of float quotients and ``int()`` truncation
(SURVEY quirk Q4), which are reproduced here by the very same built-ins.
"""
from math import ceil, sqrt
from typing import List, Tuple

import numpy as np

Bounds = Tuple[int, int, int, int, int, int]  # t0, t1, h0, h1, w0, w1


def window_sizes(t: int, h: int, w: int, num_windows=(4, 3, 3)) -> Tuple[int, int, int]:
    """(wt, wh, ww): window extent derived from the 720p-normalised grid.  window.py:32-36."""
    nt, nh, nw = num_windows
    scale = sqrt((45 * 80) / (h * w))
    rh, rw = round(h * scale), round(w * scale)
    return ceil(min(t, 30) / nt), ceil(rh / nh), ceil(rw / nw)


def _axis_regular(size: int, win: int) -> List[Tuple[int, int]]:
    n = ceil(size / win)
    out = []
    for i in range(n):
        lo, hi = i * win, min((i + 1) * win, size)
        if hi > lo:
            out.append((lo, hi))
    return out


def _axis_shifted(size: int, win: int) -> List[Tuple[int, int]]:
    s = 0.5 if win < size else 0
    n = ceil((size - s) / win)
    n = n + 1 if s > 0 else 1
    out = []
    for i in range(n):
        lo, hi = max(int((i - s) * win), 0), min(int((i - s + 1) * win), size)
        if hi > lo:
            out.append((lo, hi))
    return out


def window_bounds(t: int, h: int, w: int, num_windows=(4, 3, 3), shifted: bool = False) -> List[Bounds]:
    """All windows in the reference's order: iw outer, ih middle, it inner.  window.py:37-49,70-83."""
    wt, wh, ww = window_sizes(t, h, w, num_windows)
    axis = _axis_shifted if shifted else _axis_regular
    ts, hs, ws = axis(t, wt), axis(h, wh), axis(w, ww)
    return [(t0, t1, h0, h1, w0, w1) for (w0, w1) in ws for (h0, h1) in hs for (t0, t1) in ts]


def window_index(t: int, h: int, w: int, num_windows=(4, 3, 3), shifted: bool = False, txt_len: int = 0):
    """tgt_idx, src_idx, window_shape, cu_seqlens  — na.py:616-641 + mmattn.py:211-255 for one sample (b=1).

    tgt_idx[k]   = global token id ((t*H + h)*W + w) of packed position k (windows concatenated, row-major inside)
    src_idx      = inverse permutation (the reference's argsort of a permutation)
    window_shape = [nw, 3] extents;  cu_seqlens[k+1]-cu_seqlens[k] = Lw_k + txt_len (int32)
    """
    bounds = window_bounds(t, h, w, num_windows, shifted)
    grid = np.arange(t * h * w, dtype=np.int64).reshape(t, h, w)
    parts = [grid[t0:t1, h0:h1, w0:w1].reshape(-1) for (t0, t1, h0, h1, w0, w1) in bounds]
    tgt = np.concatenate(parts)
    src = np.empty_like(tgt)
    src[tgt] = np.arange(tgt.size, dtype=np.int64)
    shape = np.array([(t1 - t0, h1 - h0, w1 - w0) for (t0, t1, h0, h1, w0, w1) in bounds], dtype=np.int64)
    lens = shape.prod(-1) + txt_len
    cu = np.zeros(len(bounds) + 1, dtype=np.int32)
    cu[1:] = np.cumsum(lens)
    return tgt, src, shape, cu


def attention_flops(t: int, h: int, w: int, heads: int, head_dim: int = 128, txt_len: int = 58,
                    num_windows=(4, 3, 3), shifted: bool = False) -> float:
    """Algorithmic FLOPs of one window-attention call, the reference's own formula
    (src/models/dit_3b/attention.py:130-136): heads * 4 * d * sum_w (Lw + Lt)^2."""
    b = window_bounds(t, h, w, num_windows, shifted)
    s = sum(((t1 - t0) * (h1 - h0) * (w1 - w0) + txt_len) ** 2 for (t0, t1, h0, h1, w0, w1) in b)
    return float(heads) * 4.0 * head_dim * s
