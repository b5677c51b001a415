"""Window plan: host mirror of the reference's per-forward window metadata.

Replaces what the reference memoises in its ``Cache`` per window method (src/models/dit_3b/nablocks/attention/
mmattn.py:187-221): window slices (window.py:28-83), ``na.window_idx`` (na.py:616-641), window lengths, cu_seqlens.
The geometry itself is computed by the C ABI (``svr2_window_plan`` — host, integer) and the index tensors by
``svr2_window_index`` on the device; this module only owns the buffers and caches plans by (grid, window, method, Lt).
"""
import ctypes
from collections import OrderedDict
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from . import _lib

WIN_FIELDS = 8
_MAX_WINDOWS = 1 << 16


class WindowPlan:
    """Geometry of one (grid, window triple, method, Lt).  Host arrays are numpy; device copies are made lazily."""

    def __init__(self, grid: Tuple[int, int, int], window: Tuple[int, int, int], shifted: bool, txt_len: int):
        L = _lib.lib()
        T, H, W = (int(v) for v in grid)
        nt, nh, nw = (int(v) for v in window)
        counts = (ctypes.c_int64 * 8)()
        # first call sizes the table
        _lib.check(L.svr2_window_plan(T, H, W, nt, nh, nw, int(shifted), int(txt_len), None, 0, counts), "svr2_window_plan")
        n = int(counts[0])
        if n <= 0 or n > _MAX_WINDOWS:
            raise RuntimeError(f"svr2_window_plan: unexpected window count {n}")
        table = np.zeros((n, WIN_FIELDS), dtype=np.int32)
        _lib.check(L.svr2_window_plan(T, H, W, nt, nh, nw, int(shifted), int(txt_len),
                                      table.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), n, counts), "svr2_window_plan")
        self.grid, self.window, self.shifted, self.txt_len = (T, H, W), (nt, nh, nw), bool(shifted), int(txt_len)
        self.table = table
        self.n_windows = n
        self.n_tokens = int(counts[1])
        self.max_len = int(counts[2])
        self.packed_rows = int(counts[3])
        self.min_len = int(counts[5])
        n_items = ctypes.c_int64(0)
        tp = table.ctypes.data_as(ctypes.POINTER(ctypes.c_int32))
        _lib.check(L.svr2_window_attn_worklist(tp, n, self.txt_len, None, 0, ctypes.byref(n_items)), "svr2_window_attn_worklist")
        work = np.zeros((int(n_items.value), 2), dtype=np.int32)
        _lib.check(L.svr2_window_attn_worklist(tp, n, self.txt_len, work.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)),
                                               work.shape[0], ctypes.byref(n_items)), "svr2_window_attn_worklist")
        self.work = work
        _lib.check(L.svr2_window_attn_worklist_tiles(tp, n, self.txt_len, 4, None, 0, ctypes.byref(n_items)), "svr2_window_attn_worklist_tiles")
        work4 = np.zeros((int(n_items.value), 2), dtype=np.int32)
        _lib.check(L.svr2_window_attn_worklist_tiles(tp, n, self.txt_len, 4, work4.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)),
                                                     work4.shape[0], ctypes.byref(n_items)), "svr2_window_attn_worklist_tiles")
        self.work4 = work4          # the same list in units of four query tiles (CTA-pair kernel)
        self._dev: Dict[torch.device, Dict[str, torch.Tensor]] = {}

    # ---- host-side views (reference vocabulary) -----------------------------------------------------------------
    @property
    def bounds(self) -> np.ndarray:
        """[nw, 6] = t0, t1, h0, h1, w0, w1 — the reference's slice triples (window.py:37-49)."""
        t = self.table
        return np.stack([t[:, 0], t[:, 0] + t[:, 3], t[:, 1], t[:, 1] + t[:, 4], t[:, 2], t[:, 2] + t[:, 5]], axis=1)

    @property
    def window_shape(self) -> np.ndarray:
        return self.table[:, 3:6].astype(np.int64)

    @property
    def window_lens(self) -> np.ndarray:
        return self.table[:, 7].astype(np.int64)

    @property
    def cu_seqlens(self) -> np.ndarray:
        cu = np.zeros(self.n_windows + 1, dtype=np.int32)
        cu[1:] = np.cumsum(self.window_lens + self.txt_len)
        return cu

    def attention_flops(self, heads: int, head_dim: int = 128) -> float:
        """heads * 4 * d * sum_w (Lw + Lt)^2 — the reference's own formula (attention.py:130-136)."""
        s = (self.window_lens + self.txt_len).astype(np.float64)
        return float(heads) * 4.0 * head_dim * float((s * s).sum())

    # ---- device-side tables -------------------------------------------------------------------------------------
    def on(self, device: torch.device) -> Dict[str, torch.Tensor]:
        device = torch.device(device)
        d = self._dev.get(device)
        if d is None:
            d = {
                "table": torch.from_numpy(self.table).to(device),
                "work": torch.from_numpy(self.work).to(device),
                "work4": torch.from_numpy(self.work4).to(device),
            }
            self._dev[device] = d
        return d

    def row_map(self, device: torch.device) -> torch.Tensor:
        """[P, 4] int32 packed-row map (svr2_pack_rowmap), built once per plan and device: source row, local (t, h, w),
        window — what every svr2_qkv_pack_fwd CTA would otherwise search the window table for."""
        d = self.on(device)
        if "row_map" not in d:
            if torch.device(device).type != "cuda":
                raise RuntimeError("svr2_pack_rowmap runs on a CUDA device only (no CPU fallback)")
            rm = torch.empty(self.packed_rows, 4, dtype=torch.int32, device=device)
            T, H, W = self.grid
            with torch.cuda.device(device):
                stream = torch.cuda.current_stream(device).cuda_stream
                _lib.check(_lib.lib().svr2_pack_rowmap(d["table"].data_ptr(), self.n_windows, T, H, W, self.txt_len,
                                                       rm.data_ptr(), self.packed_rows, stream), "svr2_pack_rowmap")
            d["row_map"] = rm
        return d["row_map"]

    def rope_plan(self, device: torch.device, rope_kind: int, freqs_key: Tuple[float, ...], freqs: Optional[torch.Tensor]):
        """Tables of the fused qkv projection (svr2_rope_plan), once per (device, rope kind, frequency values):
        (tok_map [L + Lt, 4] int32, rope_cs [rows, n_freqs, 2] fp32 or None, rows).  ``freqs_key`` = the frequency VALUES (the
        blocks of a model hold equal copies of the checkpoint buffer, so they share one table)."""
        d = self.on(device)
        key = ("rope_plan", int(rope_kind), tuple(freqs_key))
        if key not in d:
            if torch.device(device).type != "cuda":
                raise RuntimeError("svr2_rope_plan runs on a CUDA device only (no CPU fallback)")
            T, H, W = self.grid
            n_freqs = len(freqs_key)
            base_of_n = None
            rows = 0
            if rope_kind == 1:        # 3B: table row = position value
                rows = max(self.txt_len + int(self.table[:, 3].max()), int(self.table[:, 4].max()), int(self.table[:, 5].max()),
                           self.txt_len, 1)
            elif rope_kind == 2:      # 7B: rows for linspace(-1, 1, n) of every window extent n in the plan
                ns = sorted(set(int(v) for v in self.table[:, 3:6].reshape(-1)))
                if ns[-1] > 64:
                    raise NotImplementedError(f"svr2_rope_plan: window extent {ns[-1]} > 64")
                b = np.full(65, -1, dtype=np.int32)
                for n in ns:
                    b[n] = rows
                    rows += n
                base_of_n = torch.from_numpy(b).to(device)
            tok_map = torch.empty(self.n_tokens + self.txt_len, 4, dtype=torch.int32, device=device)
            cs = torch.empty(rows, n_freqs, 2, dtype=torch.float32, device=device) if rope_kind else None
            with torch.cuda.device(device):
                stream = torch.cuda.current_stream(device).cuda_stream
                _lib.check(_lib.lib().svr2_rope_plan(
                    d["table"].data_ptr(), self.row_map(device).data_ptr(), self.n_windows, T, H, W, self.txt_len, self.packed_rows,
                    int(rope_kind), None if freqs is None else freqs.data_ptr(), n_freqs,
                    None if base_of_n is None else base_of_n.data_ptr(), rows, tok_map.data_ptr(),
                    None if cs is None else cs.data_ptr(), stream), "svr2_rope_plan")
                torch.cuda.current_stream(device).synchronize()       # freqs / base_of_n may be temporaries of the caller
            d[key] = (tok_map, cs, rows)
        return d[key]

    def index_tensors(self, device: torch.device):
        """(tgt_idx, src_idx, window_shape, cu_seqlens) on `device`, computed by svr2_window_index (bit-exact)."""
        d = self.on(device)
        if "tgt_idx" not in d:
            if torch.device(device).type != "cuda":
                raise RuntimeError("svr2_window_index runs on a CUDA device only (no CPU fallback)")
            tgt = torch.empty(self.n_tokens, dtype=torch.int64, device=device)
            src = torch.empty(self.n_tokens, dtype=torch.int64, device=device)
            wshape = torch.empty(self.n_windows, 3, dtype=torch.int64, device=device)
            cu = torch.empty(self.n_windows + 1, dtype=torch.int32, device=device)
            T, H, W = self.grid
            with torch.cuda.device(device):
                stream = torch.cuda.current_stream(device).cuda_stream
                _lib.check(_lib.lib().svr2_window_index(d["table"].data_ptr(), self.n_windows, T, H, W, self.txt_len,
                                                        tgt.data_ptr(), src.data_ptr(), wshape.data_ptr(), cu.data_ptr(),
                                                        stream), "svr2_window_index")
            d.update(tgt_idx=tgt, src_idx=src, window_shape=wshape, cu_seqlens=cu)
        return d["tgt_idx"], d["src_idx"], d["window_shape"], d["cu_seqlens"]


_PLANS: "OrderedDict[tuple, WindowPlan]" = OrderedDict()
_PLAN_CAP = 64


def get_plan(grid, window, shifted: bool, txt_len: int) -> WindowPlan:
    key = (tuple(int(v) for v in grid), tuple(int(v) for v in window), bool(shifted), int(txt_len))
    p = _PLANS.get(key)
    if p is None:
        p = WindowPlan(key[0], key[1], key[2], key[3])
        _PLANS[key] = p
        if len(_PLANS) > _PLAN_CAP:
            _PLANS.popitem(last=False)
    else:
        _PLANS.move_to_end(key)
    return p
