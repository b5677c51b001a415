"""B200-native drop-ins for the reference's ``NaMMSRTransformerBlock`` (3B and 7B).

Boundary: the reference instantiates ``get_nablock("mmdit_sr")(**kwargs)`` per layer
(src/models/dit_3b/nadit.py:135-165, dit_7b/nadit.py:109-136) and calls
``block(vid=, txt=, vid_shape=, txt_shape=, emb=, cache=)`` (dit_3b/nadit.py:221-231).  The classes below keep
* the constructor keyword arguments of src/models/dit_3b/nablocks/mmsr_block.py:31-52 and
  src/models/dit_7b/nablocks/mmsr_block.py:161-180,
* every parameter / buffer NAME and shape of the reference block (``attn.proj_qkv.vid.weight`` ... incl. the persistent
  ``attn.rope.rope.freqs`` buffer), so checkpoints load with ``load_state_dict(..., assign=True)`` unchanged,
* the forward signature and return tuple of mmsr_block.py:84-128.

What differs is how the work is done: GEMMs stay torch/cuBLAS calls; everything else is the C-ABI CUDA path
(ops.py) — 6 launches per block instead of ~60 ATen kernels plus one SDPA call per window.
"""
import math
import os
import weakref
from typing import Callable, Optional, Tuple

import torch
import torch.nn.functional as F
from torch import nn

from . import ops
from .plan import WindowPlan, get_plan

_SHIFTED = {"720pwin_by_size_bysize": False, "720pswin_by_size_bysize": True}
# The block's linear layers run on the library's own tcgen05 GEMM (csrc/gemm.cu) with the following row-wise op fused into the
# epilogue.  USE_TC_GEMM = False restores round 1's library GEMMs (cuBLAS through torch) + separate elementwise kernels — kept
# ONLY so bench.py can time both arms in one process; it is still all-CUDA and never the default.
USE_TC_GEMM = os.environ.get("SVR2_LIBRARY_GEMM", "0") != "1"
# proj_qkv + q/k RMSNorm + RoPE + window partition as ONE kernel (svr2_linear_qkv_pack_fwd).  SVR2_UNFUSED_QKV=1 restores
# svr2_linear_fwd + svr2_qkv_pack_fwd (same results; kept for A/B timing and for configurations the fused epilogue does not cover).
USE_FUSED_QKV = os.environ.get("SVR2_UNFUSED_QKV", "0") != "1"


def _linear(x, w, b=None):
    if USE_TC_GEMM and x.is_cuda and x.dim() == 2 and w.shape[0] % 64 == 0 and w.shape[1] % 8 == 0:
        return ops.linear(x.contiguous(), _bf16(w), None if b is None else _bf16(b))
    return F.linear(x, _bf16(w), None if b is None else _bf16(b))


def _factory_kind(f) -> Optional[str]:
    """The reference passes ``norm`` / ``qk_norm`` as closures of get_norm_layer(norm_type) (normalization.py:112-149)
    and ``ada`` as the AdaSingle class (modulation.py:29-32); recover the kind to reject what is not implemented."""
    if f is None or isinstance(f, str):
        return f
    cl = getattr(f, "__closure__", None)
    if cl:
        for c in cl:
            if isinstance(c.cell_contents, (str, type(None))):
                return c.cell_contents
    return getattr(f, "__name__", None)


class _Slot(nn.Module):
    """Parameter holder standing in for the reference's MMModule (src/models/dit_3b/mm.py:35-74): children are named
    ``vid`` / ``txt`` (separate weights) or ``all`` (shared) so state-dict keys match."""

    def __init__(self, make: Callable[[], nn.Module], shared: bool, vid_only: bool = False):
        super().__init__()
        self.shared_weights, self.vid_only = shared, vid_only
        if shared:
            self.all = make()
        else:
            self.vid = make()
            self.txt = None if vid_only else make()

    def pick(self, branch: str) -> nn.Module:
        return self.all if self.shared_weights else getattr(self, branch)


class _Weight(nn.Module):
    def __init__(self, dim: int):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(dim))


class _AdaParams(nn.Module):
    """AdaSingle parameters, layers=["attn","mlp"] (modulation.py:43-63)."""

    def __init__(self, dim: int):
        super().__init__()
        for layer in ("attn", "mlp"):
            setattr(self, f"{layer}_shift", nn.Parameter(torch.randn(dim) / dim ** 0.5))
            setattr(self, f"{layer}_scale", nn.Parameter(torch.randn(dim) / dim ** 0.5 + 1))
            setattr(self, f"{layer}_gate", nn.Parameter(torch.randn(dim) / dim ** 0.5))


class _Freqs(nn.Module):
    def __init__(self, freqs: torch.Tensor):
        super().__init__()
        self.register_buffer("freqs", freqs)     # persistent: it IS in the checkpoints


class _Rope(nn.Module):
    """Holds ``rope.freqs`` under the reference's name.  No ``get_axial_freqs`` attribute on purpose: the product's
    RoPE "stabilisation" / BlockSwap wrappers are hasattr-guarded (compatibility.py:388-408, blockswap.py:556-631)."""

    def __init__(self, kind: int, head_dim: int):
        super().__init__()
        if kind == ops.ROPE_3B_MM:   # RotaryEmbedding(dim=rope_dim//3, freqs_for="lang", theta=10000)  dit_3b/rope.py:77-81
            dim = head_dim // 3
            freqs = 1.0 / (10000 ** (torch.arange(0, dim, 2)[: dim // 2].float() / dim))
        else:                        # RotaryEmbedding(dim=(head_dim//2)//3, freqs_for="pixel", max_freq=256)  dit_7b/rope.py:28-32
            dim = (head_dim // 2) // 3
            freqs = torch.linspace(1.0, 256 / 2, dim // 2, dtype=torch.float32) * math.pi
        self.kind = kind
        self.mm = kind == ops.ROPE_3B_MM
        self.rope = _Freqs(freqs)


class _SwiGLU(nn.Module):      # reference mlp.py:46-62
    def __init__(self, dim: int, expand_ratio: int, multiple_of: int = 256):
        super().__init__()
        hidden = int(2 * dim * expand_ratio / 3)
        hidden = multiple_of * ((hidden + multiple_of - 1) // multiple_of)
        self.proj_in_gate = nn.Linear(dim, hidden, bias=False)
        self.proj_out = nn.Linear(hidden, dim, bias=False)
        self.proj_in = nn.Linear(dim, hidden, bias=False)

    def hidden(self, x):
        """silu(proj_in_gate(x)) * proj_in(x): ONE GEMM with two B operands and the activation in its epilogue."""
        if USE_TC_GEMM and x.is_cuda and x.dim() == 2:
            return ops.linear_swiglu(x.contiguous(), _bf16(self.proj_in_gate.weight), _bf16(self.proj_in.weight))
        return ops.swiglu(F.linear(x, _bf16(self.proj_in_gate.weight)), F.linear(x, _bf16(self.proj_in.weight)))

    def forward(self, x):
        return _linear(self.hidden(x), self.proj_out.weight)


class _GeluMLP(nn.Module):     # reference mlp.py:28-43
    def __init__(self, dim: int, expand_ratio: int):
        super().__init__()
        self.proj_in = nn.Linear(dim, dim * expand_ratio)
        self.proj_out = nn.Linear(dim * expand_ratio, dim)

    def hidden(self, x):
        """gelu_tanh(proj_in(x)): bias + activation in the GEMM epilogue; the [L, 4D] pre-activation never reaches HBM."""
        if USE_TC_GEMM and x.is_cuda and x.dim() == 2:
            return ops.linear(x.contiguous(), _bf16(self.proj_in.weight), _bf16(self.proj_in.bias), ops.ACT_GELU_TANH)
        return F.gelu(F.linear(x, _bf16(self.proj_in.weight), _bf16(self.proj_in.bias)), approximate="tanh")

    def forward(self, x):
        return _linear(self.hidden(x), self.proj_out.weight, self.proj_out.bias)


def _proj_gate_residual(x, lin: nn.Linear, e, gate_b, layer: int, resid):
    """lin(x) * (gateA[layer] + gateB) + resid — proj_out + AdaSingle "out" + residual (mmattn.py:269 / mlp.py:43,62,
    modulation.py:111-116, mmsr_block.py:114,126) as ONE GEMM; gate_b None = plain residual (3B last layer's text rows)."""
    bias = None if lin.bias is None else _bf16(lin.bias)
    if USE_TC_GEMM and gate_b is not None:
        return ops.linear_gate_residual(x.contiguous(), _bf16(lin.weight), bias, e, _bf16(gate_b), layer, resid.contiguous())
    h = _linear(x, lin.weight, lin.bias)
    return ops.gate_residual(h, resid.contiguous(), e if gate_b is not None else None, None if gate_b is None else _bf16(gate_b), layer)


_cast_cache = {}      # id(tensor) -> (weakref, version, data_ptr, device, bf16 copy)


def _bf16(t: torch.Tensor) -> torch.Tensor:
    """Parameters are used as bf16.  Checkpoints loaded with ``assign=True`` keep the file's dtype (fp16 / fp32 / fp8,
    model_loader.py:809, model_registry.py:34-49; the reference casts such parameters per call, modulation.py:95-107): they
    are cast ONCE, lazily, and the copy is reused until the parameter is replaced, modified in place or moved (BlockSwap)."""
    if t.dtype == torch.bfloat16:
        return t
    key = id(t)
    hit = _cast_cache.get(key)
    if hit is not None and hit[0]() is t and hit[1] == t._version and hit[2] == t.data_ptr() and hit[3] == t.device:
        return hit[4]
    c = t.detach().to(torch.bfloat16)
    _cast_cache[key] = (weakref.ref(t, lambda _r, k=key: _cast_cache.pop(k, None)), t._version, t.data_ptr(), t.device, c)
    return c


def _joined(a: torch.Tensor, b: torch.Tensor) -> Optional[torch.Tensor]:
    """[a; b] as ONE tensor without a copy when `b`'s rows directly follow `a`'s in the same storage, else None.
    Shared-weight layers (3B layers >= mm_layers, mm.py:59-74 routes vid and txt through the same ``.all`` module) apply
    the same row-wise function to both branches; run on the joined rows that is one GEMM / one launch instead of two
    ("vid+txt concatenated for shared-weight layers") and the weights are read once."""
    if a.dim() != 2 or b.dim() != 2 or a.shape[1] != b.shape[1] or a.dtype != b.dtype or a.device != b.device:
        return None
    if not (a.is_contiguous() and b.is_contiguous()) or a.numel() == 0 or b.numel() == 0:
        return None
    if a.untyped_storage().data_ptr() != b.untyped_storage().data_ptr() or b.storage_offset() != a.storage_offset() + a.numel():
        return None
    return torch.as_strided(a, (a.shape[0] + b.shape[0], a.shape[1]), (a.shape[1], 1))


class _GridMemo:
    """vid_shape lives on the GPU (na.py:97); reading it is a D2H sync the reference pays several times per layer.
    The same tensor OBJECT flows through all blocks of a forward, so one sync per forward is enough — also for 7B,
    whose forward disables the reference Cache (quirk Q2).  Keyed per tensor object (weak reference + version), so several
    models driven from one process — other threads, other streams — do not evict each other's entry."""
    _memo = {}      # id(vid_shape) -> (weakref, version, (grid, txt_len))

    @classmethod
    def read(cls, vid_shape: torch.Tensor, txt_shape: torch.Tensor):
        key = id(vid_shape)
        hit = cls._memo.get(key)
        if hit is not None and hit[0]() is vid_shape and hit[1] == vid_shape._version:
            return hit[2]
        vs = vid_shape.tolist()
        ts = txt_shape.tolist()
        if len(vs) != 1 or len(ts) != 1:
            # the pipeline always runs b=1 (generation_phases.py:678,701); b>1 hits reference quirk Q7
            raise NotImplementedError("seedvr2_b200 blocks support batch size 1 only")
        value = (tuple(int(v) for v in vs[0]), int(math.prod(ts[0])))
        cls._memo[key] = (weakref.ref(vid_shape, lambda _r, k=key: cls._memo.pop(k, None)), vid_shape._version, value)
        return value


class NaSwinAttentionB200(nn.Module):
    """Window attention of one block.  Reference: src/models/dit_3b/nablocks/attention/mmattn.py:145-271 and
    src/models/dit_7b/nablocks/mmsr_block.py:34-157."""

    def __init__(self, *, vid_dim: int, txt_dim: int, heads: int, head_dim: int, qk_bias: bool, qk_norm_eps: float,
                 rope_kind: int, shared_weights: bool, shared_qknorm: bool, window, window_method: str):
        super().__init__()
        if head_dim != 128:
            raise NotImplementedError("seedvr2_b200: head_dim must be 128")
        if vid_dim != txt_dim:
            raise NotImplementedError("seedvr2_b200: vid_dim != txt_dim is not supported")
        if qk_bias:
            raise NotImplementedError("seedvr2_b200: qk_bias=True is not used by the 3B/7B configs")
        if window_method not in _SHIFTED:
            raise NotImplementedError(f"seedvr2_b200: window_method {window_method!r} is not supported")
        inner = heads * head_dim
        self.heads, self.head_dim, self.eps = heads, head_dim, float(qk_norm_eps)
        self.window = tuple(int(v) for v in (window if isinstance(window, (tuple, list)) else (window,) * 3))
        self.window_method = window_method
        self.shifted = _SHIFTED[window_method]
        self.proj_qkv = _Slot(lambda: nn.Linear(vid_dim, inner * 3, bias=False), shared_weights)
        self.proj_out = _Slot(lambda: nn.Linear(inner, vid_dim), shared_weights)
        self.norm_q = _Slot(lambda: _Weight(head_dim), shared_qknorm)
        self.norm_k = _Slot(lambda: _Weight(head_dim), shared_qknorm)
        self.rope = _Rope(rope_kind, head_dim) if rope_kind != ops.ROPE_NONE else None

    def _freqs(self):
        """(fp32 frequencies on the device, their values as a tuple — read ONCE per buffer: the only host sync of the fused path)."""
        f = self.rope.rope.freqs
        memo = self.__dict__.get("_freqs_memo")
        if memo is None or memo[0] is not f or memo[1] != f._version or memo[2] != f.data_ptr():
            f32 = f if f.dtype == torch.float32 else f.float()
            memo = (f, f._version, f.data_ptr(), f32, tuple(float(v) for v in f32.tolist()))
            self.__dict__["_freqs_memo"] = memo
        return memo[3], memo[4]

    def core(self, vid: torch.Tensor, txt: torch.Tensor, plan: WindowPlan):
        """proj_qkv -> pack (norm+rope+partition) -> window attention; returns attention outputs BEFORE proj_out."""
        if USE_TC_GEMM and USE_FUSED_QKV and vid.is_cuda:
            if self.rope is not None:
                (freqs, fkey), kind = self._freqs(), self.rope.kind
            else:
                freqs, fkey, kind = None, (), ops.ROPE_NONE
            if ops.qkv_fused_supported(plan, kind, len(fkey), self.heads):
                # ONE GEMM per weight set whose epilogue normalises, rotates and scatters q, k, v into the packed window layout
                nw = (_bf16(self.norm_q.pick("vid").weight), _bf16(self.norm_k.pick("vid").weight),
                      _bf16(self.norm_q.pick("txt").weight), _bf16(self.norm_k.pick("txt").weight))
                both = _joined(vid, txt) if self.proj_qkv.shared_weights else None
                if both is not None:
                    bufs = ops.linear_qkv_pack(plan, both, _bf16(self.proj_qkv.all.weight), 0, *nw, freqs, fkey, kind, self.heads, self.eps)
                else:
                    bufs = ops.linear_qkv_pack(plan, vid.contiguous(), _bf16(self.proj_qkv.pick("vid").weight), 0, *nw, freqs, fkey,
                                               kind, self.heads, self.eps)
                    bufs = ops.linear_qkv_pack(plan, txt.contiguous(), _bf16(self.proj_qkv.pick("txt").weight), plan.n_tokens, *nw,
                                               freqs, fkey, kind, self.heads, self.eps, out=bufs)
                q, k, v, tq = bufs
                ops.replicate_text(plan, tq, q, k, v, self.heads)
                return ops.window_attn(plan, q, k, v, self.heads)
        both = _joined(vid, txt) if self.proj_qkv.shared_weights else None
        if both is not None:
            qkv = _linear(both, self.proj_qkv.all.weight)
            vqkv, tqkv = qkv[:vid.shape[0]], qkv[vid.shape[0]:]
        else:
            vqkv = _linear(vid, self.proj_qkv.pick("vid").weight)
            tqkv = _linear(txt, self.proj_qkv.pick("txt").weight)
        if self.rope is not None:
            freqs = self.rope.rope.freqs
            if freqs.dtype != torch.float32:
                freqs = freqs.float()
            kind = self.rope.kind
        else:
            freqs, kind = None, ops.ROPE_NONE
        q, k, v = ops.qkv_pack(plan, vqkv, tqkv, _bf16(self.norm_q.pick("vid").weight), _bf16(self.norm_k.pick("vid").weight),
                               _bf16(self.norm_q.pick("txt").weight), _bf16(self.norm_k.pick("txt").weight), freqs, kind,
                               self.heads, self.eps)
        return ops.window_attn(plan, q, k, v, self.heads)

    def forward(self, vid, txt, vid_shape, txt_shape, cache=None):
        grid, txt_len = _GridMemo.read(vid_shape, txt_shape)
        plan = get_plan(grid, self.window, self.shifted, txt_len)
        ov, ot = self.core(vid, txt, plan)
        # Unfused tail, kept for callers of the attention module on its own (the product's attention wrappers,
        # compatibility.py:437-522).  The block itself calls core() and fuses proj_out with the gate + residual.
        both = _joined(ov, ot) if self.proj_out.shared_weights else None
        if both is not None:
            o = _linear(both, self.proj_out.all.weight, self.proj_out.all.bias)
            return o[:ov.shape[0]], o[ov.shape[0]:]
        po_v, po_t = self.proj_out.pick("vid"), self.proj_out.pick("txt")
        return _linear(ov, po_v.weight, po_v.bias), _linear(ot, po_t.weight, po_t.bias)


class _BlockBase(nn.Module):
    """Shared forward: norm->mod->attn->gate->+res ; norm->mod->mlp->gate->+res (mmsr_block.py:98-128)."""

    txt_tail_skipped = False      # 3B last layer: ada / mlp_norm / mlp are vid_only

    def _ada(self, branch: str) -> Optional[_AdaParams]:
        return self.ada.pick(branch)

    def forward(self, vid, txt, vid_shape, txt_shape, emb, cache=None):
        if vid.dtype != torch.bfloat16:
            raise NotImplementedError(f"seedvr2_b200 blocks compute in bf16 (got {vid.dtype}); there is no other path")
        grid, txt_len = _GridMemo.read(vid_shape, txt_shape)
        if cache is not None:
            # Honour the reference Cache (src/common/cache.py): same keys the reference blocks fill.  Quirk Q1:
            # dit_3b/nadit.py:236-244 later asks for "emb_repeat_0_vid" and must find the ATTN-layer triple [L, D, 3].
            cache("vid_len", lambda: vid_shape.prod(-1))
            cache("txt_len", lambda: txt_shape.prod(-1))
            cache("emb_repeat_0_vid", lambda: emb.reshape(emb.shape[0], -1, 2, 3)[:, :, 0, :].expand(vid.shape[0], -1, -1))
        plan = get_plan(grid, self.attn.window, self.attn.shifted, txt_len)
        if plan.n_tokens != vid.shape[0] or txt_len != txt.shape[0]:
            raise RuntimeError(f"vid/txt rows {vid.shape[0]}/{txt.shape[0]} do not match vid_shape {grid} / txt_shape {txt_len}")
        e = _bf16(emb).reshape(-1).contiguous()
        vid, txt = vid.contiguous(), txt.contiguous()
        eps = self.norm_eps
        av, at = self._ada("vid"), (None if self.txt_tail_skipped else self._ada("txt"))
        L = vid.shape[0]

        attn = self.attn
        if av is at and self.mlp.shared_weights:
            # Shared-weight layer: ada / mlp (and attn) hold ONE module for both branches, so every row-wise step and every
            # GEMM runs once on the joined rows [vid; txt].  The joined tensor is free when the previous layer produced it;
            # the first shared layer pays one copy.
            x = _joined(vid, txt)
            if x is None:
                x = torch.cat([vid, txt])
            a = ops.modnorm(x, e, _bf16(av.attn_shift), _bf16(av.attn_scale), 0, eps)
            o_v, o_t = attn.core(a[:L], a[L:], plan)
            o = _joined(o_v, o_t)
            if o is None:
                o = torch.cat([o_v, o_t])
            x1 = _proj_gate_residual(o, attn.proj_out.all, e, av.attn_gate, 0, x)            # GEMM + gate + residual
            m = ops.modnorm(x1, e, _bf16(av.mlp_shift), _bf16(av.mlp_scale), 1, eps)
            x2 = _proj_gate_residual(self.mlp.all.hidden(m), self.mlp.all.proj_out, e, av.mlp_gate, 1, x1)
            return x2[:L], x2[L:], vid_shape, txt_shape

        # --- attention half
        a_v = ops.modnorm(vid, e, _bf16(av.attn_shift), _bf16(av.attn_scale), 0, eps)
        a_t = (ops.modnorm(txt, e, _bf16(at.attn_shift), _bf16(at.attn_scale), 0, eps) if at is not None
               else ops.modnorm(txt, None, None, None, 0, eps))
        o_v, o_t = attn.core(a_v, a_t, plan)
        # --- proj_out + gate + residual in one GEMM, then the MLP half's norm + modulation
        vid1 = _proj_gate_residual(o_v, attn.proj_out.pick("vid"), e, av.attn_gate, 0, vid)
        m_v = ops.modnorm(vid1, e, _bf16(av.mlp_shift), _bf16(av.mlp_scale), 1, eps)
        mlp_v = self.mlp.pick("vid")
        vid2 = _proj_gate_residual(mlp_v.hidden(m_v), mlp_v.proj_out, e, av.mlp_gate, 1, vid1)
        if at is not None:
            txt1 = _proj_gate_residual(o_t, attn.proj_out.pick("txt"), e, at.attn_gate, 0, txt)
            m_t = ops.modnorm(txt1, e, _bf16(at.mlp_shift), _bf16(at.mlp_scale), 1, eps)
            mlp_t = self.mlp.pick("txt")
            txt2 = _proj_gate_residual(mlp_t.hidden(m_t), mlp_t.proj_out, e, at.mlp_gate, 1, txt1)
        else:
            # vid_only MMModules pass txt through untouched (mm.py:69-73): txt1 = attn(txt) + txt ; txt2 = txt1 + txt1
            txt1 = _proj_gate_residual(o_t, attn.proj_out.pick("txt"), None, None, 0, txt)
            txt2 = ops.gate_residual(txt1, txt1, None, None, 1)
        return vid2, txt2, vid_shape, txt_shape


class NaMMSRTransformerBlockB200(_BlockBase):
    """SeedVR2-3B block.  Constructor kwargs of src/models/dit_3b/nablocks/mmsr_block.py:31-52."""

    def __init__(self, *, vid_dim: int, txt_dim: int, emb_dim: int, heads: int, head_dim: int, expand_ratio: int,
                 norm=None, norm_eps: float = 1e-5, ada=None, qk_bias: bool = False, qk_norm=None,
                 mlp_type: str = "swiglu", shared_weights: bool = False, rope_type: Optional[str] = "mmrope3d",
                 rope_dim: int = 128, is_last_layer: bool = False, attention_mode: str = "sdpa", **kwargs):
        super().__init__()
        for what, f in (("norm", norm), ("qk_norm", qk_norm)):
            kind = _factory_kind(f)
            if kind not in (None, "fusedrms", "_norm_layer"):
                raise NotImplementedError(f"seedvr2_b200: {what}={kind!r}; only 'fusedrms' is implemented")
        if emb_dim != 6 * vid_dim:
            raise NotImplementedError("seedvr2_b200: AdaSingle needs emb_dim == 6 * dim (modulation.py:52)")
        if mlp_type != "swiglu":
            raise NotImplementedError("seedvr2_b200: the 3B block uses mlp_type='swiglu'")
        if rope_type not in ("mmrope3d", None):
            raise NotImplementedError(f"seedvr2_b200: rope_type {rope_type!r}")
        if rope_type is not None and rope_dim != 128:
            raise NotImplementedError("seedvr2_b200: rope_dim must be 128 (configs_3b/main.yaml:36)")
        self.norm_eps = float(norm_eps)
        self.is_last_layer = is_last_layer
        self.txt_tail_skipped = bool(is_last_layer)
        self.attn_norm = nn.Module()     # RMSNorm without affine: no state
        self.attn = NaSwinAttentionB200(
            vid_dim=vid_dim, txt_dim=txt_dim, heads=heads, head_dim=head_dim, qk_bias=qk_bias, qk_norm_eps=norm_eps,
            rope_kind=ops.ROPE_3B_MM if rope_type == "mmrope3d" else ops.ROPE_NONE, shared_weights=shared_weights,
            shared_qknorm=shared_weights, window=kwargs.pop("window", None), window_method=kwargs.pop("window_method", None))
        self.mlp_norm = nn.Module()
        self.mlp = _Slot(lambda: _SwiGLU(vid_dim, expand_ratio), shared_weights, vid_only=is_last_layer)
        self.ada = _Slot(lambda: _AdaParams(vid_dim), shared_weights, vid_only=is_last_layer)


class NaMMSRTransformerBlockB200_7B(_BlockBase):
    """SeedVR2-7B block.  Constructor kwargs of src/models/dit_7b/nablocks/mmsr_block.py:161-180."""

    def __init__(self, *, vid_dim: int, txt_dim: int, emb_dim: int, heads: int, head_dim: int, expand_ratio: int,
                 norm=None, norm_eps: float = 1e-5, ada=None, qk_bias: bool = False, qk_rope: bool = True, qk_norm=None,
                 shared_qkv: bool = False, shared_mlp: bool = False, mlp_type: str = "normal", window=None,
                 window_method: Optional[str] = None, attention_mode: str = "sdpa", **kwargs):
        super().__init__()
        for what, f in (("norm", norm), ("qk_norm", qk_norm)):
            kind = _factory_kind(f)
            if kind not in (None, "fusedrms", "_norm_layer"):
                raise NotImplementedError(f"seedvr2_b200: {what}={kind!r}; only 'fusedrms' is implemented")
        if emb_dim != 6 * vid_dim:
            raise NotImplementedError("seedvr2_b200: AdaSingle needs emb_dim == 6 * dim")
        if mlp_type != "normal":
            raise NotImplementedError("seedvr2_b200: the 7B block uses mlp_type='normal' (GELU-tanh)")
        self.norm_eps = float(norm_eps)
        self.attn_norm = nn.Module()
        self.attn = NaSwinAttentionB200(
            vid_dim=vid_dim, txt_dim=txt_dim, heads=heads, head_dim=head_dim, qk_bias=qk_bias, qk_norm_eps=norm_eps,
            rope_kind=ops.ROPE_7B_PIX if qk_rope else ops.ROPE_NONE, shared_weights=shared_qkv,
            shared_qknorm=False,   # 7B never shares norm_q/norm_k (blocks/mmdit_window_block.py:63-64)
            window=window, window_method=window_method)
        self.mlp_norm = nn.Module()
        self.mlp = _Slot(lambda: _GeluMLP(vid_dim, expand_ratio), shared_mlp)
        self.ada = _Slot(lambda: _AdaParams(vid_dim), False)   # 7B ada is never shared (mmdit_window_block.py:211)
