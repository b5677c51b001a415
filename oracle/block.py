"""Functional restatement of one NaMMSR transformer block.  TEST INFRASTRUCTURE ONLY.

Reference: src/models/dit_3b/nablocks/mmsr_block.py:84-128 (3B), src/models/dit_7b/nablocks/mmsr_block.py:213-250 (7B),
attention src/models/dit_3b/nablocks/attention/mmattn.py:161-271 / dit_7b/nablocks/mmsr_block.py:68-157,
adaLN src/models/dit_3b/modulation.py:65-118, RMSNorm normalization.py:88-109, MLP mlp.py:28-62,
varlen attention attention.py:24-61 (per-window SDPA), text pooling na.py:396-417.

Written with plain torch ops in the reference's op ORDER so that running it in bf16 reproduces the reference's
rounding points and running it in fp32 gives the ground truth.  Device-agnostic: every tensor follows
``vid.device``, so the same restatement is the checker on the CPU (golden-pinned, tests/test_oracle_golden.py) and — with
``vid`` on ``cuda`` — the reference's own bf16 PyTorch path on a GPU (per-window ``F.scaled_dot_product_attention``,
attention.py:24-61) for the full-width parity tests and bench.py's GPU-vs-GPU baseline leg.  ``params`` uses the reference's
state-dict key names relative to the block (e.g. ``attn.proj_qkv.vid.weight``).
"""
from dataclasses import dataclass
from typing import Dict, Tuple

import torch
import torch.nn.functional as F

from . import geometry, rope


@dataclass
class BlockCfg:
    model: str = "3b"            # "3b" | "7b"
    dim: int = 2560
    heads: int = 20
    head_dim: int = 128
    shared: bool = False         # 3B layers >= mm_layers share vid/txt weights (".all")
    is_last: bool = False        # 3B last layer: mlp/ada/mlp_norm are vid_only
    shifted: bool = False        # window_method 720pswin (odd layers)
    window: Tuple[int, int, int] = (4, 3, 3)
    eps: float = 1e-5


def _branch(cfg: BlockCfg, module: str, branch: str) -> str:
    """Sub-module name for a branch, mirroring MMModule's .vid/.txt/.all (src/models/dit_3b/mm.py:35-74)."""
    if cfg.model == "3b":
        return "all" if cfg.shared else branch
    return branch  # 7B: shared_qkv/shared_mlp False everywhere (configs_7b/main.yaml:28-29)


def rmsnorm(x: torch.Tensor, eps: float, weight: torch.Tensor = None) -> torch.Tensor:
    var = x.pow(2).mean(dim=-1, keepdim=True)
    out = x / torch.sqrt(var + eps)
    return out * weight if weight is not None else out


def ada_triplet(emb: torch.Tensor, layer: int) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """emb [1, 6D] viewed 'b (d l g)', l=2 (attn,mlp), g=3 (shift,scale,gate).  modulation.py:75-76."""
    e = emb.reshape(emb.shape[0], -1, 2, 3)[:, :, layer, :]
    return e[..., 0], e[..., 1], e[..., 2]


def ada_in(h, emb, layer, shift_b, scale_b):
    shift_a, scale_a, _ = ada_triplet(emb, layer)
    return h * (scale_a + scale_b) + (shift_a + shift_b)


def ada_out(h, emb, layer, gate_b):
    _, _, gate_a = ada_triplet(emb, layer)
    return h * (gate_a + gate_b)


def mlp(cfg: BlockCfg, p: Dict[str, torch.Tensor], pre: str, x: torch.Tensor) -> torch.Tensor:
    if cfg.model == "3b":   # SwiGLU, no bias.  mlp.py:46-62
        g = F.linear(x, p[pre + ".proj_in_gate.weight"])
        u = F.linear(x, p[pre + ".proj_in.weight"])
        return F.linear(F.silu(g) * u, p[pre + ".proj_out.weight"])
    hmid = F.gelu(F.linear(x, p[pre + ".proj_in.weight"], p[pre + ".proj_in.bias"]), approximate="tanh")
    return F.linear(hmid, p[pre + ".proj_out.weight"], p[pre + ".proj_out.bias"])


def window_attention(cfg: BlockCfg, p: Dict[str, torch.Tensor], vid: torch.Tensor, txt: torch.Tensor,
                     grid: Tuple[int, int, int], return_internals: bool = False):
    """vid [L, D], txt [Lt, D] (already normalised+modulated) -> attention output before proj_out, [L, H*d], [Lt, H*d]."""
    T, H, W = grid
    Hh, d = cfg.heads, cfg.head_dim
    Lt = txt.shape[0]
    bv, bt = _branch(cfg, "attn", "vid"), _branch(cfg, "attn", "txt")
    vqkv = F.linear(vid, p[f"attn.proj_qkv.{bv}.weight"])
    tqkv = F.linear(txt, p[f"attn.proj_qkv.{bt}.weight"])
    tgt, src, wshape, cu = geometry.window_index(T, H, W, cfg.window, cfg.shifted, Lt)
    dev = vid.device
    tgt_t, src_t = torch.from_numpy(tgt).to(dev), torch.from_numpy(src).to(dev)
    vqkv = vqkv.index_select(0, tgt_t).reshape(-1, 3, Hh, d)
    tqkv = tqkv.reshape(-1, 3, Hh, d)
    vq, vk, vv = vqkv.unbind(1)
    tq, tk, tv = tqkv.unbind(1)
    # 7B keeps norm_q/norm_k un-shared even though it is an MMModule (mmdit_window_block.py:63-64)
    nq_v, nq_t = (bv, bt) if cfg.model == "3b" else ("vid", "txt")
    vq = rmsnorm(vq, cfg.eps, p[f"attn.norm_q.{nq_v}.weight"])
    tq = rmsnorm(tq, cfg.eps, p[f"attn.norm_q.{nq_t}.weight"])
    vk = rmsnorm(vk, cfg.eps, p[f"attn.norm_k.{nq_v}.weight"])
    tk = rmsnorm(tk, cfg.eps, p[f"attn.norm_k.{nq_t}.weight"])
    ws = [tuple(int(v) for v in r) for r in wshape]
    freqs = p["attn.rope.rope.freqs"]
    if cfg.model == "3b":
        va, ta = rope.angles_3b(ws, Lt, freqs)
        va, ta = va.to(dev), ta.to(dev)
        vq, vk = rope.rotate(vq, va), rope.rotate(vk, va)
        tq, tk = rope.rotate(tq, ta), rope.rotate(tk, ta)
    else:
        va = rope.angles_7b(ws, freqs).to(dev).to(vq.dtype)      # dit_7b/rope.py:85
        vq, vk = rope.rotate(vq, va), rope.rotate(vk, va)
    vout = torch.empty(vq.shape[0], Hh, d, dtype=vq.dtype, device=dev)
    tacc = []
    off = 0
    for k, (f, h, w) in enumerate(ws):
        n = f * h * w
        q = torch.cat([vq[off:off + n], tq]).permute(1, 0, 2).unsqueeze(0)
        kk = torch.cat([vk[off:off + n], tk]).permute(1, 0, 2).unsqueeze(0)
        v = torch.cat([vv[off:off + n], tv]).permute(1, 0, 2).unsqueeze(0)
        o = F.scaled_dot_product_attention(q, kk, v).squeeze(0).permute(1, 0, 2)
        vout[off:off + n] = o[:n]
        tacc.append(o[n:])
        off += n
    tout = torch.stack(tacc, dim=1).mean(1)              # na.py:414 mean over windows
    vout = vout.reshape(-1, Hh * d).index_select(0, src_t)
    tout = tout.reshape(-1, Hh * d)
    if return_internals:
        return vout, tout, dict(vq=vq, vk=vk, vv=vv, tq=tq, tk=tk, tv=tv, tgt=tgt, src=src, wshape=wshape, cu=cu)
    return vout, tout


def block_forward(cfg: BlockCfg, p: Dict[str, torch.Tensor], vid: torch.Tensor, txt: torch.Tensor,
                  grid: Tuple[int, int, int], emb: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """One block.  vid [L, D], txt [Lt, D], emb [1, 6D] -> (vid, txt).  Appendix A.4 of SURVEY.md."""
    bv, bt = _branch(cfg, "x", "vid"), _branch(cfg, "x", "txt")
    av, at_ = (bv, bt) if cfg.model == "3b" else ("vid", "txt")   # 7B ada is never shared (mmdit_window_block.py:211)
    a_v = ada_in(rmsnorm(vid, cfg.eps), emb, 0, p[f"ada.{av}.attn_shift"], p[f"ada.{av}.attn_scale"])
    # 3B last layer: ``ada`` / ``mlp_norm`` / ``mlp`` are vid_only MMModules (mmsr_block.py:73-81), so the txt branch
    # passes through them untouched (mm.py:69-73) — in BOTH the attn and the mlp half.
    txt_mod = not (cfg.model == "3b" and cfg.is_last)
    a_t = rmsnorm(txt, cfg.eps)
    if txt_mod:
        a_t = ada_in(a_t, emb, 0, p[f"ada.{at_}.attn_shift"], p[f"ada.{at_}.attn_scale"])
    o_v, o_t = window_attention(cfg, p, a_v, a_t, grid)
    o_v = F.linear(o_v, p[f"attn.proj_out.{bv}.weight"], p[f"attn.proj_out.{bv}.bias"])
    o_t = F.linear(o_t, p[f"attn.proj_out.{bt}.weight"], p[f"attn.proj_out.{bt}.bias"])
    vid1 = ada_out(o_v, emb, 0, p[f"ada.{av}.attn_gate"]) + vid
    txt1 = (ada_out(o_t, emb, 0, p[f"ada.{at_}.attn_gate"]) if txt_mod else o_t) + txt

    m_v = ada_in(rmsnorm(vid1, cfg.eps), emb, 1, p[f"ada.{av}.mlp_shift"], p[f"ada.{av}.mlp_scale"])
    m_v = mlp(cfg, p, f"mlp.{bv}", m_v)
    vid2 = ada_out(m_v, emb, 1, p[f"ada.{av}.mlp_gate"]) + vid1
    if not txt_mod:
        txt2 = txt1 + txt1      # norm/ada/mlp all skipped, then the residual add (mmsr_block.py:126)
    else:
        m_t = ada_in(rmsnorm(txt1, cfg.eps), emb, 1, p[f"ada.{at_}.mlp_shift"], p[f"ada.{at_}.mlp_scale"])
        m_t = mlp(cfg, p, f"mlp.{bt}", m_t)
        txt2 = ada_out(m_t, emb, 1, p[f"ada.{at_}.mlp_gate"]) + txt1
    return vid2, txt2
