"""Synthetic full-width block parameters/inputs for timing the oracle on host cores.  TEST INFRASTRUCTURE ONLY
(used by bench.py's ``cpu_baseline`` leg and ``--impl reference`` arm, never by the product).

Shapes and names follow the reference's state dict for one block: 3B = configs_3b/main.yaml:11-36
(dim 2560, 20 heads, SwiGLU hidden 6912), 7B = configs_7b/main.yaml:11-33 (dim 3072, 24 heads, GELU hidden 12288).
Magnitudes follow the reference constructors (nn.Linear default init ~ U(-1/sqrt(in), 1/sqrt(in)), AdaSingle
``randn/sqrt(D) (+1)`` modulation.py:58-63)."""
import math
from typing import Dict, Tuple

import torch

from .block import BlockCfg

MODELS = {"3b": dict(dim=2560, heads=20, mlp=6912), "7b": dict(dim=3072, heads=24, mlp=12288)}


def block_cfg(model: str, shared: bool = False, shifted: bool = False) -> BlockCfg:
    m = MODELS[model]
    return BlockCfg(model=model, dim=m["dim"], heads=m["heads"], shared=shared and model == "3b", shifted=shifted)


def block_params(cfg: BlockCfg, seed: int = 0, dtype=torch.float32) -> Dict[str, torch.Tensor]:
    g = torch.Generator().manual_seed(seed)
    D, inner, hid = cfg.dim, cfg.heads * cfg.head_dim, MODELS[cfg.model]["mlp"]

    def lin(o, i):
        return ((torch.rand(o, i, generator=g) * 2 - 1) / math.sqrt(i)).to(dtype)

    def vec(n, base=0.0, s=None):
        return (base + torch.randn(n, generator=g) * (s if s is not None else 1 / math.sqrt(n))).to(dtype)

    p: Dict[str, torch.Tensor] = {}
    branches = ["all"] if cfg.shared else ["vid", "txt"]
    for b in branches:
        p[f"attn.proj_qkv.{b}.weight"] = lin(3 * inner, D)
        p[f"attn.proj_out.{b}.weight"] = lin(D, inner)
        p[f"attn.proj_out.{b}.bias"] = vec(D, s=0.02)
        if cfg.model == "3b":
            p[f"mlp.{b}.proj_in_gate.weight"] = lin(hid, D)
            p[f"mlp.{b}.proj_in.weight"] = lin(hid, D)
            p[f"mlp.{b}.proj_out.weight"] = lin(D, hid)
        else:
            p[f"mlp.{b}.proj_in.weight"] = lin(hid, D)
            p[f"mlp.{b}.proj_in.bias"] = vec(hid, s=0.02)
            p[f"mlp.{b}.proj_out.weight"] = lin(D, hid)
            p[f"mlp.{b}.proj_out.bias"] = vec(D, s=0.02)
    for b in (branches if cfg.model == "3b" else ["vid", "txt"]):
        p[f"attn.norm_q.{b}.weight"] = vec(cfg.head_dim, 1.0, 0.1)
        p[f"attn.norm_k.{b}.weight"] = vec(cfg.head_dim, 1.0, 0.1)
        for l in ("attn", "mlp"):
            p[f"ada.{b}.{l}_shift"] = vec(D)
            p[f"ada.{b}.{l}_scale"] = vec(D, 1.0)
            p[f"ada.{b}.{l}_gate"] = vec(D)
    if cfg.model == "3b":
        dim = cfg.head_dim // 3
        p["attn.rope.rope.freqs"] = 1.0 / (10000 ** (torch.arange(0, dim, 2)[: dim // 2].float() / dim))
    else:
        dim = (cfg.head_dim // 2) // 3
        p["attn.rope.rope.freqs"] = torch.linspace(1.0, 256 / 2, dim // 2, dtype=torch.float32) * math.pi
    return p


def block_inputs(cfg: BlockCfg, grid: Tuple[int, int, int], txt_len: int = 58, seed: int = 1, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)
    T, H, W = grid
    vid = torch.randn(T * H * W, cfg.dim, generator=g).to(dtype)
    txt = torch.randn(txt_len, cfg.dim, generator=g).to(dtype)
    emb = (0.5 * torch.randn(1, 6 * cfg.dim, generator=g)).to(dtype)
    return vid, txt, emb
