"""Functional restatement of the whole one-step NaDiT forward (the glue either side of the block loop + the blocks).
TEST INFRASTRUCTURE ONLY — imported by tests/ and tests/golden/make_golden.py, never by the product.

Reference: src/models/dit_3b/nadit.py:171-248 (3B), src/models/dit_7b/nadit.py:152-191 (7B); text projection
nadit.py:117-121; patchify + Linear src/models/dit_3b/patch/patch_v1.py:76-127; TimeEmbedding
src/models/dit_3b/embedding.py:25-62 (+ diffusers get_timestep_embedding, flip_sin_to_cos=False, shift 0); output
norm + modulation dit_3b/nadit.py:233-244 including cache-key quirk Q1: ``vid_out_ada`` is handed the
ATTN-layer (shift, scale) of ``emb`` that the blocks left in the cache under "emb_repeat_0_vid".

Pinned against the real reference NaDiT by tests/golden/stack_*.pt (tests/test_oracle_golden.py).  ``params`` uses the
reference NaDiT's state-dict key names (``vid_in.proj.weight``, ``blocks.0.attn.proj_qkv.vid.weight`` ...).
"""
import math
from typing import Dict, Tuple

import torch
import torch.nn.functional as F

from . import block as oblock


def timestep_embedding(timestep: torch.Tensor, dim: int = 256) -> torch.Tensor:
    half = dim // 2
    exponent = -math.log(10000) * torch.arange(half, dtype=torch.float32) / half
    ang = timestep[:, None].float() * torch.exp(exponent)[None, :]
    return torch.cat([torch.sin(ang), torch.cos(ang)], dim=-1)


def nadit_forward(model: str, p: Dict[str, torch.Tensor], vid: torch.Tensor, txt: torch.Tensor,
                  latent_grid: Tuple[int, int, int], timestep: float, *, dim: int, heads: int, num_layers: int,
                  mm_layers: int = 0, window=(4, 3, 3)) -> torch.Tensor:
    """vid [T*2H*2W, 33], txt [Lt, txt_in_dim] -> predicted latent [T*2H*2W, 16]; dtype follows `vid`."""
    T, H2, W2 = latent_grid
    H, W = H2 // 2, W2 // 2
    dt = vid.dtype
    txt = F.linear(txt, p["txt_in.weight"], p["txt_in.bias"])
    # "(T t)(H h)(W w) c -> T H W (t h w c)", patch (1,2,2)    patch_v1.py:91
    C = vid.shape[-1]
    x = vid.view(T, H, 2, W, 2, C).permute(0, 1, 3, 2, 4, 5).reshape(T * H * W, 4 * C)
    x = F.linear(x, p["vid_in.proj.weight"], p["vid_in.proj.bias"])
    e = timestep_embedding(torch.tensor([float(timestep)])).to(dt)
    e = F.silu(F.linear(e, p["emb_in.proj_in.weight"], p["emb_in.proj_in.bias"]))
    e = F.silu(F.linear(e, p["emb_in.proj_hid.weight"], p["emb_in.proj_hid.bias"]))
    emb = F.linear(e, p["emb_in.proj_out.weight"], p["emb_in.proj_out.bias"])
    for i in range(num_layers):
        cfg = oblock.BlockCfg(model=model, dim=dim, heads=heads, shared=(model == "3b" and i >= mm_layers),
                              is_last=(model == "3b" and i == num_layers - 1), shifted=bool(i % 2), window=window)
        pre = f"blocks.{i}."
        bp = {k[len(pre):]: v for k, v in p.items() if k.startswith(pre)}
        x, txt = oblock.block_forward(cfg, bp, x, txt, (T, H, W), emb)
    if "vid_out_norm.weight" in p:
        x = oblock.rmsnorm(x, 1e-5, p["vid_out_norm.weight"])
        shift_a, scale_a, _ = oblock.ada_triplet(emb, 0)          # quirk Q1: the attn-layer pair, not an "out" pair
        x = x * (scale_a + p["vid_out_ada.out_scale"]) + (shift_a + p["vid_out_ada.out_shift"])
    x = F.linear(x, p["vid_out.proj.weight"], p["vid_out.proj.bias"])
    # "T H W (t h w c) -> (T t)(H h)(W w) c"    patch_v1.py:121
    co = x.shape[-1] // 4
    return x.view(T, H, W, 2, 2, co).permute(0, 1, 3, 2, 4, 5).reshape(T * H2 * W2, co)
