"""Deterministic parity cases shared by the golden generator (run once, in the build container, against the
real reference) and by the tests (run anywhere, without the reference).

Nothing here imports the reference or the product; it only defines shapes, seeds and how synthetic parameters and
inputs are drawn, so that `make_golden.py` and the tests see bit-identical tensors.
"""
import hashlib
import math
from dataclasses import dataclass, field
from typing import Dict, Tuple

import torch

TXT_LEN = 58  # pos_emb.pt is [58, 5120]


@dataclass
class BlockCase:
    name: str
    model: str                      # "3b" | "7b"
    dim: int
    heads: int
    grid: Tuple[int, int, int]      # token grid T, H, W (after 2x2 patching)
    shifted: bool
    shared: bool = False            # 3B only
    is_last: bool = False           # 3B only
    txt_len: int = TXT_LEN
    seed: int = 0
    head_dim: int = 128
    window: Tuple[int, int, int] = (4, 3, 3)
    # |out| stays below 2 at these scales, so one bf16 ulp (<= 0.0078) is well under the 2e-2 north-star tolerance and
    # the tolerance is meaningful; at unit scale the reference's OWN bf16-vs-fp32 gap is 0.06-0.2 (make_golden.py log).
    in_scale: float = 0.25
    emb_scale: float = 0.25

    @property
    def mlp_hidden(self) -> int:
        if self.model == "3b":      # SwiGLU hidden, reference mlp.py:53-54
            h = int(2 * self.dim * 4 / 3)
            return 256 * ((h + 255) // 256)
        return self.dim * 4


BLOCK_CASES = [
    BlockCase("3b_sep_win", "3b", 128, 2, (2, 20, 36), shifted=False, shared=False, seed=11),
    BlockCase("3b_shared_swin", "3b", 128, 2, (2, 20, 36), shifted=True, shared=True, seed=12),
    BlockCase("3b_last_swin", "3b", 128, 2, (1, 32, 32), shifted=True, shared=True, is_last=True, seed=13),
    BlockCase("7b_win", "7b", 128, 3, (3, 20, 36), shifted=False, seed=14),
    BlockCase("7b_swin", "7b", 128, 3, (5, 17, 30), shifted=True, seed=15),
]


def rope_freqs(model: str, head_dim: int = 128) -> torch.Tensor:
    """The persistent buffer ``attn.rope.rope.freqs`` as the reference constructors create it.
    3B: RotaryEmbedding(dim=128//3, freqs_for='lang', theta=10000)  (dit_3b/rope.py:77-81, rope_dim=128 in the YAML)
    7B: RotaryEmbedding(dim=(128//2)//3, freqs_for='pixel', max_freq=256) (dit_7b/rope.py:28-32, mmsr_block.py:64)"""
    if model == "3b":
        dim = head_dim // 3
        return 1.0 / (10000 ** (torch.arange(0, dim, 2)[: dim // 2].float() / dim))
    dim = (head_dim // 2) // 3
    return torch.linspace(1.0, 256 / 2, dim // 2) * math.pi


def block_param_shapes(c: BlockCase) -> Dict[str, Tuple[int, ...]]:
    """State-dict contract of one block; `make_golden.py` asserts it against the reference block."""
    D, I, Hd = c.dim, c.heads * c.head_dim, c.mlp_hidden
    out: Dict[str, Tuple[int, ...]] = {}
    if c.model == "3b":
        br = ["all"] if c.shared else ["vid", "txt"]
        br_tail = ["all"] if c.shared else (["vid"] if c.is_last else ["vid", "txt"])
        for b in br:
            out[f"attn.proj_qkv.{b}.weight"] = (3 * I, D)
            out[f"attn.proj_out.{b}.weight"] = (D, I)
            out[f"attn.proj_out.{b}.bias"] = (D,)
            out[f"attn.norm_q.{b}.weight"] = (c.head_dim,)
            out[f"attn.norm_k.{b}.weight"] = (c.head_dim,)
        for b in br_tail:
            out[f"mlp.{b}.proj_in_gate.weight"] = (Hd, D)
            out[f"mlp.{b}.proj_in.weight"] = (Hd, D)
            out[f"mlp.{b}.proj_out.weight"] = (D, Hd)
            for l in ("attn", "mlp"):
                for g in ("shift", "scale", "gate"):
                    out[f"ada.{b}.{l}_{g}"] = (D,)
        out["attn.rope.rope.freqs"] = (c.head_dim // 3 // 2,)
    else:
        for b in ("vid", "txt"):
            out[f"attn.proj_qkv.{b}.weight"] = (3 * I, D)
            out[f"attn.proj_out.{b}.weight"] = (D, I)
            out[f"attn.proj_out.{b}.bias"] = (D,)
            out[f"attn.norm_q.{b}.weight"] = (c.head_dim,)
            out[f"attn.norm_k.{b}.weight"] = (c.head_dim,)
            out[f"mlp.{b}.proj_in.weight"] = (Hd, D)
            out[f"mlp.{b}.proj_in.bias"] = (Hd,)
            out[f"mlp.{b}.proj_out.weight"] = (D, Hd)
            out[f"mlp.{b}.proj_out.bias"] = (D,)
            for l in ("attn", "mlp"):
                for g in ("shift", "scale", "gate"):
                    out[f"ada.{b}.{l}_{g}"] = (D,)
        out["attn.rope.rope.freqs"] = ((c.head_dim // 2) // 3 // 2,)
    return out


def _gen(seed: int, key: str) -> torch.Generator:
    h = int.from_bytes(hashlib.sha256(f"{seed}:{key}".encode()).digest()[:4], "little")
    return torch.Generator().manual_seed(h)


def draw_param(key: str, shape: Tuple[int, ...], seed: int, model: str) -> torch.Tensor:
    """bf16-representable fp32 values, reference-like magnitudes (Linear ~1/sqrt(in), AdaSingle modulation.py:58-63)."""
    if key.endswith("rope.freqs"):
        return rope_freqs(model)          # checkpoint data, kept fp32
    g = _gen(seed, key)
    r = torch.randn(*shape, generator=g)
    if ".norm_q." in key or ".norm_k." in key or key.endswith("_norm.weight"):
        v = 1.0 + 0.1 * r
    elif key.endswith("_scale"):
        v = 1.0 + r / math.sqrt(shape[0])
    elif key.endswith(("_shift", "_gate")):
        v = r / math.sqrt(shape[0])
    elif key.endswith(".bias"):
        v = 0.02 * r
    else:
        v = r / math.sqrt(shape[-1])
    return v.bfloat16().float()


def block_params(c: BlockCase) -> Dict[str, torch.Tensor]:
    return {k: draw_param(k, s, c.seed, c.model) for k, s in block_param_shapes(c).items()}


def block_inputs(c: BlockCase):
    """vid [L, D], txt [Lt, D], emb [1, 6D] — scaled normal, bf16-representable."""
    T, H, W = c.grid
    vid = (c.in_scale * torch.randn(T * H * W, c.dim, generator=_gen(c.seed, "in.vid"))).bfloat16().float()
    txt = (c.in_scale * torch.randn(c.txt_len, c.dim, generator=_gen(c.seed, "in.txt"))).bfloat16().float()
    emb = (c.emb_scale * torch.randn(1, 6 * c.dim, generator=_gen(c.seed, "in.emb"))).bfloat16().float()
    return vid, txt, emb


def tensor_digest(t: torch.Tensor) -> str:
    return hashlib.sha256(t.detach().contiguous().cpu().numpy().tobytes()).hexdigest()[:16]


# ---- geometry cases: (T, H, W) token grids; the five BASELINE configs + edge cases ------------------------------
GEOMETRY_GRIDS = [
    (1, 32, 32),      # C0 512^2 image
    (1, 128, 128),    # C1 2048^2 image
    (9, 90, 160),     # C2 33 f 1440p
    (31, 68, 120),    # C3 121 f 1080p
    (6, 68, 120),     # C3 per rank @ 8 GPUs
    (5, 68, 120),     # C3 last rank
    (10, 135, 240),   # C4 4K chunk
    (1, 1, 1), (1, 2, 3), (2, 3, 5), (4, 45, 80), (3, 46, 81), (7, 17, 29), (2, 20, 36), (5, 17, 30),
    (33, 9, 7), (1, 200, 3), (1, 3, 200), (30, 15, 27), (31, 15, 27), (8, 64, 64), (1, 67, 119),
]
GEOMETRY_FULL_LIMIT = 4096   # grids with <= this many tokens store full index arrays, larger ones store sha256


# ---- whole-NaDiT cases (the glue either side of the block loop) --------------------------------
@dataclass
class StackCase:
    name: str
    model: str                      # "3b" | "7b"
    dim: int
    heads: int
    num_layers: int
    mm_layers: int                  # 3B: layers < mm_layers use separate vid/txt weights (dit_3b/nadit.py:149-151)
    latent_grid: Tuple[int, int, int]   # T, 2H, 2W (before the (1,2,2) patching)
    txt_in_dim: int = 160
    txt_len: int = TXT_LEN
    timestep: float = 1000.0        # one-step trailing schedule, T=1000 (configs_3b/main.yaml:68)
    seed: int = 0
    in_scale: float = 1.0           # latents + noise are unit normal (infer.py:56,341)
    txt_scale: float = 0.25


STACK_CASES = [
    # 2 separate + 2 shared layers, last layer vid_only, win/swin alternating, vid_out_norm + vid_out_ada (quirk Q1)
    StackCase("3b_4layer", "3b", 256, 2, 4, 2, (2, 40, 72), seed=21),
    # 7B: separate weights everywhere, GELU MLP with biases, pixel RoPE on video only, no output norm
    StackCase("7b_4layer", "7b", 256, 2, 4, 0, (3, 34, 60), seed=22),
]


def stack_kwargs(c: StackCase) -> Dict:
    """Constructor arguments of the reference NaDiT (configs_3b/main.yaml:11-36, configs_7b/main.yaml:11-33), narrowed."""
    methods = ["720pwin_by_size_bysize", "720pswin_by_size_bysize"] * (c.num_layers // 2)
    common = dict(vid_in_channels=33, vid_out_channels=16, vid_dim=c.dim, txt_in_dim=c.txt_in_dim, txt_dim=c.dim,
                  emb_dim=6 * c.dim, heads=c.heads, head_dim=128, expand_ratio=4, norm="fusedrms", norm_eps=1e-5,
                  ada="single", qk_bias=False, qk_norm="fusedrms", patch_size=[1, 2, 2], num_layers=c.num_layers,
                  block_type=["mmdit_sr"] * c.num_layers, window=[(4, 3, 3)] * c.num_layers, window_method=methods)
    if c.model == "3b":
        return dict(common, mm_layers=c.mm_layers, mlp_type="swiglu", patch_type="v1", rope_type="mmrope3d",
                    rope_dim=128, msa_type=None, mca_type=None, vid_out_norm="fusedrms")
    return dict(common, qk_rope=True, shared_mlp=False, shared_qkv=False, mlp_type="normal")


def stack_params(c: StackCase, shapes: Dict[str, Tuple[int, ...]]) -> Dict[str, torch.Tensor]:
    """Deterministic parameters for every state-dict entry; `shapes` comes from the module being filled (the golden
    generator asserts the reference NaDiT and NaDiTStack expose the same names and shapes)."""
    return {k: draw_param(k, s, c.seed, c.model) for k, s in shapes.items()}


def stack_inputs(c: StackCase):
    """vid [T*2H*2W, 33] (16 noise + 16 condition + 1 mask channels, infer.py:56,341), txt [Lt, txt_in_dim]."""
    T, H2, W2 = c.latent_grid
    vid = (c.in_scale * torch.randn(T * H2 * W2, 33, generator=_gen(c.seed, "stack.vid"))).bfloat16().float()
    txt = (c.txt_scale * torch.randn(c.txt_len, c.txt_in_dim, generator=_gen(c.seed, "stack.txt"))).bfloat16().float()
    return vid, txt


def psnr(a: torch.Tensor, ref: torch.Tensor) -> float:
    """PSNR in dB of `a` against `ref`, peak = the reference's value range (the output is a latent, not [0,1] pixels)."""
    ref = ref.double()
    mse = float(((a.double() - ref) ** 2).mean())
    peak = float(ref.max() - ref.min())
    return float("inf") if mse == 0 else 10.0 * math.log10(peak * peak / mse)


# ---- one-step SR call either side of the DiT forward -----------------------------------------------------
@dataclass(frozen=True)
class SamplerCase:
    name: str
    grid: Tuple[int, int, int]          # latent grid (t, h, w) before patching
    dtype: str                          # "bf16" | "fp32": the pipeline's compute dtype (sampler state and latents)
    latent_noise_scale: float
    seed: int
    channels: int = 16


SAMPLER_CASES = [
    SamplerCase("bf16_plain", (3, 6, 10), "bf16", 0.0, 31),
    SamplerCase("bf16_image", (1, 8, 8), "bf16", 0.0, 32),
    SamplerCase("bf16_latent_noise", (3, 6, 10), "bf16", 0.25, 33),
    SamplerCase("bf16_latent_noise_image", (1, 12, 16), "bf16", 1.0, 34),
    SamplerCase("fp32_latent_noise", (2, 4, 6), "fp32", 0.5, 35),
]


def sampler_inputs(c: SamplerCase):
    """latent, base noise and augmentation noise as generation_phases.py:676-679 builds them, plus the fixed matrix of the
    stand-in DiT.  Values scaled so that x_t - pred exercises bf16 rounding."""
    dt = torch.bfloat16 if c.dtype == "bf16" else torch.float32
    g = torch.Generator().manual_seed(c.seed)
    shape = (*c.grid, c.channels)
    latent = (torch.randn(shape, generator=g) * 1.3).to(dt)
    base_noise = torch.randn(shape, generator=g).to(dt)
    aug_noise = base_noise * 0.1 + torch.randn(shape, generator=g).to(dt) * 0.05
    mix = (torch.randn(2 * c.channels + 1, c.channels, generator=g) * 0.2).to(dt)
    return latent, base_noise, aug_noise, mix


def standin_dit(mix: torch.Tensor):
    """A deterministic stand-in for the DiT: pred = vid @ mix (in the compute dtype), so the golden pins the channel order
    of the concat and the endpoint arithmetic without a model."""
    return lambda vid, timestep: (vid.float() @ mix.float()).to(vid.dtype)
