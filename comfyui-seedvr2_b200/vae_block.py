"""The video VAE's modules on the B200.

* ``ResnetBlock3DB200`` mirrors ``ResnetBlock3D`` (src/models/video_vae_v3/modules/attn_video_vae.py:253-361) as the VAE builds it (no
  time embedding, no sampler inside the block, SiLU, dropout 0): per-frame GroupNorm + SiLU as one fused pass, the two causal 3-D
  convolutions as tcgen05 implicit GEMMs, the residual add and the next GroupNorm's statistics in the convolutions' epilogues.
* ``Upsample3DB200`` / ``Downsample3DB200`` mirror the samplers (:58-175, :176-246): 1 x 1 x 1 expansion + pixel / time shuffle +
  causal convolution; zero row / column + strided causal convolution.
* ``make_causal_norm_wrapper`` / ``patch_causal_norm_wrapper`` put the stand-alone GroupNorms (conv_norm_out) on the same kernel.
* ``install_vae_blocks(model)`` swaps all of it — and the mid block's attention layer (``vae_attn.AttentionB200``) — into a built VAE in place.
Same forward signatures and memory bookkeeping (through the convolutions) as the modules they replace.  CUDA only — there is no other
path."""
from typing import Optional

import torch
from torch import nn

from . import _lib, ops
from .vae_conv import _DTYPE_CODE, CausalConv3dB200, GroupStats, to_channels_last


def groupnorm_silu(x: torch.Tensor, weight: Optional[torch.Tensor], bias: Optional[torch.Tensor], groups: int, eps: float,
                   silu: bool = True, out: Optional[torch.Tensor] = None, stats: Optional[GroupStats] = None) -> torch.Tensor:
    """x [T, H, W, C] (or [T, HW, C]) channels-last: GroupNorm over each frame's (C / groups) x H x W values, then SiLU; rounded to
    the tensor dtype after each of the two, like the reference's two tensor ops (causal_inflation_lib.py:366-409 + nn.SiLU).
    ``stats``: the partials the producing convolution gathered for x (32 groups) — the statistics pass is skipped."""
    code = _DTYPE_CODE.get(x.dtype)
    if code is None:
        raise NotImplementedError(f"seedvr2_b200.vae_block: fp16 or bf16 activations only (got {x.dtype})")
    ops._need(x, "x", x.dtype)
    T, C = x.shape[0], x.shape[-1]
    hw = x.numel() // max(T * C, 1)
    for name, t in (("weight", weight), ("bias", bias)):
        if t is not None:
            ops._need(t, name, x.dtype)
            if t.numel() != C:
                raise RuntimeError(f"seedvr2_b200.vae_block: `{name}` must have {C} elements")
    y = torch.empty_like(x) if out is None else ops._need(out, "out", x.dtype)
    if T == 0 or hw == 0:
        return y
    lib = _lib.lib()
    if stats is not None and groups == 32:
        ops._call("svr2_groupnorm_silu_fwd", 1, x.device, lib.svr2_groupnorm_silu_fwd, x.data_ptr(), ops._ptr(weight), ops._ptr(bias),
                  y.data_ptr(), stats.partials.data_ptr(), int(stats.slots), T, hw, C, 32, float(eps), int(silu), code, ops._stream(x))
        return y
    ws = torch.empty(int(lib.svr2_groupnorm_workspace_bytes(T, groups)), dtype=torch.uint8, device=x.device)
    ops._call("svr2_groupnorm_silu_fwd", 2, x.device, lib.svr2_groupnorm_silu_fwd, x.data_ptr(), ops._ptr(weight), ops._ptr(bias),
              y.data_ptr(), ws.data_ptr(), 0, T, hw, C, int(groups), float(eps), int(silu), code, ops._stream(x))
    return y


class ResnetBlock3DB200(nn.Module):
    """Drop-in for the VAE's ``ResnetBlock3D``; see the module docstring."""

    def __init__(self, norm1: nn.GroupNorm, conv1: nn.Module, norm2: nn.GroupNorm, conv2: nn.Module,
                 conv_shortcut: Optional[nn.Module] = None, output_scale_factor: float = 1.0):
        super().__init__()
        self.groups1, self.eps1 = norm1.num_groups, norm1.eps
        self.groups2, self.eps2 = norm2.num_groups, norm2.eps
        for name, n in (("norm1", norm1), ("norm2", norm2)):
            self.register_buffer(name + "_weight", None if n.weight is None else n.weight.detach().clone(), persistent=False)
            self.register_buffer(name + "_bias", None if n.bias is None else n.bias.detach().clone(), persistent=False)
        self.conv1 = CausalConv3dB200.from_reference(conv1)
        self.conv2 = CausalConv3dB200.from_reference(conv2)
        self.conv_shortcut = None if conv_shortcut is None else CausalConv3dB200.from_reference(conv_shortcut)
        self.in_channels, self.out_channels = self.conv1.in_channels, self.conv2.out_channels
        self.output_scale_factor = float(output_scale_factor)

    @classmethod
    def from_reference(cls, block: nn.Module) -> "ResnetBlock3DB200":
        """From a built ``ResnetBlock3D``.  Raises NotImplementedError for configurations the VAE does not use."""
        if getattr(block, "time_emb_proj", None) is not None or getattr(block, "upsample", None) is not None or \
                getattr(block, "downsample", None) is not None:
            raise NotImplementedError("seedvr2_b200.vae_block: blocks with a time embedding or an internal sampler are not mirrored")
        if not isinstance(block.nonlinearity, nn.SiLU):
            raise NotImplementedError("seedvr2_b200.vae_block: SiLU blocks only")
        if not isinstance(block.norm1, nn.GroupNorm) or not isinstance(block.norm2, nn.GroupNorm):
            raise NotImplementedError("seedvr2_b200.vae_block: GroupNorm blocks only")
        if block.conv1.in_channels % 64 or block.conv1.out_channels % 128 or block.conv2.out_channels % 128:
            raise NotImplementedError("seedvr2_b200.vae_block: block widths outside the kernel's channel granularity (64 in, 128 out)")
        drop = getattr(block, "dropout", None)
        if drop is not None and getattr(drop, "p", 0.0) != 0.0 and block.training:
            raise NotImplementedError("seedvr2_b200.vae_block: inference only (dropout)")
        return cls(block.norm1, block.conv1, block.norm2, block.conv2, getattr(block, "conv_shortcut", None),
                   getattr(block, "output_scale_factor", 1.0))

    def _norm_params(self, dtype: torch.dtype):
        ps = (self.norm1_weight, self.norm1_bias, self.norm2_weight, self.norm2_bias)
        if all(p is None or p.dtype == dtype for p in ps):
            return ps
        cache = self.__dict__.setdefault("_norm_cast", {})
        hit = cache.get(dtype)
        if hit is None or any(a is not None and a.device != b.device for a, b in zip(hit, ps)):
            hit = tuple(None if p is None else p.to(dtype) for p in ps)
            cache[dtype] = hit
        return hit

    def forward(self, input_tensor: torch.Tensor, temb=None, memory_state="DISABLED", **kwargs) -> torch.Tensor:
        if temb is not None:
            raise NotImplementedError("seedvr2_b200.vae_block: the VAE calls its residual blocks with temb=None")
        x = to_channels_last(input_tensor)                           # [B, T, H, W, C]: no copy for channels_last_3d input
        cd = self.conv1._compute_dtype()                             # the autocast dtype where the pipeline autocasts the VAE
        if x.dtype != cd:
            x = x.to(cd)
        n1w, n1b, n2w, n2b = self._norm_params(cd)
        B = x.shape[0]
        # statistics of the input, if the block that produced it gathered them in its last convolution's epilogue
        st_in = getattr(input_tensor, "_svr2_gn_stats", None)
        if st_in is not None and not (B == 1 and self.groups1 == 32 and st_in.valid_for(x)):
            st_in = None
        h = torch.empty_like(x)
        for b in range(B):
            groupnorm_silu(x[b], n1w, n1b, self.groups1, self.eps1, out=h[b], stats=st_in)
        h, st_h = self.conv1.forward_cl(h, memory_state, want_stats=self.groups2 == 32)
        for b in range(B):
            groupnorm_silu(h[b], n2w, n2b, self.groups2, self.eps2, out=h[b], stats=st_h)      # in place
        skip = x if self.conv_shortcut is None else self.conv_shortcut.forward_cl(x, memory_state)
        out, st_out = self.conv2.forward_cl(h, memory_state, resid=skip, want_stats=True)     # (input_tensor + hidden_states), :361
        if self.output_scale_factor != 1.0:
            out, st_out = out / self.output_scale_factor, None
        ret = out.permute(0, 4, 1, 2, 3)
        if st_out is not None:
            ret._svr2_gn_stats = st_out          # for the next block's norm1 (the VAE chains its blocks through nn.Identity)
        return ret


def pixel_shuffle3d(x: torch.Tensor, channels: int, spatial: int, temporal: int, drop_second_frame: bool = False) -> torch.Tensor:
    """x [T, H, W, spatial^2 * temporal * channels] channels-last -> [T * temporal (- 1), H * spatial, W * spatial, channels]:
    the up-sampler's rearrange "b (x y z c) f h w -> b c (f z) (h x) (w y)" (+ remove_head) as one copy kernel."""
    ops._need(x, "x", x.dtype)
    T, H, W, cr = x.shape
    if x.element_size() != 2 or cr != spatial * spatial * temporal * channels:
        raise RuntimeError("seedvr2_b200.vae_block: pixel_shuffle3d expects 2-byte elements and spatial^2 * temporal * channels inputs")
    out = torch.empty(T * temporal - int(drop_second_frame), H * spatial, W * spatial, channels, dtype=x.dtype, device=x.device)
    ops._call("svr2_pixel_shuffle3d_fwd", 1, x.device, _lib.lib().svr2_pixel_shuffle3d_fwd, x.data_ptr(), out.data_ptr(), T, H, W, channels,
              spatial, temporal, int(drop_second_frame), ops._stream(x))
    return out


class Upsample3DB200(nn.Module):
    """Drop-in for ``Upsample3D`` (attn_video_vae.py:58-175) as the VAE builds it (no norm, no transposed convolution, MAGViT-v2
    expansion): 1 x 1 x 1 channel expansion, pixel / time shuffle with the duplicated first frame removed on the first temporal
    slice, causal 3 x 3 x 3 convolution — with ``slicing`` fed as two temporal pieces, like the reference."""

    def __init__(self, upscale_conv: nn.Module, conv: nn.Module, channels: int, spatial_ratio: int, temporal_ratio: int, slicing: bool):
        super().__init__()
        self.upscale_conv = CausalConv3dB200.from_reference(upscale_conv)
        self.conv = CausalConv3dB200.from_reference(conv)
        self.channels, self.spatial_ratio, self.temporal_ratio, self.slicing = channels, spatial_ratio, temporal_ratio, slicing
        self.temporal_up = temporal_ratio == 2

    @classmethod
    def from_reference(cls, up: nn.Module) -> "Upsample3DB200":
        if getattr(up, "norm", None) is not None or getattr(up, "use_conv_transpose", False) or getattr(up, "interpolate", False) \
                or not getattr(up, "use_conv", True):
            raise NotImplementedError("seedvr2_b200.vae_block: up-sampler configuration outside the mirror")
        conv = up.conv if getattr(up, "name", "conv") == "conv" else up.Conv2d_0
        if not (_conv_supported(conv) and _conv_supported(up.upscale_conv)) or up.channels % 8:
            raise NotImplementedError("seedvr2_b200.vae_block: up-sampler convolutions outside the kernel's shapes")
        return cls(up.upscale_conv, conv, up.channels, up.spatial_ratio, up.temporal_ratio, bool(getattr(up, "slicing", False)))

    def forward(self, hidden_states: torch.Tensor, output_size=None, memory_state="DISABLED", **kwargs):
        from .vae_conv import _state_name
        x = to_channels_last(hidden_states)                          # [B, T, H, W, C]
        T = x.shape[1]
        pieces = [x[:, :T // 2], x[:, T // 2:]] if self.slicing else [x]          # :126-133
        first = _state_name(memory_state) != "ACTIVE"
        ups = []
        for i, piece in enumerate(pieces):
            if piece.shape[1] == 0:
                continue
            t = self.upscale_conv.forward_cl(piece.contiguous(), "DISABLED")       # [B, t, H, W, r C]
            drop = self.temporal_up and first and i == 0                           # remove_head, :151-153
            ups.append(torch.stack([pixel_shuffle3d(t[b], self.channels, self.spatial_ratio, self.temporal_ratio, drop)
                                    for b in range(t.shape[0])], 0) if t.shape[0] > 1 else
                       pixel_shuffle3d(t[0], self.channels, self.spatial_ratio, self.temporal_ratio, drop)[None])
        ys = self.conv._forward_pieces(ups, memory_state)
        y = ys[0] if len(ys) == 1 else torch.cat(ys, dim=1)
        return y.permute(0, 4, 1, 2, 3)


class Downsample3DB200(nn.Module):
    """Drop-in for ``Downsample3D`` (attn_video_vae.py:176-246) as the encoder builds it (no norm, convolution, padding 0): the
    zero row / column the reference appends (`pad (0, 1, 0, 1)`, :239-241) is the TMA unit's zero fill inside the strided convolution."""

    def __init__(self, conv: nn.Module, channels: int):
        super().__init__()
        self.conv = CausalConv3dB200(conv.weight, conv.bias, tuple(conv.stride), pad_hw=1)
        self.channels = channels

    @classmethod
    def from_reference(cls, down: nn.Module) -> "Downsample3DB200":
        conv = down.conv
        if getattr(down, "norm", None) is not None or not getattr(down, "use_conv", True) or getattr(down, "padding", 0) != 0 \
                or not getattr(down, "spatial_down", False) or type(conv).__name__ != "InflatedCausalConv3d":
            raise NotImplementedError("seedvr2_b200.vae_block: down-sampler configuration outside the mirror")
        if conv.in_channels % 64 or conv.out_channels % 128 or tuple(conv.padding)[1:] != (0, 0) or conv.weight.dtype not in _DTYPE_CODE:
            raise NotImplementedError("seedvr2_b200.vae_block: down-sampler convolution outside the kernel's shapes")
        return cls(conv, down.channels)

    def forward(self, hidden_states: torch.Tensor, memory_state="DISABLED", **kwargs) -> torch.Tensor:
        x = to_channels_last(hidden_states)
        return self.conv.forward_cl(x, memory_state).permute(0, 4, 1, 2, 3)


def make_causal_norm_wrapper(original):
    """A drop-in for the reference's ``causal_norm_wrapper(norm_layer, x)`` (causal_inflation_lib.py:354-409): per-frame GroupNorm of a
    5-D CUDA fp16 / bf16 tensor runs on svr2_groupnorm_silu_fwd (no activation, channels-last result — no memory-limit chunking is
    needed, the kernel has no workspace beyond 50 KB of partials); everything else goes to ``original``."""
    def causal_norm_wrapper(norm_layer: nn.Module, x: torch.Tensor) -> torch.Tensor:
        if (isinstance(norm_layer, nn.GroupNorm) and torch.is_tensor(x) and x.ndim == 5 and x.is_cuda and x.dtype in _DTYPE_CODE
                and x.shape[1] % norm_layer.num_groups == 0 and x.shape[1] // norm_layer.num_groups in (2, 4, 8, 16)
                and x.shape[1] % 8 == 0 and 256 % (x.shape[1] // 8) == 0 and norm_layer.weight is not None
                and (norm_layer.weight.dtype == x.dtype or torch.is_autocast_enabled("cuda"))):
            # under the pipeline's autocast (16-bit input, weights of the other 16-bit dtype) ATen would run this norm in fp32;
            # here it stays in the input's dtype, with the two small parameter vectors cast on the fly
            w, b_ = norm_layer.weight, norm_layer.bias
            if w.dtype != x.dtype:
                w, b_ = w.to(x.dtype), None if b_ is None else b_.to(x.dtype)
            xc = to_channels_last(x)
            y = torch.empty_like(xc)
            for b in range(xc.shape[0]):
                groupnorm_silu(xc[b], w, b_, norm_layer.num_groups, norm_layer.eps, silu=False, out=y[b])
            return y.permute(0, 4, 1, 2, 3)
        return original(norm_layer, x)
    causal_norm_wrapper._svr2_original = original
    return causal_norm_wrapper


def patch_causal_norm_wrapper() -> int:
    """Replace ``causal_norm_wrapper`` in every loaded module of the reference's VAE package (the function is imported by name into
    attn_video_vae.py, which calls it for the decoder's / encoder's output norm); returns how many module globals were rebound."""
    import sys
    n = 0
    for name, mod in list(sys.modules.items()):
        if mod is None or "video_vae_v3.modules" not in name:
            continue
        fn = getattr(mod, "causal_norm_wrapper", None)
        if callable(fn) and not hasattr(fn, "_svr2_original"):
            setattr(mod, "causal_norm_wrapper", make_causal_norm_wrapper(fn))
            n += 1
    return n


def clear_memory(model: nn.Module) -> int:
    """Drop the temporal-slice memories of every swapped convolution (what VideoAutoencoderKL.slicing_encode / slicing_decode do for
    ``InflatedCausalConv3d`` after the last slice, attn_video_vae.py:1261-1265, 1285-1289); returns how many held one."""
    n = 0
    for m in model.modules():
        if isinstance(m, CausalConv3dB200) and m.memory is not None:
            m.memory = None
            n += 1
    return n


def _wrap_slicing(model: nn.Module) -> None:
    """The reference clears its convolutions' memories by ``isinstance(m, InflatedCausalConv3d)`` at the end of slicing_encode /
    slicing_decode; the swapped modules are not instances of that class, so the two methods are wrapped to clear theirs as well
    (top-level feature maps of two frames per convolution would otherwise stay allocated until the next clip)."""
    for name in ("slicing_encode", "slicing_decode"):
        fn = getattr(model, name, None)
        if callable(fn) and not getattr(fn, "_svr2_wrapped", False):
            def wrapped(*args, _fn=fn, **kwargs):
                try:
                    return _fn(*args, **kwargs)
                finally:
                    clear_memory(model)
            wrapped._svr2_wrapped = True
            setattr(model, name, wrapped)


def _conv_supported(conv: nn.Module) -> bool:
    try:
        kt, kh, kw = conv.kernel_size
        strided_ok = (kh == 3 and tuple(conv.stride) == ((2 if kt == 3 else 1), 2, 2) and tuple(conv.padding)[1:] == (0, 0)
                      and conv.in_channels % 64 == 0 and conv.out_channels % 128 == 0)
        return ((tuple(conv.stride) == (1, 1, 1) or strided_ok) and tuple(conv.dilation) == (1, 1, 1) and conv.groups == 1 and kh == kw
                and kt in (1, 3) and kh in (1, 3) and conv.weight.dtype in _DTYPE_CODE)
    except (AttributeError, TypeError, ValueError):
        return False


def install_vae_blocks(model: nn.Module, convs: bool = True, norms: bool = True):
    """Replace, in a built (CUDA, fp16 / bf16) VAE, every ``ResnetBlock3D`` this module mirrors and — with ``convs`` — every other
    stride-1 ``InflatedCausalConv3d`` (the up-samplers' 3 x 3 x 3 convolutions — a third of the decoder's FLOPs — and the boundary
    convolutions conv_in / conv_out, which run with zero-padded channels).
    With ``norms`` the reference's ``causal_norm_wrapper`` is rebound (``patch_causal_norm_wrapper``) so that the stand-alone
    GroupNorms (conv_norm_out) run on this library too.  Returns (blocks replaced, convolutions replaced)."""
    n_blocks = n_convs = 0
    if norms:
        patch_causal_norm_wrapper()
    _wrap_slicing(model)
    for parent in list(model.modules()):
        for name, child in list(parent.named_children()):
            if type(child).__name__ != "ResnetBlock3D":
                continue
            try:
                setattr(parent, name, ResnetBlock3DB200.from_reference(child))
                n_blocks += 1
            except NotImplementedError:
                pass          # a configuration outside the mirror keeps the reference's block
    if convs:
        for parent in list(model.modules()):
            for name, child in list(parent.named_children()):
                if type(child).__name__ == "Upsample3D":
                    try:
                        setattr(parent, name, Upsample3DB200.from_reference(child))
                        n_convs += 2
                    except (NotImplementedError, AttributeError):
                        pass
                elif type(child).__name__ == "Attention" and hasattr(child, "to_q") and hasattr(child, "group_norm"):
                    try:
                        from .vae_attn import AttentionB200
                        setattr(parent, name, AttentionB200.from_reference(child))
                        n_convs += 1
                    except (NotImplementedError, AttributeError):
                        pass          # multi-head / cross-attention layers keep the reference's module
                elif type(child).__name__ == "Downsample3D":
                    try:
                        setattr(parent, name, Downsample3DB200.from_reference(child))
                        n_convs += 1
                    except (NotImplementedError, AttributeError):
                        pass
        for parent in list(model.modules()):
            if isinstance(parent, (ResnetBlock3DB200, CausalConv3dB200, Upsample3DB200, Downsample3DB200)) or type(parent).__name__ == "AttentionB200":
                continue
            for name, child in list(parent.named_children()):
                pointwise = type(child) is nn.Conv3d and tuple(child.kernel_size) == (1, 1, 1) and tuple(child.padding) == (0, 0, 0)
                if (type(child).__name__ == "InflatedCausalConv3d" or pointwise) and _conv_supported(child):
                    # `pointwise`: the up-samplers' 1 x 1 x 1 channel expansion in front of the pixel shuffle (attn_video_vae.py:94-96)
                    setattr(parent, name, CausalConv3dB200.from_reference(child))
                    n_convs += 1
    return n_blocks, n_convs
