"""The video VAE's causal 3-D convolutions on the B200.

``CausalConv3dB200`` mirrors ``InflatedCausalConv3d`` (src/models/video_vae_v3/modules/causal_inflation_lib.py:61-303) for the
VAE's convolutions — kernels (1|3, 1|3, 1|3) with stride 1 (residual blocks, up-samplers, conv_in / conv_out) and the
down-samplers' (1|3, 3, 3) with stride (1|2, 2, 2): same constructor source (an existing module or its weight / bias), same
``forward(input, memory_state)`` meaning — ``[B, C, T, H, W]`` (or a list of temporal pieces) in, ``[B, C_out, T', H', W']`` out, the
causal head taken from the first frame or from the memory the previous temporal slice left behind — and the same memory
bookkeeping (``self.memory`` holds the last ``kernel_t - stride_t`` input frames unless the state is DISABLED).
``set_memory_limit`` is accepted and ignored: the reference splits the convolution spatially to bound cuDNN's workspace
(:117-211); the implicit GEMM of csrc/conv3d.cu has no workspace.

Tensors are kept channels-last (``torch.channels_last_3d``): a ``[B, C, T, H, W]`` tensor in that format is the kernel's
``[T, H, W, C]`` layout as it stands, and the result is returned as such a view, so a chain of these modules never transposes.
CUDA only — there is no other path."""
from typing import Optional

import torch
from torch import nn

from . import _lib, ops

_DTYPE_CODE = {torch.bfloat16: 1, torch.float16: 2}


def _state_name(memory_state) -> str:
    """The reference's MemoryState enum (modules/types.py) or its name as a string."""
    name = getattr(memory_state, "name", memory_state)
    if name not in ("DISABLED", "INITIALIZING", "ACTIVE"):
        raise AssertionError(f"memory_state must be DISABLED, INITIALIZING or ACTIVE (got {memory_state!r})")      # :215 assert
    return name


def pack_conv_weight(weight: torch.Tensor) -> torch.Tensor:
    """nn.Conv3d weight [C_out, C_in, kt, kh, kw] -> [C_out, kt, kh, kw, C_in], the K axis of the implicit GEMM.  A weight that is
    still on the host (the pipeline keeps the VAE off the GPU between its phases) is re-laid out there, once, at install time —
    the convolution itself has no CPU path."""
    if weight.dtype not in _DTYPE_CODE:
        raise RuntimeError("seedvr2_b200.vae_conv: the weight must be an fp16 / bf16 tensor")
    if not weight.is_cuda:
        return weight.permute(0, 2, 3, 4, 1).contiguous()
    w = weight.contiguous()
    co, ci, kt, kh, kw = w.shape
    out = torch.empty(co, kt, kh, kw, ci, dtype=w.dtype, device=w.device)
    ops._call("svr2_conv3d_pack_weight", 1, w.device, _lib.lib().svr2_conv3d_pack_weight, w.data_ptr(), out.data_ptr(), co, ci, kt, kh,
              kw, ops._stream(w))
    return out


def transpose_2b(src: torch.Tensor, dst: torch.Tensor, rows: int, cols: int, ld_src: int, ld_dst: int, batch: int, bs_src: int,
                 bs_dst: int) -> None:
    """dst[b][j][i] = src[b][i][j] (2-byte elements): the layout change at the boundary with the reference's NCDHW modules."""
    ops._call("svr2_transpose_2b", 1, src.device, _lib.lib().svr2_transpose_2b, src.data_ptr(), dst.data_ptr(), rows, cols, ld_src, ld_dst,
              batch, bs_src, bs_dst, ops._stream(src))


def to_channels_last(t: torch.Tensor) -> torch.Tensor:
    """[B, C, T, H, W] in any layout -> contiguous [B, T, H, W, C].  A channels_last_3d tensor is returned as a view; a contiguous
    NCDHW tensor (what the reference's own modules produce) goes through svr2_transpose_2b; anything else through a strided copy."""
    x = t.permute(0, 2, 3, 4, 1)
    if x.is_contiguous():
        return x
    if t.is_contiguous() and t.is_cuda and t.element_size() == 2 and t.numel() > 0:
        B, C, T, H, W = t.shape
        out = torch.empty(B, T, H, W, C, dtype=t.dtype, device=t.device)
        n = T * H * W
        transpose_2b(t, out, C, n, n, C, B, C * n, C * n)
        return out
    return x.contiguous()


def to_ncdhw(y: torch.Tensor, channels: Optional[int] = None) -> torch.Tensor:
    """Contiguous [B, T, H, W, C_stored] -> contiguous NCDHW [B, channels, T, H, W] (the first `channels` of the stored ones)."""
    B, T, H, W, cs = y.shape
    c = cs if channels is None else channels
    n = T * H * W
    if not (y.is_contiguous() and y.is_cuda and y.element_size() == 2 and y.numel() > 0):
        return y[..., :c].permute(0, 4, 1, 2, 3).contiguous()
    out = torch.empty(B, c, T, H, W, dtype=y.dtype, device=y.device)
    transpose_2b(y, out, n, c, cs, n, B, n * cs, c * n)
    return out


class GroupStats:
    """Statistics partials of a tensor gathered by the convolution that produced it (``svr2_causal_conv3d_fwd`` gn_partials): what
    the per-frame GroupNorm over 32 groups that follows needs instead of its own pass over the tensor."""

    def __init__(self, partials: torch.Tensor, slots: int, of: torch.Tensor):
        self.partials, self.slots = partials, slots
        self._ptr, self._shape, self._version = of.data_ptr(), tuple(of.shape), of._version

    def valid_for(self, t: torch.Tensor) -> bool:
        """Still describes ``t``: same storage, same shape, not written since."""
        return t.data_ptr() == self._ptr and tuple(t.shape) == self._shape and t._version == self._version


def gn_slots() -> int:
    return int(_lib.lib().svr2_conv3d_gn_slots())


def causal_conv3d(x: torch.Tensor, weight_packed: torch.Tensor, bias: Optional[torch.Tensor] = None,
                  head: Optional[torch.Tensor] = None, resid: Optional[torch.Tensor] = None,
                  out: Optional[torch.Tensor] = None, gn_partials: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x [T, H, W, C_in] -> [T, H, W, C_out]; ``head`` = [kt - 1, H, W, C_in] frames in front of x (None: x[0] repeated);
    ``resid`` [T, H, W, C_out] is added after the convolution's own rounding; ``gn_partials`` [T, gn_slots(), 32, 3] fp32 receives the
    output's GroupNorm statistics."""
    code = _DTYPE_CODE.get(x.dtype)
    if code is None:
        raise NotImplementedError(f"seedvr2_b200.vae_conv: fp16 or bf16 activations only (got {x.dtype})")
    ops._need(x, "x", x.dtype)
    ops._need(weight_packed, "weight_packed", x.dtype)
    T, H, W, ci = x.shape
    co, kt, kh, kw, ci_w = weight_packed.shape
    if ci_w != ci or kh != kw:
        raise RuntimeError(f"seedvr2_b200.vae_conv: weight {tuple(weight_packed.shape)} does not fit input channels {ci}")
    if bias is not None:
        ops._need(bias, "bias", x.dtype)
    if head is not None:
        ops._need(head, "head", x.dtype)
        if tuple(head.shape) != (kt - 1, H, W, ci):
            raise RuntimeError(f"seedvr2_b200.vae_conv: head must be {(kt - 1, H, W, ci)}, got {tuple(head.shape)}")
    if resid is not None:
        ops._need(resid, "resid", x.dtype)
        if tuple(resid.shape) != (T, H, W, co):
            raise RuntimeError(f"seedvr2_b200.vae_conv: resid must be {(T, H, W, co)}, got {tuple(resid.shape)}")
    y = torch.empty(T, H, W, co, dtype=x.dtype, device=x.device) if out is None else ops._need(out, "out", x.dtype)
    if gn_partials is not None:
        ops._need(gn_partials, "gn_partials", torch.float32)
        if gn_partials.numel() != T * gn_slots() * 32 * 3:
            raise RuntimeError(f"seedvr2_b200.vae_conv: gn_partials must hold {T} x {gn_slots()} x 32 x 3 floats")
    ops._call("svr2_causal_conv3d_fwd", 1, x.device, _lib.lib().svr2_causal_conv3d_fwd, x.data_ptr(), ops._ptr(head),
              weight_packed.data_ptr(), ops._ptr(bias), ops._ptr(resid), y.data_ptr(), ops._ptr(gn_partials), T, H, W, ci, co, kt, kh,
              code, ops._stream(x))
    return y


def causal_conv3d_down(x: torch.Tensor, weight_packed: torch.Tensor, bias: Optional[torch.Tensor] = None,
                       head: Optional[torch.Tensor] = None, pad_hw: int = 1) -> torch.Tensor:
    """The encoder's down-sampling convolution: x [T, H, W, C_in] -> [T', H', W', C_out], kernel (kt, 3, 3), stride (2 if kt == 3 else
    1, 2, 2), ``pad_hw`` zero rows / columns after the last one (Downsample3D's pad (0, 1, 0, 1)); ``head`` [1 | 2, H, W, C_in] = the
    previous temporal slice's memory (None: the first frame twice)."""
    code = _DTYPE_CODE.get(x.dtype)
    if code is None:
        raise NotImplementedError(f"seedvr2_b200.vae_conv: fp16 or bf16 activations only (got {x.dtype})")
    ops._need(x, "x", x.dtype)
    ops._need(weight_packed, "weight_packed", x.dtype)
    T, H, W, ci = x.shape
    co, kt, kh, kw, ci_w = weight_packed.shape
    if ci_w != ci or (kh, kw) != (3, 3) or kt not in (1, 3):
        raise RuntimeError(f"seedvr2_b200.vae_conv: weight {tuple(weight_packed.shape)} is not a (1|3, 3, 3) kernel over {ci} channels")
    hf = 0
    if head is not None:
        ops._need(head, "head", x.dtype)
        hf = head.shape[0]
        if kt != 3 or tuple(head.shape[1:]) != (H, W, ci):
            raise RuntimeError(f"seedvr2_b200.vae_conv: head {tuple(head.shape)} does not fit")
    if bias is not None:
        ops._need(bias, "bias", x.dtype)
    To = (T + (hf if head is not None else 2) - 3) // 2 + 1 if kt == 3 else T
    Ho, Wo = (H + pad_hw - 3) // 2 + 1, (W + pad_hw - 3) // 2 + 1
    y = torch.empty(max(To, 0), Ho, Wo, co, dtype=x.dtype, device=x.device)
    ops._call("svr2_causal_conv3d_down_fwd", 1, x.device, _lib.lib().svr2_causal_conv3d_down_fwd, x.data_ptr(), ops._ptr(head), hf,
              weight_packed.data_ptr(), ops._ptr(bias), y.data_ptr(), T, H, W, ci, co, kt, int(pad_hw), code, ops._stream(x))
    return y


class CausalConv3dB200(nn.Module):
    """Drop-in for the VAE's ``InflatedCausalConv3d``s; see the module docstring."""

    def __init__(self, weight: torch.Tensor, bias: Optional[torch.Tensor], stride=(1, 1, 1), pad_hw: int = 0):
        super().__init__()
        co, ci, kt, kh, kw = weight.shape
        if kh != kw or kt not in (1, 3) or kh not in (1, 3):
            raise NotImplementedError(f"seedvr2_b200.vae_conv: kernel {(kt, kh, kw)} (temporal 1|3, spatial 1|3 square)")
        stride = tuple(stride)
        if stride != (1, 1, 1) and not (kh == 3 and stride == ((2 if kt == 3 else 1), 2, 2)):
            raise NotImplementedError(f"seedvr2_b200.vae_conv: stride {stride} with kernel {(kt, kh, kw)} (1,1,1 or the down-samplers' (1|2, 2, 2))")
        self.stride, self.pad_hw = stride, int(pad_hw)      # pad_hw: zero rows / columns after the last one, folded into the kernel
        self.in_channels, self.out_channels, self.kernel_size = ci, co, (kt, kh, kw)
        # the kernel's K step is 64 input channels and its N tile 128 output channels: the VAE's boundary convolutions (conv_in
        # 16 -> 512, conv_out 128 -> 3) run with zero-padded channels — zero weights, zero inputs, the extra outputs dropped
        self.ci_pad, self.co_pad = -(-ci // 64) * 64, -(-co // 128) * 128
        w, b = weight.detach(), None if bias is None else bias.detach().clone()
        if (self.ci_pad, self.co_pad) != (ci, co):
            wp = torch.zeros(self.co_pad, self.ci_pad, kt, kh, kw, dtype=w.dtype, device=w.device)
            wp[:co, :ci] = w
            w = wp
            if b is not None:
                b = torch.cat((b, torch.zeros(self.co_pad - co, dtype=b.dtype, device=b.device)))
        self.register_buffer("weight_packed", pack_conv_weight(w), persistent=False)
        self.register_buffer("bias", b, persistent=False)
        self.memory: Optional[torch.Tensor] = None       # [B, kt - 1, H, W, C] — the reference keeps [B, C, kt - 1, H, W]

    @classmethod
    def from_reference(cls, conv: nn.Module) -> "CausalConv3dB200":
        if tuple(conv.dilation) != (1, 1, 1) or conv.groups != 1:
            raise NotImplementedError("seedvr2_b200.vae_conv: undilated, ungrouped convolutions only")
        if tuple(conv.stride) != (1, 1, 1) and tuple(conv.padding)[1:] != (0, 0):
            raise NotImplementedError("seedvr2_b200.vae_conv: strided convolutions without spatial padding only (Downsample3D pads itself)")
        return cls(conv.weight, conv.bias, tuple(conv.stride))

    def set_memory_limit(self, value: float) -> None:      # causal_inflation_lib.py:77-78: no workspace to bound here
        pass

    def _compute_dtype(self) -> torch.dtype:
        """The dtype the reference's convolution would compute in: the autocast dtype where the pipeline wraps the VAE in
        ``torch.autocast`` because its inputs and the VAE's weights differ (src/core/infer.py:156-158, 232-234), else the weights'."""
        if torch.is_autocast_enabled("cuda"):
            dt = torch.get_autocast_dtype("cuda")
            if dt in _DTYPE_CODE:
                return dt
        return self.weight_packed.dtype

    def _weights_for(self, dtype: torch.dtype):
        """(packed weight, bias) in ``dtype``; the cast of a non-native dtype is made once and kept."""
        if dtype == self.weight_packed.dtype:
            return self.weight_packed, self.bias
        cache = self.__dict__.setdefault("_cast_cache", {})
        hit = cache.get(dtype)
        if hit is None or hit[0].device != self.weight_packed.device:
            hit = (self.weight_packed.to(dtype), None if self.bias is None else self.bias.to(dtype))
            cache[dtype] = hit
        return hit

    def forward(self, input, memory_state="DISABLED", resid: Optional[torch.Tensor] = None):
        """``input`` [B, C, T, H, W], or a list of such tensors = consecutive temporal pieces (the up-sampler's slicing,
        attn_video_vae.py:126-133 -> slicing_forward :243-303): each piece's causal head is the previous piece's tail."""
        if isinstance(input, (list, tuple)):
            if resid is not None:
                raise NotImplementedError("seedvr2_b200.vae_conv: a residual with a list input")
            xs = [self._to_cl(t) for t in input]
            return [self._from_cl(y) for y in self._forward_pieces(xs, memory_state)]
        r = None if resid is None else self._to_cl(resid)
        return self._from_cl(self.forward_cl(self._to_cl(input), memory_state, r))

    def _from_cl(self, y: torch.Tensor) -> torch.Tensor:
        """[B, T, H, W, C_out] -> [B, C_out, T, H, W]: a channels-last view, or (padded outputs) a contiguous copy of the real ones."""
        if self.co_pad != self.out_channels:
            return to_ncdhw(y, self.out_channels)
        return y.permute(0, 4, 1, 2, 3)

    @staticmethod
    def _to_cl(t: torch.Tensor) -> torch.Tensor:
        return to_channels_last(t)                                              # [B, T, H, W, C]: a view for channels_last_3d

    def forward_cl(self, x: torch.Tensor, memory_state="DISABLED", resid: Optional[torch.Tensor] = None, want_stats: bool = False):
        """The same on channels-last tensors: x [B, T, H, W, C_in] -> [B, T, H, W, C_out].  ``want_stats`` (batch 1, C_out 128 / 256 /
        512): returns (y, GroupStats of y) — the statistics the GroupNorm after this convolution needs."""
        if want_stats and x.shape[0] == 1 and self.out_channels in (128, 256, 512) and self.ci_pad == self.in_channels:
            part = torch.empty(x.shape[1], gn_slots(), 32, 3, dtype=torch.float32, device=x.device)
            y = self._forward_pieces([x], memory_state, resid, part)[0]
            return y, GroupStats(part, gn_slots(), y)
        y = self._forward_pieces([x], memory_state, resid)[0]
        return (y, None) if want_stats else y

    def _forward_pieces(self, xs, memory_state, resid: Optional[torch.Tensor] = None, gn_partials: Optional[torch.Tensor] = None):
        state = _state_name(memory_state)
        if state != "ACTIVE":
            self.memory = None                                                  # :216-217
        kt = self.kernel_size[0]
        if self.stride != (1, 1, 1):
            return self._forward_strided(xs, state, resid)
        cd = self._compute_dtype()
        wq, bq = self._weights_for(cd)
        xs = [x if x.dtype == cd else x.to(cd) for x in xs]
        if resid is not None and resid.dtype != cd:
            resid = resid.to(cd)
        head = self.memory if kt > 1 else None                                  # [B, kt - 1, H, W, C] or None (first frame repeated)
        if head is not None and head.dtype != cd:
            head = head.to(cd)
        if resid is not None and self.co_pad != self.out_channels:
            raise NotImplementedError("seedvr2_b200.vae_conv: a residual on a convolution with padded output channels")
        ys = []
        fresh = False                  # `head` is a tensor of its own (not a view of a piece)
        for i, x in enumerate(xs):
            B = x.shape[0]
            if self.ci_pad != self.in_channels:
                xp = torch.zeros(*x.shape[:4], self.ci_pad, dtype=x.dtype, device=x.device)
                xp[..., :self.in_channels] = x
                x = xp
            y = torch.empty(*x.shape[:4], self.co_pad, dtype=x.dtype, device=x.device)
            for b in range(B):
                causal_conv3d(x[b], wq, bq, None if head is None else head[b],
                              None if resid is None else resid[b], out=y[b], gn_partials=gn_partials)
            ys.append(y)
            if kt > 1 and (i + 1 < len(xs) or state != "DISABLED"):
                # the last kt - 1 frames of (head, piece): the next piece's head (:286-296) and, at the end, the memory
                k = x.shape[1]
                if k >= kt - 1:
                    head, fresh = x[:, -(kt - 1):], False
                    if not head.is_contiguous():
                        head, fresh = head.contiguous(), True
                elif k > 0:
                    # a piece shorter than the head (the one-frame first slice of every clip): shift the old head, append the piece
                    nh = torch.empty(B, kt - 1, *x.shape[2:], dtype=x.dtype, device=x.device)
                    if head is not None:
                        nh[:, :kt - 1 - k] = head[:, k:]
                    else:
                        nh[:, :kt - 1 - k] = x[:, :1]              # the replicated first frame (broadcast over the slot axis)
                    nh[:, kt - 1 - k:] = x
                    head, fresh = nh, True
        if kt > 1 and state != "DISABLED":
            # basic_forward :229-240, slicing_forward :273-283; a view of a piece is copied so the piece itself can be freed
            self.memory = head if fresh else (None if head is None else head.clone())
        return ys

    def _forward_strided(self, xs, state: str, resid):
        """Down-sampler form: kernel (kt, 3, 3), stride (kt == 3 ? 2 : 1, 2, 2); the memory is the last kernel_t - stride_t = 1 input
        frame (slicing_forward :262, :273-283)."""
        if resid is not None or self.ci_pad != self.in_channels or self.co_pad != self.out_channels:
            raise NotImplementedError("seedvr2_b200.vae_conv: strided convolutions take unpadded channels and no residual")
        kt = self.kernel_size[0]
        cd = self._compute_dtype()
        wq, bq = self._weights_for(cd)
        xs = [x if x.dtype == cd else x.to(cd) for x in xs]
        head = self.memory if kt > 1 else None
        if head is not None and head.dtype != cd:
            head = head.to(cd)
        ys = []
        for x in xs:
            outs = [causal_conv3d_down(x[b], wq, bq, None if head is None else head[b], self.pad_hw)
                    for b in range(x.shape[0])]
            ys.append(outs[0][None] if len(outs) == 1 else torch.stack(outs, 0))
            if kt > 1 and x.shape[1] >= 1:
                head = x[:, -1:]
        if kt > 1 and state != "DISABLED" and head is not None:
            self.memory = head.clone()
        return ys
