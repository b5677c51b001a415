"""The video VAE's causal 3-D convolution on the B200 (first slice).

``CausalConv3dB200`` mirrors ``InflatedCausalConv3d`` (src/models/video_vae_v3/modules/causal_inflation_lib.py:61-272) for the
stride-1 convolutions of the VAE's residual blocks: same constructor source (an existing module or its weight / bias), same
``forward(input, memory_state)`` meaning — ``[B, C, T, H, W]`` in, ``[B, C_out, T, H, W]`` out, the causal head taken from the
first frame or from the memory the previous temporal slice left behind — and the same memory bookkeeping (``self.memory`` holds the
last ``kernel_t - 1`` input frames unless the state is DISABLED).  ``set_memory_limit`` is accepted and ignored: the reference
splits the convolution spatially to bound cuDNN's workspace (:117-211); the implicit GEMM of csrc/conv3d.cu has no workspace.

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
    """nn.Conv3d weight [C_out, C_in, kt, kh, kw] -> [C_out, kt, kh, kw, C_in], the K axis of the implicit GEMM."""
    if not weight.is_cuda or weight.dtype not in _DTYPE_CODE:
        raise RuntimeError("seedvr2_b200.vae_conv: the weight must be a CUDA fp16 / bf16 tensor; there is no CPU path")
    w = weight.contiguous()
    co, ci, kt, kh, kw = w.shape
    out = torch.empty(co, kt, kh, kw, ci, dtype=w.dtype, device=w.device)
    ops._call("svr2_conv3d_pack_weight", 1, w.device, _lib.lib().svr2_conv3d_pack_weight, w.data_ptr(), out.data_ptr(), co, ci, kt, kh,
              kw, ops._stream(w))
    return out


def causal_conv3d(x: torch.Tensor, weight_packed: torch.Tensor, bias: Optional[torch.Tensor] = None,
                  head: Optional[torch.Tensor] = None, resid: Optional[torch.Tensor] = None,
                  out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x [T, H, W, C_in] -> [T, H, W, C_out]; ``head`` = [kt - 1, H, W, C_in] frames in front of x (None: x[0] repeated);
    ``resid`` [T, H, W, C_out] is added after the convolution's own rounding."""
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
    ops._call("svr2_causal_conv3d_fwd", 1, x.device, _lib.lib().svr2_causal_conv3d_fwd, x.data_ptr(), ops._ptr(head),
              weight_packed.data_ptr(), ops._ptr(bias), ops._ptr(resid), y.data_ptr(), T, H, W, ci, co, kt, kh, code, ops._stream(x))
    return y


class CausalConv3dB200(nn.Module):
    """Drop-in for a stride-1 ``InflatedCausalConv3d``; see the module docstring."""

    def __init__(self, weight: torch.Tensor, bias: Optional[torch.Tensor]):
        super().__init__()
        co, ci, kt, kh, kw = weight.shape
        if kh != kw or kt not in (1, 3) or kh not in (1, 3):
            raise NotImplementedError(f"seedvr2_b200.vae_conv: kernel {(kt, kh, kw)} (temporal 1|3, spatial 1|3 square, stride 1)")
        self.in_channels, self.out_channels, self.kernel_size = ci, co, (kt, kh, kw)
        self.register_buffer("weight_packed", pack_conv_weight(weight.detach()), persistent=False)
        self.register_buffer("bias", None if bias is None else bias.detach().clone(), persistent=False)
        self.memory: Optional[torch.Tensor] = None       # [B, kt - 1, H, W, C] — the reference keeps [B, C, kt - 1, H, W]

    @classmethod
    def from_reference(cls, conv: nn.Module) -> "CausalConv3dB200":
        if tuple(conv.stride) != (1, 1, 1) or tuple(conv.dilation) != (1, 1, 1) or conv.groups != 1:
            raise NotImplementedError("seedvr2_b200.vae_conv: stride-1, undilated, ungrouped convolutions only")
        return cls(conv.weight, conv.bias)

    def set_memory_limit(self, value: float) -> None:      # causal_inflation_lib.py:77-78: no workspace to bound here
        pass

    def forward(self, input: torch.Tensor, memory_state="DISABLED", resid: Optional[torch.Tensor] = None) -> torch.Tensor:
        x = input.permute(0, 2, 3, 4, 1)                                        # [B, T, H, W, C]: contiguous for channels_last_3d
        if not x.is_contiguous():
            x = x.contiguous()
        r = None
        if resid is not None:
            r = resid.permute(0, 2, 3, 4, 1)
            r = r if r.is_contiguous() else r.contiguous()
        return self.forward_cl(x, memory_state, r).permute(0, 4, 1, 2, 3)       # [B, C_out, T, H, W], channels-last strides

    def forward_cl(self, x: torch.Tensor, memory_state="DISABLED", resid: Optional[torch.Tensor] = None) -> torch.Tensor:
        """The same on channels-last tensors: x [B, T, H, W, C_in] -> [B, T, H, W, C_out]."""
        state = _state_name(memory_state)
        if state != "ACTIVE":
            self.memory = None                                                  # :216-217
        kt = self.kernel_size[0]
        B = x.shape[0]
        y = torch.empty(*x.shape[:4], self.out_channels, dtype=x.dtype, device=x.device)
        for b in range(B):
            head = self.memory[b] if (self.memory is not None and kt > 1) else None
            causal_conv3d(x[b], self.weight_packed, self.bias, head, None if resid is None else resid[b], out=y[b])
        if kt > 1 and state != "DISABLED":
            # the last kt - 1 frames of (head, input): basic_forward :229-240, slicing_forward :273-283
            if x.shape[1] >= kt - 1:
                self.memory = x[:, -(kt - 1):].clone()
            else:
                prev = self.memory if self.memory is not None else x[:, :1].expand(-1, kt - 1, -1, -1, -1)
                self.memory = torch.cat((prev, x), dim=1)[:, -(kt - 1):].contiguous()
        return y
