"""GPU post-processing of a decoded batch: the pipeline's default colour correction ("wavelet") and the conversion to output frames.

Mirrors the reference's post-processing of one batch (src/core/generation_phases.py:1286-1337 with ``color_correction="wavelet"``,
src/utils/color_fix.py:122-246) and the writer's uint8 / BGR conversion (inference_cli.py:691-695,734-738); SURVEY §8f row N4.
Same argument meaning as ``wavelet_reconstruction(content_feat, style_feat)``: ``sample`` carries the details (the decoded,
upscaled frames), ``style`` the colours (the transformed input frames), both ``[T, C, H, W]`` in ``[-1, 1]``, any strides, fp32 or
bf16 (the pipeline's compute dtype).  CUDA only — there is no other path."""
from typing import Optional, Tuple

import torch

from . import _lib, ops

LEVELS = 5      # wavelet_decomposition(levels=5), color_fix.py:160


def _dtype_code(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return 0
    if t.dtype == torch.bfloat16:
        return 1
    raise NotImplementedError(f"seedvr2_b200.postprocess: fp32 or bf16 frames only (got {t.dtype})")


def _decompose(img: torch.Tensor, want_high: bool) -> Tuple[Optional[torch.Tensor], torch.Tensor]:
    """(high, low) of wavelet_decomposition; ``img`` [T, C, H, W] with arbitrary strides."""
    T, C, H, W = img.shape
    code = _dtype_code(img)
    dev = img.device
    low_a = torch.empty(T, C, H, W, dtype=img.dtype, device=dev)
    low_b = torch.empty_like(low_a) if LEVELS > 1 else None
    high = torch.empty_like(low_a) if want_high else None
    max_safe = max(1, min(H, W) // 8)          # wavelet_blur: radius capped at 1/8 of the smaller side
    src, st = img, img.stride()
    dst = low_a
    for lvl in range(LEVELS):
        radius = min(2 ** lvl, max_safe)
        ops._call("svr2_wavelet_level_fwd", 1, dev, _lib.lib().svr2_wavelet_level_fwd, src.data_ptr(), st[0], st[1], st[2], st[3],
                  dst.data_ptr(), None if high is None else high.data_ptr(), int(lvl == 0), T, C, H, W, radius, code, ops._stream(img))
        src, st = dst, dst.stride()
        dst = low_b if dst is low_a else low_a
    return high, src


@torch._dynamo.disable
def wavelet_color_fix(sample: torch.Tensor, style: torch.Tensor, want_float: bool = True, want_uint8: bool = False,
                      bgr: bool = False) -> Tuple[Optional[torch.Tensor], Optional[torch.Tensor]]:
    """-> (frames [T, H, W, C] in [0, 1] in the input dtype or None, frames_u8 [T, H, W, C] uint8 or None).

    frames = clamp(wavelet_reconstruction(sample, style), -1, 1) * 0.5 + 0.5 laid out as the pipeline's final video;
    frames_u8 = (frames.float() * 255).astype(uint8), channels swapped to BGR on request (what the reference's writer hands to cv2)."""
    if not sample.is_cuda or not style.is_cuda:
        raise RuntimeError("seedvr2_b200.postprocess: frames must live on a CUDA device; there is no CPU path")
    if sample.dim() != 4 or sample.shape != style.shape:
        # the reference resizes a mismatching style image with bilinear interpolation (color_fix.py:207-221); the pipeline never
        # produces one (both come from the same transform), so it is not restated
        raise NotImplementedError(f"wavelet_color_fix: sample {tuple(sample.shape)} and style {tuple(style.shape)} must be equal [T, C, H, W]")
    if style.dtype != sample.dtype:
        style = style.to(sample.dtype)
    if not (want_float or want_uint8):
        raise ValueError("wavelet_color_fix: nothing requested")
    T, C, H, W = sample.shape
    high, _ = _decompose(sample, True)
    _, low_style = _decompose(style, False)
    out = torch.empty(T, H, W, C, dtype=sample.dtype, device=sample.device) if want_float else None
    out_u8 = torch.empty(T, H, W, C, dtype=torch.uint8, device=sample.device) if want_uint8 else None
    ops._call("svr2_colorfix_finish_fwd", 1, sample.device, _lib.lib().svr2_colorfix_finish_fwd, high.data_ptr(), low_style.data_ptr(),
              None if out is None else out.data_ptr(), None if out_u8 is None else out_u8.data_ptr(), int(bgr), T, C, H, W,
              _dtype_code(sample), ops._stream(sample))
    return out, out_u8


def _histogram_match(source: torch.Tensor, reference: torch.Tensor) -> torch.Tensor:
    """_histogram_matching_channel (color_fix.py:473-513) for equal element counts: rank-wise replacement of the source values by
    the reference values.  A global sort + scatter — torch's radix sort, as in the reference."""
    flat = source.reshape(-1)
    _, idx = torch.sort(flat)
    ref_sorted, _ = torch.sort(reference.reshape(-1))
    out = torch.empty_like(flat)
    out.scatter_(0, idx, ref_sorted)
    return out.reshape(source.shape)


@torch._dynamo.disable
def lab_color_fix(sample: torch.Tensor, style: torch.Tensor, luminance_weight: float = 0.8, want_float: bool = True,
                  want_uint8: bool = False, bgr: bool = False) -> Tuple[Optional[torch.Tensor], Optional[torch.Tensor]]:
    """``lab_color_transfer(sample, style, luminance_weight)`` (color_fix.py:249-367, the CLI's default ``--color_correction lab``)
    followed by the batch normalisation and, on request, the writer's uint8 conversion — same return convention as
    ``wavelet_color_fix``.  RGB only (the pipeline strips and re-attaches alpha around the colour correction)."""
    if not sample.is_cuda or not style.is_cuda:
        raise RuntimeError("seedvr2_b200.postprocess: frames must live on a CUDA device; there is no CPU path")
    if sample.dim() != 4 or sample.shape != style.shape or sample.shape[1] != 3:
        raise NotImplementedError(f"lab_color_fix: sample {tuple(sample.shape)} and style {tuple(style.shape)} must be equal [T, 3, H, W]")
    if style.dtype != sample.dtype:
        style = style.to(sample.dtype)
    code = _dtype_code(sample)
    T, C, H, W = sample.shape
    dev = sample.device
    high, _ = _decompose(sample, True)                    # step 1: wavelet reconstruction (content details + style colours)
    _, low_style = _decompose(style, False)
    lab_c = torch.empty(T, 3, H, W, dtype=torch.float32, device=dev)
    lab_s = torch.empty_like(lab_c)
    L = _lib.lib()
    ops._call("svr2_rgb_to_lab_fwd", 1, dev, L.svr2_rgb_to_lab_fwd, high.data_ptr(), low_style.data_ptr(), lab_c.data_ptr(), T, H, W, code,
              ops._stream(sample))
    sc = style.contiguous()
    ops._call("svr2_rgb_to_lab_fwd", 1, dev, L.svr2_rgb_to_lab_fwd, sc.data_ptr(), None, lab_s.data_ptr(), T, H, W, code, ops._stream(sample))
    del high, low_style
    a = _histogram_match(lab_c[:, 1], lab_s[:, 1])
    b = _histogram_match(lab_c[:, 2], lab_s[:, 2])
    if luminance_weight < 1.0:
        mL = _histogram_match(lab_c[:, 0], lab_s[:, 0])
        rL = lab_c[:, 0].mul(luminance_weight).add_(mL.mul(1.0 - luminance_weight))
    else:
        rL = lab_c[:, 0]
    lab = torch.stack([rL, a, b], dim=1)
    out = torch.empty(T, H, W, 3, dtype=sample.dtype, device=dev) if want_float else None
    out_u8 = torch.empty(T, H, W, 3, dtype=torch.uint8, device=dev) if want_uint8 else None
    ops._call("svr2_lab_to_rgb_finish_fwd", 1, dev, L.svr2_lab_to_rgb_finish_fwd, lab.data_ptr(), None if out is None else out.data_ptr(),
              None if out_u8 is None else out_u8.data_ptr(), int(bgr), T, H, W, code, ops._stream(sample))
    return out, out_u8
