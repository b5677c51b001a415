"""Temporal segment split across GPUs and overlap crossfade.  TEST INFRASTRUCTURE ONLY.

Reference: inference_cli.py:1087-1114 (frame ranges per GPU), :1163-1198 (sequential blend of the per-GPU results),
src/core/generation_utils.py:284-312 (``blend_overlapping_frames``: Hann crossfade for overlap >= 3, linear below).
"""
from typing import List, Tuple

import torch


def segment_ranges(total_frames: int, num_devices: int, overlap: int, start_frame: int = 0) -> List[Tuple[int, int]]:
    """[start, end) frames each rank READS; every rank but the last reads `overlap` extra frames, clamped."""
    base, rem = total_frames // num_devices, total_frames % num_devices
    out, cur = [], start_frame
    for idx in range(num_devices):
        n = base + (1 if idx < rem else 0)
        end = cur + n
        if idx < num_devices - 1 and overlap > 0:
            end = min(end + overlap, start_frame + total_frames)
        out.append((cur, end))
        cur += n
    return out


def batch_ranges(total_frames: int, batch_size: int, overlap: int) -> List[Tuple[int, int]]:
    """Latent batches inside one segment, generation_phases.py:259-282 line by line (step, reset, early break)."""
    step = batch_size - overlap if overlap > 0 else batch_size
    if step <= 0:
        step = batch_size
        overlap = 0
    num = 0
    out = []
    for idx in range(0, total_frames, step):
        end_idx = min(idx + batch_size, total_frames)
        if idx > 0 and end_idx - idx <= overlap:
            break
        num += 1
        out.append((idx, end_idx))
    return out


def blend_weights(overlap: int, dtype=torch.float32) -> torch.Tensor:
    """Weight of the PREVIOUS segment's tail per overlapping frame."""
    if overlap >= 3:
        t = torch.linspace(0.0, 1.0, steps=overlap, dtype=dtype)
        u = ((t - 1.0 / 3.0) / (2.0 / 3.0 - 1.0 / 3.0)).clamp(0.0, 1.0)
        return 0.5 + 0.5 * torch.cos(torch.pi * u)
    return torch.linspace(1.0, 0.0, steps=overlap, dtype=dtype)


def blend(prev_tail: torch.Tensor, cur_head: torch.Tensor, overlap: int) -> torch.Tensor:
    w_prev = blend_weights(overlap, prev_tail.dtype).view(overlap, *([1] * (prev_tail.ndim - 1)))
    return prev_tail * w_prev + cur_head * (1.0 - w_prev)


def stitch(chunks: List[torch.Tensor], overlap: int) -> torch.Tensor:
    """Sequential concatenation with crossfade, exactly the parent-process loop of inference_cli.py:1163-1198."""
    if overlap <= 0 or len(chunks) == 1:
        return torch.cat([c.float() for c in chunks], dim=0)
    result = None
    for idx, c in enumerate(chunks):
        c = c.float()
        if idx == 0:
            result = c
        elif c.shape[0] > overlap and result.shape[0] >= overlap:
            blended = blend(result[-overlap:], c[:overlap], overlap)
            result = torch.cat([result[:-overlap], blended, c[overlap:]], dim=0)
        elif c.shape[0] > overlap:
            result = torch.cat([result, c[overlap:]], dim=0)
    return result
