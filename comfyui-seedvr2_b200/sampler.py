"""The reference's one-step super-resolution call around the DiT forward, as stream-ordered device work.

Mirrors, for the path the product actually takes (task "sr", cfg scale 1, ONE sampling step — generation_phases.py:594-598):
  * `latent_noise_coeffs`  — the scalar side of `_add_noise` (generation_phases.py:685-693): t = 1000*scale in the compute
    dtype, `timestep_transform` (infer.py:260-289), A = 1 - t/T, B = t/T (schedules/lerp.py:45-49).  Host scalar math on
    0-d/1-element CPU tensors with the reference's dtypes, so its promotions (bf16 -> fp32 after the transform) carry over.
  * `sr_input`             — get_condition + torch.cat([x_t, cond], -1) (infer.py:54-78,341), one kernel.
  * `one_step_timestep`    — UniformTrailingSamplingTimesteps(T, steps=1) = [T] (timesteps/sampling/trailing.py:39-50).
  * `upscale_latent`       — runner.inference(noises, conditions, texts) for b = 1: DiT forward at t = T, then the Euler
    endpoint x_0 = (x_t - B pred)/(A + B) with A = 0, B = 1 (samplers/euler.py:45-66, schedules/base.py:107-109).
All tensors stay on the device; nothing here touches oracle/ or has a CPU path.
"""
from typing import Callable, Optional, Tuple

import torch

from . import ops

T_SCHEDULE = 1000.0            # configs_3b/main.yaml:66-68, configs_7b/main.yaml (lerp schedule, T = 1000.0)


def timestep_transform(t: torch.Tensor, shape: Tuple[int, int, int], T: float = T_SCHEDULE, vt: int = 4, vs: int = 8) -> torch.Tensor:
    """infer.py:260-289 on host tensors.  `shape` is whatever triple the caller hands the reference as "latents_shapes"."""
    s = torch.tensor([list(shape)], dtype=torch.long)
    frames = (s[:, 0] - 1) * vt + 1
    heights, widths = s[:, 1] * vs, s[:, 2] * vs

    def lin(x1, y1, x2, y2):
        m = (y2 - y1) / (x2 - x1)
        return lambda x: m * x + (y1 - m * x1)

    img, vid = lin(256 * 256, 1.0, 1024 * 1024, 3.2), lin(256 * 256 * 37, 1.0, 1280 * 720 * 145, 5.0)
    shift = torch.where(frames > 1, vid(heights * widths * frames), img(heights * widths))
    t = t / T
    t = shift * t / (1 + (shift - 1) * t)
    return t * T


def latent_noise_coeffs(latent_shape: Tuple[int, ...], latent_noise_scale: float, dtype: torch.dtype = torch.bfloat16,
                        T: float = T_SCHEDULE) -> Optional[Tuple[float, float]]:
    """(A, B) of latent_blur = A*latent + B*aug_noise, or None when the scale is 0 (the reference returns the latent itself).
    The reference passes x.shape[1:] = (h, w, c) of the [t, h, w, c] latent to timestep_transform (generation_phases.py:689)
    — reproduced, not corrected."""
    if latent_noise_scale == 0.0:
        return None
    t = torch.tensor([1000.0], dtype=dtype) * latent_noise_scale
    t = timestep_transform(t, tuple(latent_shape[1:]), T)
    if t.dtype != torch.float32:
        raise NotImplementedError("latent_noise_coeffs: the reference's transform yields fp32 here; anything else is untested")
    A, B = 1 - (t / T), t / T
    return float(A), float(B)


def sr_input(noise: torch.Tensor, latent: torch.Tensor, aug_noise: Optional[torch.Tensor] = None,
             latent_noise_scale: float = 0.0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """DiT input rows [t*h*w, 2c+1] from noise / latent / aug_noise [t, h, w, c] (bf16, on the device)."""
    ab = latent_noise_coeffs(tuple(latent.shape), latent_noise_scale, latent.dtype)
    if ab is None:
        return ops.sr_input(noise, latent, None, out=out)
    if aug_noise is None:
        raise RuntimeError("sr_input: latent_noise_scale > 0 needs aug_noise")
    return ops.sr_input(noise, latent, aug_noise, ab[0], ab[1], out=out)


def one_step_timestep(dtype: torch.dtype = torch.bfloat16, T: float = T_SCHEDULE) -> torch.Tensor:
    """The single trailing timestep, in the sampler dtype (trailing.py:39-50 with steps = 1, shift = 1): tensor(T)."""
    t = torch.arange(1.0, 0.0, -1.0).to(dtype)
    return (t * T)[0]


def upscale_latent(dit: Callable[..., torch.Tensor], noise: torch.Tensor, latent: torch.Tensor, txt: torch.Tensor,
                   aug_noise: Optional[torch.Tensor] = None, latent_noise_scale: float = 0.0,
                   T: float = T_SCHEDULE) -> torch.Tensor:
    """x_0 [t, h, w, c] of one SR step.  `dit(vid, txt, vid_shape, txt_shape, timestep, grid=...)` is NaDiTStack.forward
    (or any callable with its signature) returning the prediction rows [t*h*w, c]."""
    t_, h_, w_, c_ = noise.shape
    vid = sr_input(noise, latent, aug_noise, latent_noise_scale)
    ts = one_step_timestep(noise.dtype, T)
    timestep = torch.full((1,), float(ts), dtype=torch.float32, device=noise.device)
    txt_shape = torch.tensor([[txt.shape[0]]], dtype=torch.long)
    pred = dit(vid, txt, None, txt_shape, timestep, grid=(t_, h_, w_))
    Bt = ts / T                                          # 0-d tensors in the sampler dtype, as schedules/lerp.py:45-49
    A, B = float(1 - Bt), float(Bt)
    x0 = ops.v_lerp_endpoint(noise.reshape(-1, c_), pred.contiguous(), A, B)
    return x0.view(t_, h_, w_, c_)
