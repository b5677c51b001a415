"""TEST INFRASTRUCTURE ONLY — never imported by the product path.

CPU restatement of the step either side of the DiT forward in the reference's one-step super-resolution call (SURVEY §8f
row N2): the SR condition, the channel concat that forms the DiT input, and the one-step Euler endpoint of the lerp
schedule.  Plain torch ops on the caller's dtypes, so PyTorch's own type promotion reproduces the reference's rounding
points (the sampler state is bf16 on the GPU path: generation_phases.py:598 passes compute_dtype).

Pinned by tests/golden/sampler_*.pt, produced by the real VideoDiffusionInfer / EulerSampler / LinearInterpolationSchedule
(tests/golden/make_golden.py sampler).
"""
from typing import Callable, Tuple

import torch


def timestep_transform(t: torch.Tensor, latent_shape: torch.Tensor, T: float = 1000.0, vt: int = 4, vs: int = 8) -> torch.Tensor:
    """Resolution-dependent timestep shift, src/core/infer.py:260-289 (config `timesteps.transform: True`,
    configs_3b/main.yaml:80).  latent_shape: int64 [b, 3] latent grid (t, h, w) before patching."""
    frames = (latent_shape[:, 0] - 1) * vt + 1
    heights = latent_shape[:, 1] * vs
    widths = latent_shape[:, 2] * vs

    def lin(x1, y1, x2, y2):
        m = (y2 - y1) / (x2 - x1)
        b = y1 - m * x1
        return lambda x: m * x + b

    img = lin(256 * 256, 1.0, 1024 * 1024, 3.2)
    vid = lin(256 * 256 * 37, 1.0, 1280 * 720 * 145, 5.0)
    shift = torch.where(frames > 1, vid(heights * widths * frames), img(heights * widths))
    t = t / T
    t = shift * t / (1 + (shift - 1) * t)
    return t * T


def _expand(t: torch.Tensor, ndim: int) -> torch.Tensor:
    """src/common/diffusion/utils.py:23-29"""
    return t.reshape(t.shape + (1,) * (ndim - t.ndim))


def schedule_forward(x_0: torch.Tensor, x_T: torch.Tensor, t: torch.Tensor, T: float = 1000.0) -> torch.Tensor:
    """x_t = A(t) x_0 + B(t) x_T with A = 1 - t/T, B = t/T: schedules/base.py:82-87, schedules/lerp.py:45-49."""
    t = _expand(t, x_0.ndim)
    return (1 - (t / T)) * x_0 + (t / T) * x_T


def add_noise(latent: torch.Tensor, aug_noise: torch.Tensor, latent_noise_scale: float, T: float = 1000.0) -> torch.Tensor:
    """Latent-space augmentation of the condition, src/core/generation_phases.py:685-693.  latent: [t, h, w, c]."""
    if latent_noise_scale == 0.0:
        return latent
    t = torch.tensor([1000.0], dtype=latent.dtype) * latent_noise_scale
    shape = torch.tensor(latent.shape[1:])[None]          # (sic) the reference passes (h, w, c) as the "latent shape"
    t = timestep_transform(t, shape, T)
    return schedule_forward(latent, aug_noise, t, T)


def sr_condition(noise: torch.Tensor, latent_blur: torch.Tensor) -> torch.Tensor:
    """task == "sr" branch of get_condition, src/core/infer.py:54-78: [t, h, w, c + 1] = (blurred latent, mask of ones)."""
    t, h, w, c = noise.shape
    cond = torch.zeros([t, h, w, c + 1], dtype=noise.dtype)
    cond[..., :-1] = latent_blur[:]
    cond[..., -1:] = 1.0
    return cond


def v_lerp_endpoint(pred: torch.Tensor, x_t: torch.Tensor, t: torch.Tensor, T: float = 1000.0) -> torch.Tensor:
    """pred_x_0 of the v_lerp prediction type, schedules/base.py:96-110 (EulerSampler.get_endpoint, samplers/base.py:88-98)."""
    t = _expand(t, x_t.ndim)
    A, B = 1 - (t / T), t / T
    return (x_t - B * pred) / (A + B)


def one_step(dit: Callable[[torch.Tensor, torch.Tensor], torch.Tensor], noise: torch.Tensor, condition: torch.Tensor,
             T: float = 1000.0) -> Tuple[torch.Tensor, torch.Tensor]:
    """runner.inference with cfg scale 1 and ONE sampling step (generation_phases.py:594-598): trailing timesteps for
    steps = 1 are [T] (timesteps/sampling/trailing.py:39-50), the Euler loop body never runs and the endpoint is taken at
    t = T (samplers/euler.py:45-66).  noise [t,h,w,c], condition [t,h,w,c+1] -> (DiT input [t*h*w, 2c+1], x_0 [t,h,w,c]).
    `dit(vid, timestep)` returns the prediction [t*h*w, c]."""
    t = (torch.arange(1.0, 0.0, -1.0).to(noise.dtype) * T)[0]           # tensor(1000.) in the sampler dtype
    x = noise.reshape(-1, noise.shape[-1])                              # na.flatten, na.py:69-103 (b = 1)
    cond = condition.reshape(-1, condition.shape[-1])
    vid = torch.cat([x, cond], dim=-1)                                  # infer.py:341
    pred = dit(vid, t.repeat(1))
    x0 = v_lerp_endpoint(pred, x, t, T)
    return vid, x0.reshape(noise.shape)
