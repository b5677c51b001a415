"""CPU: host-side scalar logic of the one-step SR call (comfyui-seedvr2_b200/sampler.py) against the oracle and the golden
vectors of the real reference.  No compute call into the library."""
import os

import pytest
import torch

import cases

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_latent_noise_coefficients_follow_the_oracle(svr2):
    """Host scalar math of the augmentation (timestep transform with the reference's (h, w, c) argument quirk)."""
    from oracle import sampler as osamp
    for shape, scale in (((3, 6, 10, 16), 0.25), ((1, 12, 16, 16), 1.0), ((9, 180, 320, 16), 0.1), ((1, 256, 256, 16), 0.7)):
        A, B = svr2.sampler.latent_noise_coeffs(shape, scale)
        t = torch.tensor([1000.0], dtype=torch.bfloat16) * scale
        t = osamp.timestep_transform(t, torch.tensor(shape[1:])[None])
        assert A == float(1 - t / 1000.0) and B == float(t / 1000.0)
    assert svr2.sampler.latent_noise_coeffs((3, 6, 10, 16), 0.0) is None


def test_one_step_timestep_is_the_reference_one(svr2):
    for c in cases.SAMPLER_CASES:
        gold = torch.load(os.path.join(GOLD, f"sampler_{c.name}.pt"))
        dt = torch.bfloat16 if c.dtype == "bf16" else torch.float32
        ts = svr2.sampler.one_step_timestep(dt)
        assert ts.dtype == dt and [float(ts)] == gold["timestep"].tolist()


def test_host_tensors_are_refused(svr2):
    a = torch.zeros(2, 2, 2, 16, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="no CPU path"):
        svr2.sampler.sr_input(a, a)
    with pytest.raises(RuntimeError, match="no CPU path"):
        svr2.ops.v_lerp_endpoint(a, a)
