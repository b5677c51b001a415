"""The one-step SR call around the DiT forward (comfyui-seedvr2_b200/sampler.py, csrc/sampler.cu) through the C ABI, against
the golden vectors of the real VideoDiffusionInfer / EulerSampler (tests/golden/sampler_*.pt) and the oracle.  Byte-exact."""
import os

import pytest
import torch

import cases

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
BF16 = [c for c in cases.SAMPLER_CASES if c.dtype == "bf16"]


@pytest.mark.parametrize("case", BF16, ids=lambda c: c.name)
def test_one_step_call_matches_reference_golden(svr2, cuda, case):
    gold = torch.load(os.path.join(GOLD, f"sampler_{case.name}.pt"))
    latent, base_noise, aug_noise, mix = cases.sampler_inputs(case)
    seen = {}
    fake = cases.standin_dit(mix)

    def dit(vid, txt, vid_shape, txt_shape, timestep, grid=None):
        # the stand-in runs on the host so that its own arithmetic is the golden's, bit for bit
        seen["vid"], seen["timestep"], seen["grid"] = vid.cpu(), timestep.cpu(), grid
        return fake(vid.cpu(), timestep.cpu()).to(cuda)

    x0 = svr2.sampler.upscale_latent(dit, base_noise.to(cuda), latent.to(cuda), torch.zeros(58, 8, device=cuda).bfloat16(),
                                     aug_noise.to(cuda), case.latent_noise_scale)
    assert torch.equal(seen["vid"], gold["dit_input"]), "DiT input (noise | condition | mask) differs from the reference"
    assert seen["timestep"].tolist() == gold["timestep"].tolist() and tuple(seen["grid"]) == case.grid
    assert x0.dtype == gold["x0"].dtype and torch.equal(x0.cpu(), gold["x0"])


@pytest.mark.parametrize("grid", [(9, 180, 320), (1, 256, 256), (1, 3, 5)])
def test_full_size_properties(svr2, cuda, grid):
    """BASELINE sizes (C2: 518 400 latent rows): slices of the DiT input are the inputs themselves, the mask is all ones, the
    augmented condition equals the fp32 torch expression rounded once, and the endpoint equals torch's bf16 x_t - pred."""
    t, h, w = grid
    g = torch.Generator(device="cuda").manual_seed(11)
    noise, latent, aug = (torch.randn(t, h, w, 16, generator=g, device=cuda).bfloat16() for _ in range(3))
    vid = svr2.sampler.sr_input(noise, latent)
    assert vid.shape == (t * h * w, 33)
    assert torch.equal(vid[:, :16], noise.view(-1, 16)) and torch.equal(vid[:, 16:32], latent.view(-1, 16))
    assert bool((vid[:, 32] == 1).all())
    A, B = svr2.sampler.latent_noise_coeffs(tuple(latent.shape), 0.3)
    vid = svr2.sampler.sr_input(noise, latent, aug, 0.3)
    want = (torch.tensor([A], device=cuda) * latent + torch.tensor([B], device=cuda) * aug).bfloat16()   # 1-d fp32 coefficients promote: fp32 math, one rounding
    assert torch.equal(vid[:, 16:32], want.view(-1, 16)) and torch.equal(vid[:, :16], noise.view(-1, 16))
    pred = torch.randn(t * h * w, 16, generator=g, device=cuda).bfloat16()
    x0 = svr2.ops.v_lerp_endpoint(noise.view(-1, 16), pred)
    assert torch.equal(x0, noise.view(-1, 16) - pred)
    # general coefficients: every op rounds to bf16 like ATen's bf16 kernels
    Ab, Bb = torch.tensor(0.25, device=cuda).bfloat16(), torch.tensor(0.75, device=cuda).bfloat16()
    x0 = svr2.ops.v_lerp_endpoint(noise.view(-1, 16), pred, float(Ab), float(Bb))
    assert torch.equal(x0, (noise.view(-1, 16) - Bb * pred) / (Ab + Bb))


def test_upscale_latent_through_the_stack(svr2, cuda):
    """The composed call on a narrow 3B-architecture NaDiTStack equals concat -> forward -> x_t - pred done by hand."""
    torch.manual_seed(3)
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        with torch.device(cuda):
            st = svr2.NaDiTStack("3b", num_layers=2, vid_dim=256, heads=2, mm_layers=1, txt_in_dim=64).eval()
    finally:
        torch.set_default_dtype(prev)
    grid = (2, 40, 72)
    g = torch.Generator(device="cuda").manual_seed(12)
    noise, latent = (torch.randn(*grid, 16, generator=g, device=cuda).bfloat16() for _ in range(2))
    txt = torch.randn(58, 64, generator=g, device=cuda).bfloat16()
    x0 = svr2.sampler.upscale_latent(st.forward, noise, latent, txt)
    vid = torch.cat([noise.view(-1, 16), latent.view(-1, 16), torch.ones(noise.numel() // 16, 1, device=cuda).bfloat16()], -1)
    pred = st(vid, txt, None, torch.tensor([[58]]), torch.full((1,), 1000.0, device=cuda), grid=grid)
    assert torch.equal(x0.view(-1, 16), noise.view(-1, 16) - pred)


def test_errors(svr2, cuda):
    a = torch.zeros(2, 2, 2, 16, device=cuda).bfloat16()
    with pytest.raises(RuntimeError):
        svr2.ops.sr_input(a.cpu(), a.cpu())                      # no CPU path
    with pytest.raises(NotImplementedError):
        svr2.ops.sr_input(a.float(), a.float())                  # bf16 only
    with pytest.raises(RuntimeError):
        svr2.sampler.sr_input(a, a, None, 0.5)                   # augmentation without aug_noise
    with pytest.raises(RuntimeError):
        svr2.ops.v_lerp_endpoint(a, a, 1.0, -1.0)                # A + B == 0
