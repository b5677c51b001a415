"""Host-side logic of the VAE binding without a GPU: install_vae_blocks on a host-resident module tree composed from
the reference's real parts — what gets replaced, the packed / padded weight layouts, the rebinding of causal_norm_wrapper, the
slicing wrappers — and the loud failure of the compute path off the GPU."""
import pytest
import torch
from torch import nn

from oracle import refstage

if not refstage.available():
    pytest.skip("reference not staged", allow_module_level=True)

import seedvr2_b200 as svr2      # noqa: E402
from vae_standins import Decoder3D, Downsample3D, reference_parts      # noqa: E402


def test_install_on_a_host_resident_decoder():
    torch.manual_seed(0)
    dec = Decoder3D(16, (128, 128, 128, 128)).eval().half()
    ref_w = dec.conv_in.weight.detach().clone()
    up_w = dec.up_samplers[0].conv.weight.detach().clone()
    n_blocks, n_convs = svr2.vae_block.install_vae_blocks(dec)
    assert (n_blocks, n_convs) == (14, 8)
    names = {type(m).__name__ for m in dec.modules()}
    assert "InflatedCausalConv3d" not in names and {"ResnetBlock3DB200", "Upsample3DB200", "CausalConv3dB200"} <= names
    # conv_in: 16 -> 128 channels, K padded to 64 with zero weights, packed [C_out, kt, kh, kw, C_in]
    ci = dec.conv_in
    assert (ci.in_channels, ci.ci_pad, ci.out_channels, ci.co_pad) == (16, 64, 128, 128)
    assert tuple(ci.weight_packed.shape) == (128, 3, 3, 3, 64)
    assert torch.equal(ci.weight_packed[..., :16], ref_w.permute(0, 2, 3, 4, 1)) and float(ci.weight_packed[..., 16:].abs().max()) == 0
    # conv_out: 128 -> 3, N padded to 128 with zero rows and zero bias
    co = dec.conv_out
    assert (co.out_channels, co.co_pad) == (3, 128) and float(co.weight_packed[3:].abs().max()) == 0 and float(co.bias[3:].abs().max()) == 0
    up = dec.up_samplers[0]
    assert (up.spatial_ratio, up.temporal_ratio, up.channels) == (2, 2, 128)
    assert torch.equal(up.conv.weight_packed, up_w.permute(0, 2, 3, 4, 1).contiguous())
    assert tuple(up.upscale_conv.weight_packed.shape) == (8 * 128, 1, 1, 1, 128)
    # no state-dict entries: the packed weights are derived buffers, installing comes after loading
    assert not any("weight_packed" in k for k in dec.state_dict())
    # the compute path is CUDA-only and says so
    with pytest.raises(RuntimeError, match="no CPU path"):
        dec(torch.zeros(1, 16, 1, 4, 4, dtype=torch.float16), memory_state="DISABLED")


def test_down_sampler_and_norm_rebinding_on_the_host():
    import src.models.video_vae_v3.modules.causal_inflation_lib as lib
    _, original, _, _ = reference_parts()
    holder = nn.Sequential(Downsample3D(128, temporal_down=True).half(), Downsample3D(256, temporal_down=False).half())
    assert svr2.vae_block.install_vae_blocks(holder) == (0, 2)
    assert holder[0].conv.stride == (2, 2, 2) and holder[0].conv.pad_hw == 1 and holder[1].conv.stride == (1, 2, 2)
    fn = lib.causal_norm_wrapper
    assert getattr(fn, "_svr2_original", None) is original          # rebound once, the original kept
    # off the GPU / in fp32 the replacement defers to the original
    gn = nn.GroupNorm(32, 64)
    x = torch.randn(1, 64, 2, 5, 6)
    assert torch.equal(fn(gn, x), original(gn, x))


def test_slicing_methods_are_wrapped_once():
    class Vae(nn.Module):
        def __init__(self):
            super().__init__()
            self.calls = 0

        def slicing_decode(self, z):
            self.calls += 1
            return z

    vae = Vae()
    svr2.vae_block.install_vae_blocks(vae)
    first = vae.slicing_decode
    svr2.vae_block.install_vae_blocks(vae)
    assert vae.slicing_decode is first and getattr(first, "_svr2_wrapped", False)
    assert vae.slicing_decode(3) == 3 and vae.calls == 1
    assert svr2.vae_block.clear_memory(vae) == 0
