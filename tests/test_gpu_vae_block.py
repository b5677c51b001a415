"""svr2_groupnorm_silu_fwd and the VAE residual block (comfyui-seedvr2_b200/vae_block.py) on the GPU.

GroupNorm + SiLU: against ATen's own two ops on the same GPU in the same dtype (the reference's path: causal_norm_wrapper ->
nn.GroupNorm per frame, then nn.SiLU) — equal except where a statistics difference of a few fp32 ulps flips a rounding, bounded
at one ulp of the output on a small fraction — and against the fp32 oracle.
Residual block: against the reference's block composed from its real, importable parts (InflatedCausalConv3d through cuDNN,
causal_norm_wrapper, nn.SiLU; diffusers' base class is not in the image) in fp16 over the VAE's temporal slices, both measured
against the fp32 oracle (oracle/vae_block.py)."""
import pytest
import torch
import torch.nn.functional as F
from torch import nn

from oracle import refstage
from oracle import vae_block as ovb
from vae_standins import DecoderTop, ResnetBlock3D, Upsample3D

pytestmark = pytest.mark.gpu


def _ref_gn_silu(x_thwc, w, b, groups, eps, silu=True):
    """ATen group_norm per frame + silu in the tensor dtype; x [T, H, W, C]."""
    y = F.group_norm(x_thwc.permute(0, 3, 1, 2).contiguous(), groups, w, b, eps)
    if silu:
        y = F.silu(y)
    return y.permute(0, 2, 3, 1)


@pytest.mark.parametrize("C", [64, 128, 256, 512])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16], ids=["fp16", "bf16"])
@pytest.mark.parametrize("shape", [(3, 37, 23), (1, 64, 64), (2, 135, 240)], ids=["3x37x23", "1x64x64", "2x135x240"])
def test_groupnorm_silu_matches_aten(svr2, cuda, C, dtype, shape):
    T, H, W = shape
    g = torch.Generator(device="cuda").manual_seed(C + T)
    x = (torch.randn(T, H, W, C, generator=g, device=cuda) * 1.7 + 0.4 * torch.randn(1, 1, 1, C, generator=g, device=cuda)).to(dtype)
    w = (1 + 0.3 * torch.randn(C, generator=g, device=cuda)).to(dtype)
    b = (0.2 * torch.randn(C, generator=g, device=cuda)).to(dtype)
    ulp = 2.0 ** (-10 if dtype == torch.float16 else -7)
    # the normalisation alone: ATen's output, except where a statistics difference of a few fp32 ulps flips the rounding
    got = svr2.vae_block.groupnorm_silu(x, w, b, 32, 1e-6, silu=False)
    want = _ref_gn_silu(x, w, b, 32, 1e-6, silu=False)
    exact = _ref_gn_silu(x.float(), w.float(), b.float(), 32, 1e-6, silu=False)
    torch.cuda.synchronize()
    assert got.shape == want.shape and got.dtype == dtype
    d = (got.float() - want.float()).abs()
    # never more than one flipped rounding; where a * x + b cancels to ~0 the fp32 evaluation error (a few 1e-7 of |a x|) is
    # what is left, so the bound carries an absolute term on the operand scale
    assert bool((d <= ulp * 1.01 * want.float().abs() + 1e-6 * (1 + x.float().abs())).all())
    assert float((d > 0).float().mean()) < 0.05, float((d > 0).float().mean())      # a flipped rounding of one group's rstd at most
    e_got, e_want = float((got.float() - exact).abs().max()), float((want.float() - exact).abs().max())
    assert e_got <= 1.05 * e_want + 1e-6, (e_got, e_want)
    # the fused activation: bit-equal to ATen's silu of the normalised tensor (same fp32 formula, same rounding point)
    fused = svr2.vae_block.groupnorm_silu(x, w, b, 32, 1e-6, silu=True)
    assert torch.equal(fused, F.silu(got))
    # and end to end against the reference's two ops, on the scale of the activation's input (silu is steep in relative terms
    # for large negative inputs: one flipped rounding of y = -10 moves silu(y) by 0.8 %)
    want_act = F.silu(want)
    lim = ulp * 1.01 * (want_act.float().abs() + want.float().abs()) + 1e-6 * (1 + x.float().abs())
    assert bool(((fused.float() - want_act.float()).abs() <= lim).all())


def test_groupnorm_variants(svr2, cuda):
    g = torch.Generator(device="cuda").manual_seed(9)
    x = (torch.randn(2, 40, 56, 128, generator=g, device=cuda) * 0.5 + 60.0).half()      # |mean| >> std: shifted sums must hold
    w = torch.ones(128, device=cuda).half()
    b = torch.zeros(128, device=cuda).half()
    got = svr2.vae_block.groupnorm_silu(x, w, b, 32, 1e-6, silu=False)
    # ATen rounds the mean to fp16 (60 +- 0.03) before subtracting it, so "exact" is not the target here: ATen's own output is
    want = _ref_gn_silu(x, w, b, 32, 1e-6, silu=False)
    d = (got.float() - want.float()).abs()
    assert bool((d <= 2.0 ** -10 * 1.01 * want.float().abs() + 1e-6 * (1 + x.float().abs())).all())
    assert float((d > 0).float().mean()) < 0.05
    exact = _ref_gn_silu(x.float(), None, None, 32, 1e-6, silu=False)
    assert float((got.float() - exact).abs().max()) <= 1.05 * float((want.float() - exact).abs().max())
    # no affine, no activation; in place
    plain = svr2.vae_block.groupnorm_silu(x, None, None, 32, 1e-6, silu=False)
    assert torch.equal(plain, got)
    y = x.clone()
    svr2.vae_block.groupnorm_silu(y, w, b, 32, 1e-6, silu=True, out=y)
    assert torch.equal(y, svr2.vae_block.groupnorm_silu(x, w, b, 32, 1e-6))
    # deterministic: the partials are merged in a fixed order
    assert torch.equal(svr2.vae_block.groupnorm_silu(x, w, b, 32, 1e-6), svr2.vae_block.groupnorm_silu(x, w, b, 32, 1e-6))
    with pytest.raises(RuntimeError, match="channels per group"):
        svr2.vae_block.groupnorm_silu(torch.zeros(1, 4, 4, 96, dtype=torch.float16, device=cuda), None, None, 32, 1e-6)
    with pytest.raises(NotImplementedError):
        svr2.vae_block.groupnorm_silu(x.float(), None, None, 32, 1e-6)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16], ids=["fp16", "bf16"])
def test_fused_silu_is_atens_silu_bit_for_bit(svr2, cuda, dtype):
    """The activation takes a fast path (ex2 / rcp approximations) wherever its rounding provably equals the exact formula's and
    the exact path elsewhere: 67 M values covering the activation's whole useful input range (|y| up to ~25), bit-equal."""
    g = torch.Generator(device="cuda").manual_seed(4)
    x = torch.randn(2, 512, 512, 128, generator=g, device=cuda).to(dtype)
    w = torch.linspace(0.05, 5.0, 128, device=cuda).to(dtype)          # per-channel gain: outputs from ~0.05 sigma to 5 sigma wide
    b = torch.linspace(-2.0, 2.0, 128, device=cuda).to(dtype)
    y = svr2.vae_block.groupnorm_silu(x, w, b, 32, 1e-6, silu=False)
    fused = svr2.vae_block.groupnorm_silu(x, w, b, 32, 1e-6, silu=True)
    assert float(y.float().abs().max()) > 20
    assert torch.equal(fused, F.silu(y))


@pytest.mark.parametrize("ci,co,full", [(128, 128, True), (256, 128, True), (128, 256, False)], ids=["128", "256to128", "128to256_half"])
def test_resnet_block_matches_reference_parts(svr2, cuda, ci, co, full):
    if not refstage.available():
        pytest.skip("reference not staged")
    refstage.activate()
    from src.models.video_vae_v3.modules.causal_inflation_lib import causal_norm_wrapper, init_causal_conv3d
    from src.models.video_vae_v3.modules.types import MemoryState
    torch.manual_seed(ci + co)
    ref = ResnetBlock3D(ci, co, full).eval().cuda().half()
    mine = svr2.vae_block.ResnetBlock3DB200.from_reference(ref)
    params = {k: v.float() for k, v in ref.state_dict().items()}
    x = (torch.randn(1, ci, 9, 40, 56, device=cuda) * 1.5).half().contiguous(memory_format=torch.channels_last_3d)
    # the fp32 oracle must not run its convolutions in TF32 (cuDNN's default for fp32 inputs)
    with torch.no_grad(), torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
        outs_r, outs_m, outs_o, mem = [], [], [], None
        for i, (a, b) in enumerate(((0, 4), (4, 5), (5, 9))):
            st = MemoryState.INITIALIZING if i == 0 else MemoryState.ACTIVE
            outs_r.append(ref(x[:, :, a:b].contiguous(), memory_state=st))
            outs_m.append(mine(x[:, :, a:b], None, memory_state=st))
            o, mem = ovb.resnet_block3d(x[:, :, a:b].float(), params, memory=mem)
            outs_o.append(o)
        r, m, o = torch.cat(outs_r, 2).float(), torch.cat(outs_m, 2).float(), torch.cat(outs_o, 2)
        one_shot, _ = ovb.resnet_block3d(x.float(), params)
        assert float((o - one_shot).abs().max()) < 1e-4                      # slicing does not change the function
        e_ref, e_mine = float((r - o).abs().max()), float((m - o).abs().max())
        rms_ref, rms_mine = float((r - o).pow(2).mean().sqrt()), float((m - o).pow(2).mean().sqrt())
        assert m.shape == r.shape
        assert rms_mine <= 1.1 * rms_ref + 1e-6, ("rms", rms_mine, rms_ref, e_mine, e_ref)
        assert e_mine <= 1.5 * e_ref + 1e-3, ("max", e_mine, e_ref, rms_mine, rms_ref)
        # one-shot call with the memory off, and its equality with the streamed result
        whole = mine(x, None, memory_state=MemoryState.DISABLED)
        assert torch.equal(whole, torch.cat(outs_m, 2))
    # the installer finds the block by its class name and swaps it in place
    up_conv = init_causal_conv3d(co, co, 3, padding=1, inflation_mode="tail").cuda().half()          # an up-sampler's convolution
    strided = init_causal_conv3d(co, co, 3, stride=2, padding=1, inflation_mode="tail").cuda().half()  # a down-sampler's: not mirrored
    holder = nn.Sequential(ref, nn.Sequential(up_conv, strided))
    assert svr2.vae_block.install_vae_blocks(holder) == (1, 1)
    assert isinstance(holder[0], svr2.vae_block.ResnetBlock3DB200) and isinstance(holder[1][0], svr2.vae_conv.CausalConv3dB200)
    assert holder[1][1] is strided


@pytest.mark.parametrize("co", [128, 256, 512])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16], ids=["fp16", "bf16"])
def test_conv_epilogue_statistics_feed_the_groupnorm(svr2, cuda, co, dtype):
    """The convolution's epilogue gathers the (count, mean, M2) partials of its output; merged, they are the per-frame group
    statistics ATen computes, and the GroupNorm that consumes them equals the one that makes its own pass."""
    g = torch.Generator(device="cuda").manual_seed(co)
    T, H, W, ci = 3, 37, 50, 128                      # ragged tiles: rows / columns past the frame's edge must not be counted
    x = (torch.randn(T, H, W, ci, generator=g, device=cuda) + 0.5).to(dtype)
    w = (torch.randn(co, ci, 3, 3, 3, generator=g, device=cuda) * (27 * ci) ** -0.5).to(dtype)
    b = (torch.randn(co, generator=g, device=cuda) * 2 + 3).to(dtype)               # |mean| > std: the pivot matters
    resid = torch.randn(T, H, W, co, generator=g, device=cuda).to(dtype)
    wp = svr2.vae_conv.pack_conv_weight(w)
    slots = svr2.vae_conv.gn_slots()
    part = torch.full((T, slots, 32, 3), float("nan"), device=cuda)                 # the call zeroes what it does not write
    y = svr2.vae_conv.causal_conv3d(x, wp, b, None, resid, gn_partials=part)
    assert torch.equal(y, svr2.vae_conv.causal_conv3d(x, wp, b, None, resid))       # gathering changes nothing in the output
    n, mean, m2 = part[..., 0].double(), part[..., 1].double(), part[..., 2].double()
    assert bool(torch.isfinite(part).all())
    cnt = n.sum(1)                                                                  # [T, 32]
    assert bool((cnt == H * W * (co // 32)).all())
    gmean = (n * mean).sum(1) / cnt
    gm2 = (m2 + n * (mean - gmean[:, None]) ** 2).sum(1)
    yg = y.double().reshape(T, H * W, 32, co // 32)
    assert float((gmean - yg.mean((1, 3))).abs().max()) < 1e-5
    var_ref = yg.var((1, 3), unbiased=False)
    assert float(((gm2 / cnt) / var_ref - 1).abs().max()) < 1e-5
    gw = (1 + 0.2 * torch.randn(co, generator=g, device=cuda)).to(dtype)
    gb = (0.2 * torch.randn(co, generator=g, device=cuda)).to(dtype)
    own = svr2.vae_block.groupnorm_silu(y, gw, gb, 32, 1e-6)
    stats = svr2.vae_conv.GroupStats(part, slots, y)
    fed = svr2.vae_block.groupnorm_silu(y, gw, gb, 32, 1e-6, stats=stats)
    d = (own.float() - fed.float()).abs()
    ulp = 2.0 ** (-10 if dtype == torch.float16 else -7)
    assert float((d > 0).float().mean()) < 0.05
    assert bool((d <= ulp * 1.01 * (own.float().abs() + 1) + 1e-6).all())
    y.add_(1)                                                                       # written since: the statistics no longer apply
    assert not stats.valid_for(y)


def test_block_chain_reuses_epilogue_statistics(svr2, cuda):
    if not refstage.available():
        pytest.skip("reference not staged")
    refstage.activate()
    from src.models.video_vae_v3.modules.causal_inflation_lib import causal_norm_wrapper, init_causal_conv3d
    from src.models.video_vae_v3.modules.types import MemoryState
    torch.manual_seed(1)
    refs = [ResnetBlock3D(256, 128).eval().cuda().half(), ResnetBlock3D(128, 128).eval().cuda().half()]
    mine = [svr2.vae_block.ResnetBlock3DB200.from_reference(r) for r in refs]
    ident = nn.Identity()
    x = torch.randn(1, 256, 3, 40, 56, device=cuda).half().contiguous(memory_format=torch.channels_last_3d)
    with torch.no_grad():
        n0 = svr2.ops.launch_count()
        h = ident(mine[0](x, None, memory_state=MemoryState.DISABLED))
        n1 = svr2.ops.launch_count()
        out = mine[1](h, None, memory_state=MemoryState.DISABLED)
        n2 = svr2.ops.launch_count()
        # block 0: norm1 makes its own statistics pass (2 kernels), norm2 is fed by conv1 (1), three convolutions -> 6;
        # block 1: both norms are fed (1 + 1), two convolutions -> 4
        assert (n1 - n0, n2 - n1) == (6, 4), (n1 - n0, n2 - n1)
        r = refs[1](refs[0](x.contiguous(), memory_state=MemoryState.DISABLED), memory_state=MemoryState.DISABLED)
        params = [{k: v.float() for k, v in rr.state_dict().items()} for rr in refs]
        with torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
            o, _ = ovb.resnet_block3d(ovb.resnet_block3d(x.float(), params[0])[0], params[1])
        rms_ref, rms_mine = float((r.float() - o).pow(2).mean().sqrt()), float((out.float() - o).pow(2).mean().sqrt())
        assert rms_mine <= 1.1 * rms_ref + 1e-6, (rms_mine, rms_ref)


@pytest.mark.parametrize("slicing", [False, True], ids=["whole", "sliced_upsampler"])
def test_decoder_top_half_with_installed_modules(svr2, cuda, slicing):
    """The decoder's upper half (two up blocks, spatial up-sampler, output norm and convolution) built from the reference's real
    parts, against a copy with `install_vae_blocks` applied, over two temporal slices with the memory carried: 6 residual blocks,
    the up-sampler's expansion and (list-fed) convolution and the padded conv_out all run on this library's kernels."""
    if not refstage.available():
        pytest.skip("reference not staged")
    import copy
    from vae_standins import reference_parts
    _, _, _, MemoryState = reference_parts()
    torch.manual_seed(7)
    ref = DecoderTop(128, 128, 128, slicing=slicing).eval().cuda().half()      # narrow: the test is about wiring, not width
    mine = copy.deepcopy(ref)
    n_blocks, n_convs = svr2.vae_block.install_vae_blocks(mine)
    mine.use_rebound_norm = True
    assert (n_blocks, n_convs) == (6, 3)                                      # expansion, up-sampler convolution, conv_out
    x = torch.randn(1, 128, 6, 24, 40, device=cuda).half()
    with torch.no_grad():
        outs_r, outs_m = [], []
        for i, (a, b) in enumerate(((0, 4), (4, 6))):
            st = MemoryState.INITIALIZING if i == 0 else MemoryState.ACTIVE
            outs_r.append(ref(x[:, :, a:b].contiguous(), memory_state=st))
            outs_m.append(mine(x[:, :, a:b].contiguous(), memory_state=st))
        r, m = torch.cat(outs_r, 2).float(), torch.cat(outs_m, 2).float()
    assert m.shape == r.shape == (1, 3, 6, 48, 80)
    # fp16 against fp16 through 14 convolutions: compare on the output's scale
    scale = float(r.abs().max())
    assert float((r - m).abs().max()) <= 0.02 * scale, (float((r - m).abs().max()), scale)
    assert float((r - m).pow(2).mean().sqrt()) <= 2e-3 * scale


@pytest.mark.parametrize("temporal_up,slicing", [(False, False), (True, False), (True, True), (False, True)],
                         ids=["spatial", "spatial+temporal", "spatial+temporal_sliced", "spatial_sliced"])
def test_upsampler_mirror(svr2, cuda, temporal_up, slicing):
    """Upsample3D from the reference's parts (einops rearrange, remove_head, the real causal convolution fed a tensor or a list)
    against the mirror (1 x 1 x 1 GEMM, shuffle kernel with the head removal folded in, implicit-GEMM convolution) over two
    temporal slices; the shuffle itself is bit-equal to einops."""
    if not refstage.available():
        pytest.skip("reference not staged")
    from einops import rearrange
    from vae_standins import reference_parts
    _, _, remove_head, MemoryState = reference_parts()
    torch.manual_seed(3)
    ref = Upsample3D(128, temporal_up=temporal_up, spatial_up=True, slicing=slicing).eval().cuda().half()
    nn.init.normal_(ref.upscale_conv.weight, 0, 128 ** -0.5)          # not the identity initialisation: every (x, y, z) plane differs
    mine = svr2.vae_block.Upsample3DB200.from_reference(ref)
    z = 2 if temporal_up else 1
    t = torch.randn(3, 10, 12, 4 * z * 128, device=cuda).half()
    want = rearrange(t.permute(3, 0, 1, 2)[None], "b (x y z c) f h w -> b c (f z) (h x) (w y)", x=2, y=2, z=z)
    assert torch.equal(svr2.vae_block.pixel_shuffle3d(t, 128, 2, z).permute(3, 0, 1, 2)[None], want)
    if temporal_up:
        assert torch.equal(svr2.vae_block.pixel_shuffle3d(t, 128, 2, z, True).permute(3, 0, 1, 2)[None], remove_head(want))
    x = torch.randn(1, 128, 7, 20, 24, device=cuda).half()
    with torch.no_grad():
        outs_r, outs_m = [], []
        for i, (a, b) in enumerate(((0, 4), (4, 7))):
            st = MemoryState.INITIALIZING if i == 0 else MemoryState.ACTIVE
            outs_r.append(ref(x[:, :, a:b].contiguous(), memory_state=st))
            outs_m.append(mine(x[:, :, a:b].contiguous(), memory_state=st))
        r, m = torch.cat(outs_r, 2).float(), torch.cat(outs_m, 2).float()
    assert m.shape == r.shape == (1, 128, 7 * z - (1 if temporal_up else 0), 40, 48)
    scale = float(r.abs().max())
    assert float((r - m).abs().max()) <= 2.0 ** -8 * scale, (float((r - m).abs().max()), scale)
    if mine.conv.kernel_size[0] > 1:
        assert torch.equal(mine.conv.memory.permute(0, 4, 1, 2, 3).float(), ref.conv.memory.float()) or \
            float((mine.conv.memory.permute(0, 4, 1, 2, 3).float() - ref.conv.memory.float()).abs().max()) <= 2.0 ** -9 * scale


def test_whole_decoder_standin_psnr(svr2, cuda):
    """Decoder3D.forward from the reference's parts (conv_in 16 -> C, mid resnets, four up blocks with temporal + spatial up-samplers,
    output norm, conv_out C -> 3; narrow) with everything `install_vae_blocks` covers swapped in, streamed over three temporal slices:
    decoded pixels against the reference's modules >= 45 dB on the [-1, 1] range (SURVEY §8c's bar, here on fp16 vs fp16)."""
    if not refstage.available():
        pytest.skip("reference not staged")
    import copy
    from vae_standins import Decoder3D, reference_parts
    _, _, _, MemoryState = reference_parts()
    torch.manual_seed(11)
    ref = Decoder3D(16, (128, 128, 128, 128), attention=True).eval().cuda().half()
    mine = copy.deepcopy(ref)
    n_blocks, n_convs = svr2.vae_block.install_vae_blocks(mine)
    mine.use_rebound_norm = True
    assert (n_blocks, n_convs) == (14, 9)            # 2 mid + 12 up blocks; conv_in, mid attention, 3 up-samplers x 2, conv_out
    assert not any(type(m).__name__ in ("Attention", "InflatedCausalConv3d", "ResnetBlock3D", "Upsample3D") for m in mine.modules())
    zs = [torch.randn(1, 16, t, 9, 12, device=cuda).half() for t in (1, 2, 1)]
    with torch.no_grad():
        outs_r, outs_m = [], []
        for i, z in enumerate(zs):
            st = MemoryState.INITIALIZING if i == 0 else MemoryState.ACTIVE
            outs_r.append(ref(z, memory_state=st))
            outs_m.append(mine(z, memory_state=st))
    r, m = torch.cat(outs_r, 2).float(), torch.cat(outs_m, 2).float()
    assert m.shape == r.shape == (1, 3, 1 + 8 + 4, 72, 96)           # 4n + 1 frames, 8 x the latent's side
    r = r / r.abs().max()                                             # random weights: bring the output to the pixel range
    m = m / torch.cat(outs_r, 2).float().abs().max()
    psnr = float(10 * torch.log10(4.0 / (r - m).pow(2).mean()))
    assert psnr >= 45.0, psnr


@pytest.mark.parametrize("temporal", [False, True], ids=["spatial", "spatial+temporal"])
@pytest.mark.parametrize("c,hw", [(128, (37, 50)), (256, (40, 56)), (512, (17, 33))], ids=["128ch_odd", "256ch_even", "512ch_small"])
def test_downsampler_mirror(svr2, cuda, temporal, c, hw):
    """The encoder's Downsample3D from the reference's parts (F.pad + the real strided InflatedCausalConv3d through cuDNN) against the
    mirror (one strided TMA box per tap, the padding as zero fill) over three temporal slices, against the fp32 oracle."""
    if not refstage.available():
        pytest.skip("reference not staged")
    from oracle import vae_conv as ovc
    from vae_standins import Downsample3D, reference_parts
    _, _, _, MemoryState = reference_parts()
    torch.manual_seed(c)
    ref = Downsample3D(c, temporal_down=temporal).eval().cuda().half()
    holder = nn.Sequential(ref)
    assert svr2.vae_block.install_vae_blocks(holder) == (0, 1)
    mine = holder[0]
    assert isinstance(mine, svr2.vae_block.Downsample3DB200)
    H, W = hw
    x = torch.randn(1, c, 11, H, W, device=cuda).half()
    with torch.no_grad():
        mem, outs_r, outs_m, outs_o = None, [], [], []
        for i, (a, b) in enumerate(((0, 5), (5, 9), (9, 11))):
            st = MemoryState.INITIALIZING if i == 0 else MemoryState.ACTIVE
            outs_r.append(ref(x[:, :, a:b].contiguous(), memory_state=st))
            outs_m.append(mine(x[:, :, a:b].contiguous(), memory_state=st))
            o, mem = ovc.causal_conv3d_down(x[:, :, a:b].float(), ref.conv.weight.float(), ref.conv.bias.float(), mem)
            outs_o.append(o)
        r, m, o = torch.cat(outs_r, 2).float(), torch.cat(outs_m, 2).float(), torch.cat(outs_o, 2)
    assert m.shape == r.shape == o.shape == (1, c, 6 if temporal else 11, (H - 2) // 2 + 1, (W - 2) // 2 + 1)
    e_ref, e_mine = float((r - o).abs().max()), float((m - o).abs().max())
    assert e_mine <= max(1.05 * e_ref, 2.0 ** -11 * float(o.abs().max()) * 1.01), (e_mine, e_ref)


def test_causal_norm_wrapper_rebinding(svr2, cuda):
    """patch_causal_norm_wrapper rebinds the reference module's global; the replacement equals the original on a 5-D fp16 GroupNorm
    (within ATen-vs-ours rounding flips), defers everything else to it, and is idempotent."""
    if not refstage.available():
        pytest.skip("reference not staged")
    refstage.activate()
    import src.models.video_vae_v3.modules.causal_inflation_lib as lib
    svr2.vae_block.patch_causal_norm_wrapper()
    assert svr2.vae_block.patch_causal_norm_wrapper() == 0                     # already rebound
    fn, orig = lib.causal_norm_wrapper, lib.causal_norm_wrapper._svr2_original
    assert fn is not orig
    gn = nn.GroupNorm(32, 128, eps=1e-6).cuda().half()
    nn.init.normal_(gn.weight, 1.0, 0.2)
    nn.init.normal_(gn.bias, 0.0, 0.2)
    x = torch.randn(2, 128, 3, 20, 24, device=cuda).half()
    with torch.no_grad():
        a, b = fn(gn, x), orig(gn, x)
        assert a.shape == b.shape
        d = (a.float() - b.float()).abs()
        assert float((d > 0).float().mean()) < 0.05 and bool((d <= 2.0 ** -10 * 1.01 * b.float().abs() + 1e-6 * (1 + x.float().abs())).all())
        # everything else goes to the original: fp32 input, LayerNorm
        x32 = x.float()
        gn32 = nn.GroupNorm(32, 128).cuda()
        assert torch.equal(fn(gn32, x32), orig(gn32, x32))
        ln = nn.LayerNorm(128).cuda().half()
        assert torch.equal(fn(ln, x), orig(ln, x))


def test_memories_are_cleared_after_the_last_slice(svr2, cuda):
    """VideoAutoencoderKL.slicing_decode clears `InflatedCausalConv3d.memory` by isinstance after the last slice (attn_video_vae.py:
    1285-1289); install_vae_blocks wraps the method so that the swapped convolutions drop theirs too."""
    if not refstage.available():
        pytest.skip("reference not staged")
    from vae_standins import reference_parts
    _, _, _, MemoryState = reference_parts()

    class Vae(nn.Module):          # the two attributes of VideoAutoencoderKL that matter here
        def __init__(self):
            super().__init__()
            self.decoder = nn.Sequential()
            self.decoder.add_module("block", ResnetBlock3D(128, 128).eval().cuda().half())

        def slicing_decode(self, z):
            outs = [self.decoder.block(z[:, :, :3], memory_state=MemoryState.INITIALIZING),
                    self.decoder.block(z[:, :, 3:], memory_state=MemoryState.ACTIVE)]
            return torch.cat(outs, 2)

    vae = Vae()
    assert svr2.vae_block.install_vae_blocks(vae) == (1, 0)
    z = torch.randn(1, 128, 5, 16, 16, device=cuda).half()
    with torch.no_grad():
        out = vae.slicing_decode(z)
    assert out.shape == (1, 128, 5, 16, 16)
    assert all(m.memory is None for m in vae.modules() if isinstance(m, svr2.vae_conv.CausalConv3dB200))
    with torch.no_grad():          # while slicing, the memory is there
        vae.decoder.block(z[:, :, :3], memory_state=MemoryState.INITIALIZING)
    assert svr2.vae_block.clear_memory(vae) == 2


def test_block_under_autocast_like_the_pipeline(svr2, cuda):
    """The pipeline wraps the VAE in torch.autocast when its inputs (bf16 compute dtype) and the VAE's weights (fp16) differ
    (src/core/infer.py:156-158, 232-234): the swapped block then computes in the autocast dtype like the reference's convolutions,
    with weights cast once."""
    if not refstage.available():
        pytest.skip("reference not staged")
    from vae_standins import reference_parts
    _, _, _, MemoryState = reference_parts()
    torch.manual_seed(5)
    ref = ResnetBlock3D(256, 128).eval().cuda().half()
    mine = svr2.vae_block.ResnetBlock3DB200.from_reference(ref)
    x = torch.randn(1, 256, 5, 24, 40, device=cuda).bfloat16()
    params = {k: v.float() for k, v in ref.state_dict().items()}
    with torch.no_grad():
        with torch.autocast("cuda", torch.bfloat16):
            r = torch.cat([ref(x[:, :, :3], memory_state=MemoryState.INITIALIZING), ref(x[:, :, 3:], memory_state=MemoryState.ACTIVE)], 2)
            m = torch.cat([mine(x[:, :, :3], None, memory_state=MemoryState.INITIALIZING), mine(x[:, :, 3:], None, memory_state=MemoryState.ACTIVE)], 2)
        with torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
            o, _ = ovb.resnet_block3d(x.float(), params)
    assert m.dtype == torch.bfloat16 and m.shape == r.shape
    rms_ref, rms_mine = float((r.float() - o).pow(2).mean().sqrt()), float((m.float() - o).pow(2).mean().sqrt())
    # the reference keeps GroupNorm and SiLU in fp32 under autocast; the fused pass rounds twice to bf16 in between
    assert rms_mine <= 1.5 * rms_ref + 1e-6, (rms_mine, rms_ref)
    assert mine.conv1.memory.dtype == torch.bfloat16


def test_row_softmax(svr2, cuda):
    for dtype in (torch.float16, torch.bfloat16):
        for rows, n, pitch in ((37, 1000, 1024), (5, 32400, 32512), (3, 8, 8)):
            g = torch.Generator(device="cuda").manual_seed(n)
            s = (torch.randn(rows, pitch, generator=g, device=cuda) * 4).to(dtype)
            want = torch.zeros(rows, pitch, device=cuda)
            want[:, :n] = torch.softmax(s[:, :n].float(), -1)
            from seedvr2_b200 import vae_attn
            got = vae_attn.row_softmax_(s.clone(), n)
            assert float(got[:, n:].abs().max()) == 0.0 if pitch > n else True
            ulp = 2.0 ** (-10 if dtype == torch.float16 else -7)
            assert bool(((got.float() - want).abs() <= ulp * want + 1e-7).all())


@pytest.mark.parametrize("hw", [(24, 40), (17, 23), (135, 240)], ids=["24x40", "17x23_unaligned", "1080p_latent"])
def test_mid_block_attention_mirror(svr2, cuda, hw):
    """AttentionB200 (two GEMMs on the convolution kernel around the row softmax) against the stand-in of diffusers' layer (torch
    SDPA), fp16, called the way UNetMidBlock3D does — on the NHWC-strided "(b f) c h w" view of a channels-last clip; both against
    the layer evaluated in fp32.  Parity unpinned (diffusers is absent): the stand-in restates AttnProcessor2_0."""
    from vae_standins import Attention
    from seedvr2_b200 import vae_attn
    torch.manual_seed(2)
    H, W = hw
    C, F_ = 512, 2
    ref = Attention(C).eval().cuda().half()
    for lin in (ref.to_q, ref.to_k):
        nn.init.normal_(lin.weight, 0, 2.0 * C ** -0.5)          # logits of a few units: a softmax that is not flat
    holder = nn.Sequential(ref)
    assert svr2.vae_block.install_vae_blocks(holder) == (0, 1)
    mine = holder[0]
    assert isinstance(mine, vae_attn.AttentionB200)
    clip = torch.randn(1, F_, H, W, C, device=cuda).half()                                   # channels-last clip [b, f, h, w, c]
    x = clip.permute(0, 4, 1, 2, 3).permute(0, 2, 1, 3, 4).reshape(F_, C, H, W)             # "(b f) c h w" as a strided view
    assert x.data_ptr() == clip.data_ptr()
    with torch.no_grad():
        r = ref(x.contiguous())
        m = mine(x, temb=None)
        exact = Attention(C).eval().cuda().float()
        exact.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
        o = exact(x.float())
    assert m.shape == r.shape == (F_, C, H, W) and m.dtype == torch.float16
    e_ref, e_mine = float((r.float() - o).abs().max()), float((m.float() - o).abs().max())
    rms_ref, rms_mine = float((r.float() - o).pow(2).mean().sqrt()), float((m.float() - o).pow(2).mean().sqrt())
    # S is rounded to fp16 before the softmax here (SDPA keeps it in fp32): allow twice the reference's own error
    assert rms_mine <= 2.0 * rms_ref + 1e-5, (rms_mine, rms_ref)
    assert e_mine <= 3.0 * e_ref + 2e-3, (e_mine, e_ref)
