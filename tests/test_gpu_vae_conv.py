"""svr2_causal_conv3d_fwd (csrc/conv3d.cu) through comfyui-seedvr2_b200/vae_conv.py: the VAE's causal 3-D convolution.

Checked against (1) the oracle (oracle/vae_conv.py) evaluated in fp32 on the same fp16 / bf16 inputs, with a bound derived from
the two rounding sources — the output's own rounding (half an ulp of the result) and the fp32 accumulation over K = 27 C_in terms (the tensor
pipe aligns and truncates partial sums: measured up to 1e-6 of sum |w||x| at K = 13,824; bound 3e-6); (2) the reference's own InflatedCausalConv3d run on the same GPU through cuDNN, over the temporal slices and
memory states of the VAE's slicing (staged copy, baseline/_ref); (3) at 1080p, a crop of the result against the oracle on the
crop's receptive field."""
import pytest
import torch
import torch.nn.functional as F

from oracle import refstage
from oracle import vae_conv as ovc

pytestmark = pytest.mark.gpu

ULP = {torch.float16: 2.0 ** -11, torch.bfloat16: 2.0 ** -8}      # half an ulp, relative


def _bound(x, w, b, dtype, mem=None):
    """|ours - fp32 oracle| <= half an ulp of the rounded result (+ the bias add's) + accumulation-order noise."""
    mag, _ = ovc.causal_conv3d(x.float().abs(), w.float().abs(), None if b is None else b.float().abs(),
                               None if mem is None else mem.float().abs())
    return lambda ref: ULP[dtype] * 1.01 * ref.abs() + 3e-6 * mag + 1e-7


def _make(ci, co, kt, ks, T, H, W, dtype, seed=0, bias=True):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(1, ci, T, H, W, generator=g, device="cuda").to(dtype).contiguous(memory_format=torch.channels_last_3d)
    w = (torch.randn(co, ci, kt, ks, ks, generator=g, device="cuda") * (ci * kt * ks * ks) ** -0.5).to(dtype)
    b = (torch.randn(co, generator=g, device="cuda") * 0.5).to(dtype) if bias else None
    return x, w, b


CASES = [
    # ci, co, kt, ks, T, H, W
    (128, 128, 3, 3, 3, 24, 40),        # ragged: H, W not multiples of the 16 x 16 patch; N tile 128
    (128, 128, 3, 3, 1, 16, 16),        # one frame: the whole causal head is the replicated first frame
    (256, 256, 3, 3, 2, 33, 17),        # N tile 256, odd extents
    (128, 256, 3, 3, 2, 20, 36),        # widening conv1 of a block
    (512, 512, 3, 3, 2, 9, 50),         # deepest level, fewer rows than one patch
    (256, 128, 1, 3, 4, 18, 30),        # time_receptive_field "half": (1, 3, 3)
    (128, 256, 1, 1, 3, 21, 19),        # conv_shortcut: 1 x 1 x 1
    (64, 128, 3, 3, 2, 16, 32),         # one 64-channel K chunk per tap
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "x".join(map(str, c)))
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16], ids=["fp16", "bf16"])
def test_conv_matches_oracle(svr2, cuda, case, dtype):
    ci, co, kt, ks, T, H, W = case
    x, w, b = _make(ci, co, kt, ks, T, H, W, dtype, seed=sum(case))
    ref, _ = ovc.causal_conv3d(x.float(), w.float(), b.float())
    conv = svr2.vae_conv.CausalConv3dB200(w, b)
    got = conv(x, "DISABLED")
    torch.cuda.synchronize()
    assert got.shape == ref.shape and got.dtype == dtype
    assert got.is_contiguous(memory_format=torch.channels_last_3d)
    d = (got.float() - ref).abs()
    lim = _bound(x, w, b, dtype)(ref)
    assert bool((d <= lim).all()), f"max excess {float((d - lim).max()):.3e} at |ref| {float(ref.abs().max()):.2f}"


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16], ids=["fp16", "bf16"])
def test_head_bias_and_residual(svr2, cuda, dtype):
    ci, co, T, H, W = 128, 128, 3, 20, 24
    x, w, b = _make(ci, co, 3, 3, T, H, W, dtype, seed=11)
    g = torch.Generator(device="cuda").manual_seed(12)
    head = torch.randn(2, H, W, ci, generator=g, device="cuda").to(dtype)
    resid = torch.randn(T, H, W, co, generator=g, device="cuda").to(dtype)
    wp = svr2.vae_conv.pack_conv_weight(w)
    assert torch.equal(wp, w.permute(0, 2, 3, 4, 1).contiguous())
    xt = x[0].permute(1, 2, 3, 0)
    assert xt.is_contiguous()
    mem = head.permute(3, 0, 1, 2)[None]
    ref, _ = ovc.causal_conv3d(x.float(), w.float(), None, mem.float())
    plain = svr2.vae_conv.causal_conv3d(xt, wp, None, head)                     # no bias
    d = (plain.permute(3, 0, 1, 2)[None].float() - ref).abs()
    assert bool((d <= _bound(x, w, None, dtype, mem)(ref)).all())
    with_b = svr2.vae_conv.causal_conv3d(xt, wp, b, head)
    fused = svr2.vae_conv.causal_conv3d(xt, wp, b, head, resid)
    # the residual is added to the ROUNDED convolution output, like the reference's separate tensor add: bit-exact
    assert torch.equal(fused, with_b + resid)
    # bias enters before the rounding (cuDNN / ATen add it in fp32)
    ref_b = ref + b.float().view(1, -1, 1, 1, 1)
    d = (with_b.permute(3, 0, 1, 2)[None].float() - ref_b).abs()
    assert bool((d <= _bound(x, w, b, dtype, mem)(ref_b)).all())


def test_module_streams_like_reference(svr2, cuda):
    """The reference's InflatedCausalConv3d (fp16, cuDNN) and the mirror over the VAE's temporal slices: outputs within fp16
    rounding of each other, memories bit-equal, and the streamed result equal to the one-shot one."""
    if not refstage.available():
        pytest.skip("reference not staged")
    refstage.activate()
    from src.models.video_vae_v3.modules.causal_inflation_lib import init_causal_conv3d
    from src.models.video_vae_v3.modules.types import MemoryState
    torch.manual_seed(3)
    for kernel, pad, ci, co in ((3, 1, 128, 128), ((1, 3, 3), (0, 1, 1), 128, 256), (3, 1, 256, 256)):
        ref = init_causal_conv3d(ci, co, kernel_size=kernel, stride=1, padding=pad, inflation_mode="tail").eval().cuda().half()
        mine = svr2.vae_conv.CausalConv3dB200.from_reference(ref)
        x = torch.randn(1, ci, 9, 40, 56, device="cuda").half().contiguous(memory_format=torch.channels_last_3d)
        with torch.no_grad():
            outs_r, outs_m = [], []
            for i, (a, b) in enumerate(((0, 4), (4, 5), (5, 9))):
                st = MemoryState.INITIALIZING if i == 0 else MemoryState.ACTIVE
                outs_r.append(ref(x[:, :, a:b].contiguous(), memory_state=st))
                outs_m.append(mine(x[:, :, a:b], st))
                if ref.kernel_size[0] > 1:
                    assert torch.equal(mine.memory.permute(0, 4, 1, 2, 3), ref.memory)
            # the up-sampler hands the convolution a LIST of temporal pieces (attn_video_vae.py:126-133)
            ref.memory = None
            pieces = [x[:, :, :4].contiguous(), x[:, :, 4:].contiguous()]
            lr = ref([p.clone() for p in pieces], memory_state=MemoryState.INITIALIZING)
            lm = mine([x[:, :, :4], x[:, :, 4:]], MemoryState.INITIALIZING)
            assert isinstance(lm, list) and len(lm) == 2 and torch.equal(torch.cat(lm, 2), torch.cat(outs_m, 2))
            assert float((torch.cat(lr, 2).float() - torch.cat(lm, 2).float()).abs().max()) <= 2.0 ** -9 * float(torch.cat(lr, 2).abs().max())
            if ref.kernel_size[0] > 1:
                assert torch.equal(mine.memory.permute(0, 4, 1, 2, 3), ref.memory)
            r, m = torch.cat(outs_r, 2).float(), torch.cat(outs_m, 2).float()
            one_shot = mine(x, MemoryState.DISABLED)
            assert mine.memory is None
            assert torch.equal(torch.cat(outs_m, 2), one_shot)               # slicing changes nothing, bit for bit
            exact, _ = ovc.causal_conv3d(x.float(), ref.weight.float(), ref.bias.float())
            e_ref, e_mine = float((r - exact).abs().max()), float((m - exact).abs().max())
            assert e_mine <= max(1.05 * e_ref, 2.0 ** -11 * float(exact.abs().max()) * 1.01), (kernel, e_mine, e_ref)
            assert float((r - m).abs().max()) <= 2.0 ** -9 * float(exact.abs().max())


def test_1080p_crop_against_oracle(svr2, cuda):
    """Full-size frame (the decoder's top level: 128 channels at 1080 x 1920): crops at the corners, the edges and the centre
    against the oracle on the crop's receptive field."""
    ci = co = 128
    T, H, W = 3, 1080, 1920
    x, w, b = _make(ci, co, 3, 3, T, H, W, torch.float16, seed=21)
    conv = svr2.vae_conv.CausalConv3dB200(w, b)
    got = conv(x, "DISABLED")
    torch.cuda.synchronize()
    for (h0, w0) in ((0, 0), (0, W - 40), (H - 24, 0), (H - 24, W - 40), (530, 950), (1064, 1000)):
        h1, w1 = min(h0 + 24, H), min(w0 + 40, W)
        ha, wa = max(h0 - 1, 0), max(w0 - 1, 0)
        hb, wb = min(h1 + 1, H), min(w1 + 1, W)
        xc = x[:, :, :, ha:hb, wa:wb].float()
        # zero padding only where the crop touches the frame's edge; elsewhere the halo row / column is real data
        ref, _ = ovc.causal_conv3d(xc, w.float(), b.float())
        ref = ref[:, :, :, h0 - ha:h0 - ha + (h1 - h0), w0 - wa:w0 - wa + (w1 - w0)]
        mag, _ = ovc.causal_conv3d(xc.abs(), w.float().abs(), b.float().abs())
        mag = mag[:, :, :, h0 - ha:h0 - ha + (h1 - h0), w0 - wa:w0 - wa + (w1 - w0)]
        d = (got[:, :, :, h0:h1, w0:w1].float() - ref).abs()
        assert bool((d <= 2.0 ** -11 * 1.01 * ref.abs() + 3e-6 * mag + 1e-7).all()), (h0, w0, float(d.max()))


@pytest.mark.parametrize("ci,co", [(16, 512), (128, 3), (3, 128)], ids=["conv_in", "conv_out", "encoder_in"])
def test_boundary_convolutions_run_with_padded_channels(svr2, cuda, ci, co):
    """conv_in (16 -> 512) and conv_out (128 -> 3) of the decoder, conv_in (3 -> 128) of the encoder: zero-padded to the kernel's
    channel granularity inside the module, same function, NCDHW-contiguous result; the memory bookkeeping is unchanged."""
    if not refstage.available():
        pytest.skip("reference not staged")
    refstage.activate()
    from src.models.video_vae_v3.modules.causal_inflation_lib import init_causal_conv3d
    from src.models.video_vae_v3.modules.types import MemoryState
    torch.manual_seed(ci)
    ref = init_causal_conv3d(ci, co, 3, padding=1, inflation_mode="tail").eval().cuda().half()
    mine = svr2.vae_conv.CausalConv3dB200.from_reference(ref)
    x = torch.randn(1, ci, 6, 33, 40, device="cuda").half()
    with torch.no_grad():
        outs_r, outs_m = [], []
        for i, (a, b) in enumerate(((0, 3), (3, 6))):
            st = MemoryState.INITIALIZING if i == 0 else MemoryState.ACTIVE
            outs_r.append(ref(x[:, :, a:b].contiguous(), memory_state=st))
            outs_m.append(mine(x[:, :, a:b].contiguous(), st))
        r, m = torch.cat(outs_r, 2), torch.cat(outs_m, 2)
        exact, _ = ovc.causal_conv3d(x.float(), ref.weight.float(), ref.bias.float())
    assert m.shape == r.shape == (1, co, 6, 33, 40)
    assert m.is_contiguous() if co % 128 else m.is_contiguous(memory_format=torch.channels_last_3d)
    e_ref, e_mine = float((r.float() - exact).abs().max()), float((m.float() - exact).abs().max())
    assert e_mine <= max(1.05 * e_ref, 2.0 ** -11 * float(exact.abs().max()) * 1.01), (e_mine, e_ref)


def test_argument_validation(svr2, cuda):
    x = torch.zeros(2, 8, 8, 48, dtype=torch.float16, device="cuda")
    w = torch.zeros(128, 3, 3, 3, 48, dtype=torch.float16, device="cuda")
    with pytest.raises(RuntimeError, match="c_in"):
        svr2.vae_conv.causal_conv3d(x, w)
    x = torch.zeros(2, 8, 8, 64, dtype=torch.float16, device="cuda")
    w = torch.zeros(96, 3, 3, 3, 64, dtype=torch.float16, device="cuda")
    with pytest.raises(RuntimeError, match="c_out"):
        svr2.vae_conv.causal_conv3d(x, w)
    with pytest.raises(NotImplementedError):
        svr2.vae_conv.causal_conv3d(x.float(), w.float())
    # a VAE that is still on the host can be installed into (weights re-laid out there); running it needs the GPU
    w_cpu = torch.randn(128, 64, 3, 3, 3).half()
    assert torch.equal(svr2.vae_conv.pack_conv_weight(w_cpu).cuda(), svr2.vae_conv.pack_conv_weight(w_cpu.cuda()))
    conv = svr2.vae_conv.CausalConv3dB200(w_cpu, None)
    with pytest.raises(RuntimeError, match="no CPU path"):
        conv(torch.zeros(1, 64, 2, 8, 8, dtype=torch.float16))
    y = conv.cuda()(torch.zeros(1, 64, 2, 8, 8, dtype=torch.float16, device="cuda"))
    assert y.shape == (1, 128, 2, 8, 8) and float(y.abs().max()) == 0.0


@pytest.mark.parametrize("shape", [(1, 128, 3, 40, 56), (2, 16, 2, 33, 40), (1, 512, 1, 135, 240), (1, 24, 3, 7, 9)],
                         ids=["128ch", "batch2_16ch", "512ch_135x240", "unaligned"])
def test_layout_transposes(svr2, cuda, shape):
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn(*shape, generator=g, device="cuda").half()
    cl = svr2.vae_conv.to_channels_last(x)
    assert cl.is_contiguous() and torch.equal(cl, x.permute(0, 2, 3, 4, 1).contiguous())
    assert svr2.vae_conv.to_channels_last(cl.permute(0, 4, 1, 2, 3)).data_ptr() == cl.data_ptr()        # already channels-last: a view
    back = svr2.vae_conv.to_ncdhw(cl)
    assert back.is_contiguous() and torch.equal(back, x)
    k = min(3, shape[1])
    assert torch.equal(svr2.vae_conv.to_ncdhw(cl, k), x[:, :k].contiguous())                                # conv_out's channel slice


def test_kernels_match_golden_vectors_of_the_real_module(svr2, cuda):
    """tests/golden/vae_conv.pt — outputs of the real InflatedCausalConv3d (fp32 CPU) over streamed slices, stride 1 and the
    down-samplers' strides — against the CUDA kernels in fp16 through the module mirrors, memories carried by the mirrors."""
    import os
    gold = torch.load(os.path.join(os.path.dirname(__file__), "golden", "vae_conv.pt"))
    for name, c in gold.items():
        w, b, x = c["weight"].cuda(), c["bias"].cuda(), c["x"].cuda()
        mine = svr2.vae_conv.CausalConv3dB200(w, b, tuple(c["stride"]), pad_hw=c["pad_hw"])
        for i, ((a, e), y_ref) in enumerate(zip(c["slices"], c["y"])):
            y = mine(x[:, :, a:e].contiguous(), "INITIALIZING" if i == 0 else "ACTIVE")
            y_ref = y_ref.cuda()
            assert y.shape == y_ref.shape, name
            d = (y.float() - y_ref).abs()
            lim = 2.0 ** -11 * 1.01 * y_ref.abs() + 3e-6 * 40 + 1e-7          # sum |w||x| <~ 40 at these sizes
            assert bool((d <= lim).all()), (name, i, float(d.max()))


def test_no_writes_outside_the_outputs(svr2, cuda):
    """Every N3 kernel writes into the middle of a guard-filled buffer (ragged shapes, so tiles hang over every edge): the guards
    before and after the output stay untouched.  (compute-sanitizer is not available on the GPU pool.)"""
    lib = svr2._lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    GUARD = 4096

    def guarded(n_elems, dtype):
        buf = torch.full((n_elems + 2 * GUARD,), 7, dtype=torch.int16, device="cuda").view(dtype) if dtype != torch.float32 else \
            torch.full((n_elems + 2 * GUARD,), 7.0, dtype=torch.float32, device="cuda")
        return buf, buf[GUARD:GUARD + n_elems]

    def check(buf, n_elems, what):
        ref = torch.full((GUARD,), 7, dtype=torch.int16, device="cuda").view(buf.dtype) if buf.dtype != torch.float32 else \
            torch.full((GUARD,), 7.0, device="cuda")
        assert torch.equal(buf[:GUARD], ref) and torch.equal(buf[GUARD + n_elems:], ref), what

    g = torch.Generator(device="cuda").manual_seed(2)
    T, H, W, ci, co = 3, 21, 27, 128, 128
    x = torch.randn(T, H, W, ci, generator=g, device="cuda").half()
    wp = svr2.vae_conv.pack_conv_weight((torch.randn(co, ci, 3, 3, 3, generator=g, device="cuda") * 0.03).half())
    b = torch.randn(co, generator=g, device="cuda").half()
    # stride-1 convolution with statistics
    ybuf, y = guarded(T * H * W * co, torch.float16)
    pbuf, part = guarded(T * svr2.vae_conv.gn_slots() * 32 * 3, torch.float32)
    svr2.vae_conv.causal_conv3d(x, wp, b, out=y.view(T, H, W, co), gn_partials=part.view(T, -1, 32, 3))
    torch.cuda.synchronize()
    check(ybuf, y.numel(), "conv output")
    check(pbuf, part.numel(), "conv statistics")
    # down-sampling convolution
    To, Ho, Wo = (T + 2 - 3) // 2 + 1, (H + 1 - 3) // 2 + 1, (W + 1 - 3) // 2 + 1
    dbuf, d = guarded(To * Ho * Wo * co, torch.float16)
    svr2._lib.check(lib.svr2_causal_conv3d_down_fwd(x.data_ptr(), None, 0, wp.data_ptr(), b.data_ptr(), d.data_ptr(), T, H, W, ci, co, 3, 1, 2, st),
                    "down")
    torch.cuda.synchronize()
    check(dbuf, d.numel(), "down-sampling convolution")
    # GroupNorm (both activation paths write through the same indexing; the arithmetic one here)
    gbuf, gy = guarded(T * H * W * co, torch.float16)
    svr2.vae_block.groupnorm_silu(x, b, b, 32, 1e-6, out=gy.view(T, H, W, co))
    torch.cuda.synchronize()
    check(gbuf, gy.numel(), "groupnorm")
    # transposes (aligned and scalar paths) and the pixel shuffle
    for rows, cols in ((128, 21 * 27 * 3 - 5), (128, 1680), (24, 189)):
        src = torch.randn(rows, cols, generator=g, device="cuda").half()
        tbuf, t = guarded(rows * cols, torch.float16)
        svr2._lib.check(lib.svr2_transpose_2b(src.data_ptr(), t.data_ptr(), rows, cols, cols, rows, 1, 0, 0, st), "transpose")
        torch.cuda.synchronize()
        check(tbuf, t.numel(), f"transpose {rows}x{cols}")
        assert torch.equal(t.view(cols, rows), src.t().contiguous())
    src = torch.randn(3, 5, 7, 8 * 64, generator=g, device="cuda").half()
    sbuf, s = guarded(5 * 10 * 14 * 64, torch.float16)
    svr2._lib.check(lib.svr2_pixel_shuffle3d_fwd(src.data_ptr(), s.data_ptr(), 3, 5, 7, 64, 2, 2, 1, st), "shuffle")
    torch.cuda.synchronize()
    check(sbuf, s.numel(), "pixel shuffle")
