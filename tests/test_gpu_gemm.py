"""GPU parity of the tcgen05 linear layers (csrc/gemm.cu) through the C ABI.

Checker: the oracle's op sequence for the same step (oracle/block.py: F.linear, F.silu * up, F.gelu(tanh), ada_out + residual)
— the reference's own ops (mlp.py:28-62, modulation.py:111-116, mmsr_block.py:114,126) — evaluated in fp32 from the bf16 inputs
with the reference's bf16 rounding points applied.  Tolerance: one bf16 ulp of the output magnitude for the plain linear (fp32
accumulation order differs from any other GEMM), 2e-2 absolute (north-star block tolerance) for the fused forms."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _mk(M, N, K, seed, dev, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = (torch.randn(M, K, generator=g, device=dev) * scale).bfloat16()
    w = (torch.randn(N, K, generator=g, device=dev) / K ** 0.5).bfloat16()
    b = (torch.randn(N, generator=g, device=dev) * 0.1).bfloat16()
    return x, w, b


def _ulp(t):
    """One bf16 ulp at the magnitude of each element (8 significant bits)."""
    return torch.exp2(torch.floor(torch.log2(t.abs().clamp_min(2.0 ** -120))) - 7)


def _lin32(x, w, b=None):
    torch.backends.cuda.matmul.allow_tf32 = False
    y = x.float() @ w.float().t()
    return y + b.float() if b is not None else y


# (M, N, K): ragged row counts, single tiles, K tails, the architectures' real widths
SHAPES = [(1, 64, 64), (58, 256, 128), (300, 192, 72), (257, 640, 320), (1024, 512, 2560), (16442, 2560, 2560), (4099, 7680, 2560),
          (2050, 3072, 12288)]


@pytest.mark.parametrize("M,N,K", SHAPES)
@pytest.mark.parametrize("with_bias", [False, True])
def test_linear_matches_fp32(svr2, cuda, M, N, K, with_bias):
    x, w, b = _mk(M, N, K, 1, cuda)
    y = svr2.ops.linear(x, w, b if with_bias else None)
    ref = _lin32(x, w, b if with_bias else None)
    err = (y.float() - ref).abs()
    tol = ref.abs().clamp_min(1.0) * 2.0 ** -8 + 1e-3          # half a bf16 ulp (rounding) + accumulation-order noise
    assert bool((err <= tol).all()), f"max err {float(err.max())} at {tuple(ref.shape)}"
    # and against the library GEMM the round-1 path used: at most one bf16 ulp apart
    lib = F.linear(x, w, b if with_bias else None)
    assert float((y.float() - lib.float()).abs().max()) <= float(ref.abs().max()) * 2.0 ** -7 + 1e-3


@pytest.mark.parametrize("M,N,K", [(58, 256, 128), (300, 384, 192), (16442, 6912, 2560), (1000, 6912, 2560)])
def test_linear_swiglu(svr2, cuda, M, N, K):
    x, wg, _ = _mk(M, N, K, 2, cuda)
    _, wu, _ = _mk(M, N, K, 3, cuda)
    y = svr2.ops.linear_swiglu(x, wg, wu)
    g, u = _lin32(x, wg).bfloat16(), _lin32(x, wu).bfloat16()          # the two linears' bf16 outputs (mlp.py:60-62)
    sg = F.silu(g.float()).bfloat16().float()
    ref = (sg * u.float()).bfloat16().float()
    # Derived bound: fp32 accumulation order may flip the last bf16 bit of each GEMM output (g and u), and the product is
    # rounded once more: |d| <= (silu'(g) ulp(g) + ulp(silu)) |u| + ulp(u) |silu(g)| + ulp(out), silu' <= 1.1 — silu(g) is a
    # bf16 tensor too (mlp.py:60), so a moved g can move its rounding by one more ulp.  Where |out| <= 1 this is
    # below the north-star 2e-2; the large products (|out| ~ 10, one ulp = 0.0625) are why an absolute bound cannot hold.
    tol = (1.1 * _ulp(g.float()) + _ulp(sg)) * u.float().abs() + _ulp(u.float()) * sg.abs() + _ulp(ref) + 1e-3
    err = (y.float() - ref).abs()
    assert bool((err <= tol).all()), float((err / tol).max())
    small = ref.abs() <= 1.0
    assert float(err[small].max()) <= 2e-2
    # vs the unfused library path of round 1 (two GEMMs + svr2_swiglu_fwd): same rounding points, different accumulation order
    old = svr2.ops.swiglu(F.linear(x, wg), F.linear(x, wu)).float()
    frac = float((y.float() != old).float().mean())
    assert bool(((y.float() - old).abs() <= 2 * tol).all()) and frac < 0.02, frac       # either side may carry the flips


@pytest.mark.parametrize("M,N,K", [(58, 256, 128), (2050, 12288, 3072), (700, 512, 64)])
def test_linear_gelu(svr2, cuda, M, N, K):
    x, w, b = _mk(M, N, K, 4, cuda)
    y = svr2.ops.linear(x, w, b, activation=svr2.ops.ACT_GELU_TANH)
    h = _lin32(x, w, b).bfloat16()                                      # proj_in's bf16 output (mlp.py:40-42)
    ref = F.gelu(h.float(), approximate="tanh").bfloat16().float()
    # one last-bit flip of h (accumulation order) times gelu' (<= 1.13), plus the output rounding
    tol = 1.13 * _ulp(h.float()) + _ulp(ref) + 1e-3
    err = (y.float() - ref).abs()
    assert bool((err <= tol).all()), float((err / tol).max())
    assert float(err[ref.abs() <= 1.0].max()) <= 2e-2
    assert float((y.float() != ref).float().mean()) < 0.02


@pytest.mark.parametrize("M,N,K,layer", [(58, 256, 128, 0), (16442, 2560, 2560, 0), (16442, 2560, 6912, 1), (2050, 3072, 12288, 1),
                                         (333, 192, 320, 1)])
@pytest.mark.parametrize("with_bias", [False, True])
def test_linear_gate_residual(svr2, cuda, M, N, K, layer, with_bias):
    x, w, b = _mk(M, N, K, 5, cuda)
    g = torch.Generator(device="cuda").manual_seed(6)
    emb = (torch.randn(6 * N, generator=g, device=cuda) * 0.5).bfloat16()
    gate_b = (torch.randn(N, generator=g, device=cuda) / N ** 0.5).bfloat16()
    resid = torch.randn(M, N, generator=g, device=cuda).bfloat16()
    bias = b if with_bias else None
    out = svr2.ops.linear_gate_residual(x, w, bias, emb, gate_b, layer, resid)
    h = _lin32(x, w, bias).bfloat16()
    gate = (emb.view(N, 2, 3)[:, layer, 2].float() + gate_b.float()).bfloat16()       # modulation.py:75-76,111-116
    hg = (h.float() * gate.float()).bfloat16().float()
    ref = (hg + resid.float()).bfloat16().float()
    # one last-bit flip of h times |gate|, the product's own rounding, the sum's rounding
    tol = _ulp(h.float()) * gate.float().abs() + _ulp(hg) + _ulp(ref) + 1e-3
    err = (out.float() - ref).abs()
    assert bool((err <= tol).all()), float((err / tol).max())
    calm = (ref.abs() <= 1.0) & (hg.abs() <= 1.0) & (h.float().abs() <= 2.0)      # no cancellation of large intermediates
    assert float(err[calm].max()) <= 2e-2
    # vs the unfused path of round 1: library GEMM + svr2_gate_residual_fwd
    old = svr2.ops.gate_residual(F.linear(x, w, bias), resid, emb, gate_b, layer).float()
    assert bool(((out.float() - old).abs() <= 2 * tol).all())
    assert float((out.float() != old).float().mean()) < 0.03


def test_linear_rejects_bad_shapes(svr2, cuda):
    x = torch.zeros(4, 64, dtype=torch.bfloat16, device=cuda)
    with pytest.raises(RuntimeError):
        svr2.ops.linear(x, torch.zeros(60, 64, dtype=torch.bfloat16, device=cuda))           # N % 64
    with pytest.raises(RuntimeError):
        svr2.ops.linear(x, torch.zeros(64, 32, dtype=torch.bfloat16, device=cuda))           # K mismatch
    with pytest.raises(NotImplementedError):
        svr2.ops.linear(x.float(), torch.zeros(64, 64, device=cuda))                         # fp32: no other path


# ---------------------------------------------------------------------------------------- fused qkv projection (svr2_linear_qkv_pack_fwd)
def _freqs(kind, svr2):
    import math
    if kind == svr2.ops.ROPE_3B_MM:
        dim = 128 // 3
        return 1.0 / (10000 ** (torch.arange(0, dim, 2)[: dim // 2].float() / dim))
    if kind == svr2.ops.ROPE_7B_PIX:
        return torch.linspace(1.0, 128.0, 10) * math.pi
    return None


@pytest.mark.parametrize("grid,shifted,heads,K,kind,joined", [
    ((2, 20, 36), False, 2, 256, 1, True), ((2, 20, 36), True, 2, 256, 1, False), ((3, 20, 36), True, 4, 128, 2, False),
    ((5, 17, 30), False, 2, 192, 2, False), ((1, 32, 32), True, 2, 256, 0, True), ((1, 128, 128), False, 20, 2560, 1, True),
    ((9, 68, 120), True, 20, 2560, 1, False), ((6, 68, 120), True, 24, 3072, 2, False), ((10, 135, 240), False, 24, 3072, 2, False)],
    ids=["3b_small_joined", "3b_small_swin_separate", "7b_small_swin", "7b_small_win", "norope_joined", "3b_c1_joined", "3b_1080p33f_swin",
         "7b_c3r_swin", "7b_c4_win"])
def test_linear_qkv_pack_equals_linear_then_qkv_pack(svr2, cuda, grid, shifted, heads, K, kind, joined):
    """The fused projection (GEMM epilogue: q/k RMSNorm + RoPE + window scatter, text rows replicated afterwards) against the
    two-kernel path it replaces (svr2_linear_fwd + svr2_qkv_pack_fwd, itself checked against the oracle in test_gpu_kernels.py):
    v rows bit-equal; q / k rows equal except where the fp32 order of the row's sum of squares moves one bf16 rounding."""
    ops = svr2.ops
    txt_len = 58
    plan = svr2.get_plan(grid, (4, 3, 3), shifted, txt_len)
    L, N = plan.n_tokens, 3 * heads * 128
    g = torch.Generator(device="cuda").manual_seed(11)
    x = torch.randn(L + txt_len, K, generator=g, device=cuda).bfloat16()
    w_vid = (torch.randn(N, K, generator=g, device=cuda) / K ** 0.5).bfloat16()
    w_txt = w_vid if joined else (torch.randn(N, K, generator=g, device=cuda) / K ** 0.5).bfloat16()
    nw = [(1.0 + 0.1 * torch.randn(128, generator=g, device=cuda)).bfloat16() for _ in range(4)]
    f = _freqs(kind, svr2)
    fdev = None if f is None else f.to(cuda)
    fkey = () if f is None else tuple(float(v) for v in f.tolist())
    assert ops.qkv_fused_supported(plan, kind, len(fkey), heads)
    # reference path of the repo: dense projection, then the pack kernel
    qkv_v = ops.linear(x[:L].contiguous(), w_vid)
    qkv_t = ops.linear(x[L:].contiguous(), w_txt)
    q0, k0, v0 = ops.qkv_pack(plan, qkv_v, qkv_t, nw[0], nw[1], nw[2], nw[3], fdev, kind, heads, 1e-5)
    if joined:
        bufs = ops.linear_qkv_pack(plan, x, w_vid, 0, *nw, fdev, fkey, kind, heads, 1e-5)
    else:
        bufs = ops.linear_qkv_pack(plan, x[:L].contiguous(), w_vid, 0, *nw, fdev, fkey, kind, heads, 1e-5)
        bufs = ops.linear_qkv_pack(plan, x[L:].contiguous(), w_txt, L, *nw, fdev, fkey, kind, heads, 1e-5, out=bufs)
    q1, k1, v1, tq = bufs
    ops.replicate_text(plan, tq, q1, k1, v1, heads)
    torch.cuda.synchronize()
    assert torch.equal(v1, v0), "v rows must be bit-equal (a gather of the projection's bf16 output)"
    for name, a, b in (("q", q1, q0), ("k", k1, k0)):
        d = (a.float() - b.float()).abs()
        frac = float((a != b).float().mean())
        # a rotated element is a difference of two products and can cancel to ~0: ulps are taken at the head row's magnitude
        ulps = float((d / _ulp(b.float().abs().amax(dim=-1, keepdim=True))).max())
        print(f"{name}: {frac * 100:.4f} % of elements differ, max {ulps:.1f} bf16 ulp")
        assert frac < 5e-3 and ulps <= 2.0, (name, frac, ulps)
    # second call into the same buffers: nothing stale, same bits
    q2, k2, v2, tq2 = ops.linear_qkv_pack(plan, x if joined else x[:L].contiguous(), w_vid, 0, *nw, fdev, fkey, kind, heads, 1e-5)
    torch.cuda.synchronize()
    vid_rows = torch.from_numpy(__import__("numpy").concatenate(
        [__import__("numpy").arange(r, r + n) for r, n in zip(plan.table[:, 6], plan.table[:, 7])])).to(cuda)
    assert torch.equal(q2[vid_rows], q1[vid_rows]) and torch.equal(v2[vid_rows], v1[vid_rows])


def test_linear_qkv_pack_rejects(svr2, cuda):
    plan = svr2.get_plan((1, 8, 8), (4, 3, 3), False, 58)
    x = torch.zeros(64 + 58, 64, dtype=torch.bfloat16, device=cuda)
    w3 = torch.zeros(3 * 3 * 128, 64, dtype=torch.bfloat16, device=cuda)
    nw = [torch.ones(128, dtype=torch.bfloat16, device=cuda)] * 4
    with pytest.raises(RuntimeError):       # odd head count: a 256-column tile would straddle q | k
        svr2.ops.linear_qkv_pack(plan, x, w3, 0, *nw, None, (), 0, 3, 1e-5)
    assert not svr2.ops.qkv_fused_supported(plan, 1, 7, 2)        # a frequency count the epilogue is not built for
