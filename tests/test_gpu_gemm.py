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
