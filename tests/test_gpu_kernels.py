"""GPU parity tests, kernel by kernel, every call through the C ABI (ctypes -> libseedvr2_b200.so).
The checker is the oracle (oracle/), itself pinned to the reference by tests/test_oracle_golden.py."""
import hashlib
import json
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import cases
from oracle import block as oblock
from oracle import geometry, partition, rope

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GEO = json.load(open(os.path.join(ROOT, "tests", "golden", "geometry.json")))
BF16_ULP_AT_4 = 2.0 ** -6     # one bf16 ulp for |x| in [2, 4)


def _sha(t: torch.Tensor) -> str:
    return hashlib.sha256(np.ascontiguousarray(t.cpu().numpy()).tobytes()).hexdigest()


# ----------------------------------------------------------------------------------------------- window index
@pytest.mark.parametrize("key", sorted(GEO))
def test_window_index_bit_exact(svr2, cuda, key):
    """Bit-exact vs the reference's na.window_idx at every BASELINE grid (sha256 of the full int64 arrays)."""
    rec = GEO[key]
    grid, method = key.split(":")
    T, H, W = (int(v) for v in grid.split("x"))
    plan = svr2.WindowPlan((T, H, W), (4, 3, 3), method == "swin", cases.TXT_LEN)
    tgt, src, wshape, cu = plan.index_tensors(cuda)
    torch.cuda.synchronize()
    assert tgt.dtype == torch.int64 and src.dtype == torch.int64 and cu.dtype == torch.int32
    assert _sha(tgt) == rec["tgt_sha256"]
    assert _sha(src) == rec["src_sha256"]
    assert wshape.cpu().tolist() == rec["window_shape"]
    assert cu.cpu().tolist() == rec["cu_seqlens_txt58"]
    # size-independent property: src is the inverse permutation of tgt
    assert torch.equal(src[tgt], torch.arange(T * H * W, device=cuda))


# ----------------------------------------------------------------------------------------------- adaLN kernels
def _ada_inputs(L, D, seed, dev):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(L, D, generator=g).bfloat16()
    r = torch.randn(L, D, generator=g).bfloat16()
    emb = (0.5 * torch.randn(1, 6 * D, generator=g)).bfloat16()
    shift, gate = (torch.randn(D, generator=g) / D ** 0.5).bfloat16(), (torch.randn(D, generator=g) / D ** 0.5).bfloat16()
    scale = (1 + torch.randn(D, generator=g) / D ** 0.5).bfloat16()
    return [t.to(dev) for t in (x, r, emb, shift, scale, gate)]


def _close_bf16(a, b, what, max_ulp_frac=2e-3):
    """Same op order and rounding points -> equal except where fp32 summation order flips a bf16 rounding."""
    a, b = a.float(), b.float()
    d = (a - b).abs()
    tol = 2.0 ** -7 * b.abs().clamp(min=1.0)        # <= 1 bf16 ulp of the value
    assert bool((d <= tol).all()), f"{what}: max diff {float(d.max())}"
    assert float((d > 0).float().mean()) <= max_ulp_frac, f"{what}: {float((d > 0).float().mean())} of elements differ"


@pytest.mark.parametrize("L,D", [(58, 128), (1000, 2560), (333, 3072), (5, 1024), (64, 4096)])
@pytest.mark.parametrize("layer", [0, 1])
def test_modnorm(svr2, cuda, L, D, layer):
    x, _, emb, shift, scale, _ = _ada_inputs(L, D, 100 + layer, cuda)
    y = svr2.ops.modnorm(x, emb.reshape(-1), shift, scale, layer, 1e-5)
    ref = oblock.ada_in(oblock.rmsnorm(x, 1e-5), emb, layer, shift, scale)      # torch bf16 ops, the reference's order
    _close_bf16(y, ref, "modnorm")
    y0 = svr2.ops.modnorm(x, None, None, None, layer, 1e-5)
    _close_bf16(y0, oblock.rmsnorm(x, 1e-5), "rmsnorm only")


@pytest.mark.parametrize("L,D", [(58, 128), (1000, 2560), (4099, 256)])
def test_modnorm_affine(svr2, cuda, L, D):
    """Output norm (RMSNorm with weight) + output modulation with the attn-layer pair (quirk Q1), dit_3b/nadit.py:233-244."""
    x, _, emb, shift, scale, _ = _ada_inputs(L, D, 300, cuda)
    w = (1 + 0.1 * torch.randn(D, generator=torch.Generator().manual_seed(7))).bfloat16().to(cuda)
    y = svr2.ops.modnorm_affine(x, w, emb.reshape(-1), shift, scale, 0, 1e-5)
    ref = oblock.ada_in(oblock.rmsnorm(x, 1e-5, w), emb, 0, shift, scale)
    _close_bf16(y, ref, "modnorm_affine")
    _close_bf16(svr2.ops.modnorm_affine(x, w, None, None, None, 0, 1e-5), oblock.rmsnorm(x, 1e-5, w), "affine rmsnorm only")


@pytest.mark.parametrize("L,D", [(58, 128), (1000, 2560), (333, 3072), (4099, 256)])
def test_gate_residual(svr2, cuda, L, D):
    x, r, emb, shift, scale, gate = _ada_inputs(L, D, 7, cuda)
    for layer in (0, 1):
        out = svr2.ops.gate_residual(x, r, emb.reshape(-1), gate, layer)
        ref = oblock.ada_out(x, emb, layer, gate) + r
        assert torch.equal(out, ref), f"gate_residual layer {layer}: max diff {float((out.float() - ref.float()).abs().max())}"
    assert torch.equal(svr2.ops.gate_residual(x, r, None, None, 0), x + r)
    out, y = svr2.ops.gate_residual_modnorm(x, r, emb.reshape(-1), gate, shift, scale, 1e-5)
    ref_out = oblock.ada_out(x, emb, 0, gate) + r
    assert torch.equal(out, ref_out)
    _close_bf16(y, oblock.ada_in(oblock.rmsnorm(ref_out, 1e-5), emb, 1, shift, scale), "fused modnorm")


@pytest.mark.parametrize("L,Hd", [(58, 6912), (1000, 6912), (3, 8), (4099, 256)])
def test_swiglu(svr2, cuda, L, Hd):
    """silu(gate) * up vs the reference's two bf16 ATen ops (mlp.py:60-62).  The kernel uses ex2.approx / fast division, so
    a bf16 rounding may flip on a few elements: <= 1 bf16 ulp, < 0.1 % of elements."""
    g = torch.Generator().manual_seed(L)
    a = (2 * torch.randn(L, Hd, generator=g)).bfloat16().to(cuda)
    b = torch.randn(L, Hd, generator=g).bfloat16().to(cuda)
    out = svr2.ops.swiglu(a, b)
    ref = F.silu(a) * b
    _close_bf16(out, ref, "swiglu", max_ulp_frac=1e-3)


# ----------------------------------------------------------------------------------------------- qkv pack
def _pack_case(model, grid, shifted, heads, seed, txt_len=cases.TXT_LEN):
    g = torch.Generator().manual_seed(seed)
    T, H, W = grid
    L = T * H * W
    qkv_v = torch.randn(L, 3 * heads * 128, generator=g).bfloat16()
    qkv_t = torch.randn(txt_len, 3 * heads * 128, generator=g).bfloat16()
    w = [(1 + 0.1 * torch.randn(128, generator=g)).bfloat16() for _ in range(4)]
    return qkv_v, qkv_t, w, cases.rope_freqs(model)


def _oracle_pack(model, grid, shifted, heads, qkv_v, qkv_t, w, freqs, txt_len):
    """Expected packed q,k,v from the oracle's op sequence (bf16 on CPU)."""
    T, H, W = grid
    tgt, src, wshape, cu = geometry.window_index(T, H, W, (4, 3, 3), shifted, txt_len)
    vq, vk, vv = qkv_v.index_select(0, torch.from_numpy(tgt)).reshape(-1, 3, heads, 128).unbind(1)
    tq, tk, tv = qkv_t.reshape(-1, 3, heads, 128).unbind(1)
    vq, vk = oblock.rmsnorm(vq, 1e-5, w[0]), oblock.rmsnorm(vk, 1e-5, w[1])
    tq, tk = oblock.rmsnorm(tq, 1e-5, w[2]), oblock.rmsnorm(tk, 1e-5, w[3])
    ws = [tuple(int(v) for v in r) for r in wshape]
    if model == "3b":
        va, ta = rope.angles_3b(ws, txt_len, freqs)
        vq, vk, tq, tk = rope.rotate(vq, va), rope.rotate(vk, va), rope.rotate(tq, ta), rope.rotate(tk, ta)
    else:
        va = rope.angles_7b(ws, freqs).to(vq.dtype)
        vq, vk = rope.rotate(vq, va), rope.rotate(vk, va)
    qs, ks, vs, off = [], [], [], 0
    for (f, h, w_) in ws:
        n = f * h * w_
        qs += [vq[off:off + n], tq]; ks += [vk[off:off + n], tk]; vs += [vv[off:off + n], tv]
        off += n
    return torch.cat(qs), torch.cat(ks), torch.cat(vs)


@pytest.mark.parametrize("model,grid,shifted,heads", [
    ("3b", (2, 20, 36), False, 2), ("3b", (1, 32, 32), True, 3), ("7b", (3, 20, 36), False, 2), ("7b", (5, 17, 30), True, 3),
    ("3b", (1, 1, 1), False, 1), ("7b", (1, 2, 3), True, 1),
])
def test_qkv_pack(svr2, cuda, model, grid, shifted, heads):
    qkv_v, qkv_t, w, freqs = _pack_case(model, grid, shifted, heads, 5)
    plan = svr2.get_plan(grid, (4, 3, 3), shifted, cases.TXT_LEN)
    kind = svr2.ops.ROPE_3B_MM if model == "3b" else svr2.ops.ROPE_7B_PIX
    q, k, v = svr2.ops.qkv_pack(plan, qkv_v.to(cuda), qkv_t.to(cuda), *[t.to(cuda) for t in w], freqs.to(cuda), kind, heads, 1e-5)
    eq, ek, ev = _oracle_pack(model, grid, shifted, heads, qkv_v, qkv_t, w, freqs, cases.TXT_LEN)
    assert torch.equal(v.cpu(), ev), "v rows are a pure gather: must be bit-exact"
    for got, exp, nm in ((q, eq, "q"), (k, ek, "k")):
        d = (got.cpu().float() - exp.float()).abs()
        # cosf/sinf (device) vs the host libm differ by <= 1 ulp fp32 -> a bf16 rounding flips on a few elements
        assert float(d.max()) <= 2 * BF16_ULP_AT_4, f"{nm}: max diff {float(d.max())}"
        assert float((d > 0).float().mean()) < 5e-3, f"{nm}: {float((d > 0).float().mean())} of elements differ"


@pytest.mark.parametrize("model,grid,shifted,heads", [("3b", (1, 32, 32), True, 3), ("7b", (5, 17, 30), True, 3), ("3b", (2, 20, 36), False, 2)])
def test_qkv_pack_row_map_equals_table_search(svr2, cuda, model, grid, shifted, heads):
    """The plan's packed-row map (svr2_pack_rowmap) is an integer restatement of the per-CTA window search: both paths of
    svr2_qkv_pack_fwd must give bit-identical q, k, v; the map itself is checked against the oracle's window index."""
    qkv_v, qkv_t, w, freqs = _pack_case(model, grid, shifted, heads, 7)
    plan = svr2.get_plan(grid, (4, 3, 3), shifted, cases.TXT_LEN)
    kind = svr2.ops.ROPE_3B_MM if model == "3b" else svr2.ops.ROPE_7B_PIX
    dq, dt = qkv_v.to(cuda), qkv_t.to(cuda)
    dw = [t.to(cuda) for t in w]
    fr = freqs.to(cuda)
    q1, k1, v1 = svr2.ops.qkv_pack(plan, dq, dt, *dw, fr, kind, heads, 1e-5)
    L = svr2._lib.lib()
    P = plan.packed_rows
    q0, k0, v0 = (torch.empty(P, heads, 128, dtype=torch.bfloat16, device=cuda) for _ in range(3))
    T, H, W = grid
    rc = L.svr2_qkv_pack_fwd(dq.data_ptr(), dt.data_ptr(), *[t.data_ptr() for t in dw], fr.data_ptr(), fr.numel(), kind,
                             plan.on(cuda)["table"].data_ptr(), None, plan.n_windows, T, H, W, plan.txt_len, heads, 1e-5,
                             q0.data_ptr(), k0.data_ptr(), v0.data_ptr(), P, torch.cuda.current_stream().cuda_stream)
    svr2._lib.check(rc, "svr2_qkv_pack_fwd")
    torch.cuda.synchronize()
    for a, b, nm in ((q0, q1, "q"), (k0, k1, "k"), (v0, v1, "v")):
        assert torch.equal(a, b), f"{nm}: row-map path differs from the table-search path"
    rm = plan.row_map(cuda).cpu().numpy()
    tgt = geometry.window_index(T, H, W, plan.window, plan.shifted, plan.txt_len)[0]      # video token of every window slot
    vid_rows = rm[:, 0] >= 0
    assert np.array_equal(rm[vid_rows, 0], tgt), "video rows of the map are na.window_idx in packed order"
    txt = -(rm[~vid_rows, 0] + 1)
    assert np.array_equal(txt, np.tile(np.arange(plan.txt_len), plan.n_windows)), "text rows repeat 0..Lt-1 per window"


# ----------------------------------------------------------------------------------------------- window attention
def _sdpa_reference(plan, q, k, v, heads):
    """fp32 per-window attention on the packed layout (the reference's pytorch_varlen_attention semantics)."""
    cu = plan.cu_seqlens
    T, H, W = plan.grid
    out_v = torch.zeros(plan.n_tokens, heads * 128)
    tacc = torch.zeros(plan.txt_len, heads * 128)
    tgt = geometry.window_index(T, H, W, plan.window, plan.shifted, plan.txt_len)[0]
    voff = 0
    for i in range(plan.n_windows):
        a, b = int(cu[i]), int(cu[i + 1])
        qi, ki, vi = (t[a:b].float().permute(1, 0, 2).unsqueeze(0) for t in (q, k, v))
        o = F.scaled_dot_product_attention(qi, ki, vi).squeeze(0).permute(1, 0, 2).reshape(b - a, heads * 128)
        n = b - a - plan.txt_len
        out_v[torch.from_numpy(tgt[voff:voff + n])] = o[:n]
        tacc += o[n:].bfloat16().float()       # per-window text rows are bf16 before the mean (na.py:414)
        voff += n
    return out_v, tacc / plan.n_windows


ATTN_CASES = [((2, 20, 36), False, 2), ((1, 32, 32), True, 3), ((5, 17, 30), True, 2), ((1, 1, 1), False, 1),
              ((9, 30, 54), False, 1), ((3, 46, 81), True, 1)]


@pytest.mark.parametrize("grid,shifted,heads", ATTN_CASES)
def test_window_attn_generations_bit_identical(svr2, cuda, grid, shifted, heads):
    """The two product kernels — generation 4 (two query tiles per CTA, whole key tiles) and generation 8 (CTA pairs, one query
    tile per CTA, two softmax warpgroups on alternate key tiles) — use the same exponent reference, lazy-rescale rule and
    summation order: their outputs must be identical bit for bit, so the length-based dispatch cannot move a result."""
    ops = svr2.ops
    plan = svr2.get_plan(grid, (4, 3, 3), shifted, cases.TXT_LEN)
    g = torch.Generator().manual_seed(11)
    q, k, v = (torch.randn(plan.packed_rows, heads, 128, generator=g).bfloat16().to(cuda) for _ in range(3))

    def run(impl, q):
        ov, ot = ops.window_attn(plan, q, k, v, heads, impl=impl)
        torch.cuda.synchronize()
        return ov.clone(), ot.clone()

    base = run(ops.ATTN_GEN4, q)
    for other in (ops.ATTN_GEN8, ops.ATTN_GEN8, ops.ATTN_AUTO):      # repeated launches must also agree with each other
        ov, ot = run(other, q)
        assert torch.equal(base[0], ov), f"video rows differ between generation 4 and kernel {other}"
        assert torch.equal(base[1], ot), f"pooled text rows differ between generation 4 and kernel {other}"
    # rows whose maximum jumps from key tile to key tile exercise the lazy rescale of O, where generation 8 hands the
    # reference exponent and the row sum from warpgroup to warpgroup
    q2 = q.clone()
    q2[: plan.packed_rows // 2] *= 6.0
    base = run(ops.ATTN_GEN4, q2)
    ov, ot = run(ops.ATTN_GEN8, q2)
    assert torch.equal(base[0], ov) and torch.equal(base[1], ot), "generation 8 differs from generation 4 under lazy rescales"


def test_window_attn_experiments_not_in_product_build(svr2, cuda):
    """Earlier generations live under csrc/experiments/ and are NOT part of the shipped library: naming one is an error."""
    plan = svr2.get_plan((1, 8, 8), (4, 3, 3), False, cases.TXT_LEN)
    q = torch.zeros(plan.packed_rows, 1, 128, dtype=torch.bfloat16, device=cuda)
    with pytest.raises(NotImplementedError):
        svr2.ops.window_attn(plan, q, q, q, 1, impl=7)


@pytest.mark.parametrize("grid,shifted,heads", ATTN_CASES + [((9, 68, 120), True, 4), ((1, 128, 128), False, 4)])
def test_window_attn_generation10_matches_generation4(svr2, cuda, grid, shifted, heads):
    """Generation 10 (two threads per query row, 64 key columns each) computes every exponential with generation 4's arithmetic
    and the exact same row max; only the row sum is added in a different order (two partial sums per row), so outputs may differ
    by one bf16 rounding of O / l — never more — and repeated launches are bit-identical."""
    ops = svr2.ops
    plan = svr2.get_plan(grid, (4, 3, 3), shifted, cases.TXT_LEN)
    try:
        z = torch.zeros(plan.packed_rows, heads, 128, dtype=torch.bfloat16, device=cuda)
        ops.window_attn(plan, z, z, z, heads, impl=ops.ATTN_GEN10)
    except NotImplementedError:
        pytest.skip("generation 10 is an experiment kernel: built only with `make -C comfyui-seedvr2_b200/csrc EXPERIMENTS=1`")
    g = torch.Generator().manual_seed(13)
    q, k, v = (torch.randn(plan.packed_rows, heads, 128, generator=g).bfloat16().to(cuda) for _ in range(3))
    for scale in (1.0, 6.0):                     # 6.0: row maxima that jump between key tiles exercise the lazy rescale
        qq = q.clone()
        qq[: plan.packed_rows // 2] *= scale
        a_v, a_t = ops.window_attn(plan, qq, k, v, heads, impl=ops.ATTN_GEN4)
        b_v, b_t = ops.window_attn(plan, qq, k, v, heads, impl=ops.ATTN_GEN10)
        c_v, c_t = ops.window_attn(plan, qq, k, v, heads, impl=ops.ATTN_GEN10)
        torch.cuda.synchronize()
        assert torch.equal(b_v, c_v) and torch.equal(b_t, c_t), "generation 10 is not deterministic"
        d = (a_v.float() - b_v.float()).abs()
        ulp = torch.exp2(torch.floor(torch.log2(a_v.float().abs().clamp_min(2.0 ** -100))) - 7)
        assert bool((d <= ulp).all()), float((d / ulp).max())
        assert float((a_v != b_v).float().mean()) < 0.02
        # pooled text rows: a mean over windows of per-window bf16 rows, each of which may carry one such rounding
        assert float((a_t.float() - b_t.float()).abs().max()) <= 2.0 ** -8 * float(a_v.float().abs().max())


@pytest.mark.parametrize("impl", ["auto", "gen8", "gen4", "gen10", "mma"])
@pytest.mark.parametrize("grid,shifted,heads", ATTN_CASES)
def test_window_attn_vs_fp32(svr2, cuda, impl, grid, shifted, heads):
    code = {"auto": svr2.ops.ATTN_AUTO, "gen8": svr2.ops.ATTN_GEN8, "gen4": svr2.ops.ATTN_GEN4, "gen10": svr2.ops.ATTN_GEN10,
            "mma": svr2.ops.ATTN_MMA}[impl]
    plan = svr2.get_plan(grid, (4, 3, 3), shifted, cases.TXT_LEN)
    if impl == "gen10":
        try:
            z = torch.zeros(plan.packed_rows, heads, 128, dtype=torch.bfloat16, device=cuda)
            svr2.ops.window_attn(plan, z, z, z, heads, impl=code)
        except NotImplementedError:
            pytest.skip("generation 10 is an experiment kernel (make EXPERIMENTS=1)")
    g = torch.Generator().manual_seed(3)
    q, k, v = (torch.randn(plan.packed_rows, heads, 128, generator=g).bfloat16() for _ in range(3))
    q = q * 1.5      # sharper softmax
    ov, ot = svr2.ops.window_attn(plan, q.to(cuda), k.to(cuda), v.to(cuda), heads, impl=code)
    torch.cuda.synchronize()
    ev, et = _sdpa_reference(plan, q, k, v, heads)
    dv, dt = (ov.cpu().float() - ev).abs().max(), (ot.cpu().float() - et).abs().max()
    assert float(dv) <= 2e-2 and float(dt) <= 2e-2, f"{impl}: vid {float(dv)} txt {float(dt)}"   # north-star tolerance


def _sdpa_reference_gpu(plan, q, k, v, heads):
    """The oracle's per-window attention (oracle/block.py:window_attention, i.e. the reference's pytorch_varlen_attention
    loop, attention.py:24-61 + the text mean na.py:414) in fp32 ON THE GPU, so that BASELINE-size grids finish in seconds.
    Plain torch ops on the packed layout — an independent restatement, none of the library's kernels."""
    dev = q.device
    cu = plan.cu_seqlens
    tgt = torch.from_numpy(geometry.window_index(*plan.grid, plan.window, plan.shifted, plan.txt_len)[0]).to(dev)
    out_v = torch.zeros(plan.n_tokens, heads * 128, device=dev)
    tacc = torch.zeros(plan.txt_len, heads * 128, device=dev)
    voff = 0
    for i in range(plan.n_windows):
        a, b = int(cu[i]), int(cu[i + 1])
        qi, ki, vi = (t[a:b].float().permute(1, 0, 2) for t in (q, k, v))          # [heads, n, 128]
        s = torch.baddbmm(torch.zeros((), device=dev), qi, ki.transpose(1, 2), beta=0, alpha=128 ** -0.5)
        o = torch.bmm(torch.softmax(s, dim=-1), vi).permute(1, 0, 2).reshape(b - a, heads * 128)
        n = b - a - plan.txt_len
        out_v[tgt[voff:voff + n]] = o[:n]
        tacc += o[n:].bfloat16().float()       # per-window text rows are bf16 before the mean (na.py:414)
        voff += n
    return out_v, tacc / plan.n_windows


# BASELINE grids at the real head counts (3B: 20 heads, 7B: 24 heads): C1 win/swin, C2 win/swin, C3 per-rank, C3, C4
FULL_ATTN_CASES = [((1, 128, 128), False, 20), ((1, 128, 128), True, 20), ((9, 90, 160), False, 20), ((9, 90, 160), True, 20),
                   ((6, 68, 120), True, 24), ((31, 68, 120), False, 24), ((10, 135, 240), True, 24)]


@pytest.mark.parametrize("grid,shifted,heads", FULL_ATTN_CASES)
def test_window_attn_full_size_vs_fp32(svr2, cuda, grid, shifted, heads):
    """svr2_window_attn_fwd at the BASELINE token grids with the architectures' head counts against the fp32 restatement of
    the reference's per-window SDPA loop (run on the GPU): max-abs <= 2e-2 (north-star tolerance) on every video row and
    on the pooled text rows.  Also: V = ones gives exactly-normalised rows (masking + tile coverage + scatter coverage)."""
    torch.backends.cuda.matmul.allow_tf32 = False
    plan = svr2.get_plan(grid, (4, 3, 3), shifted, cases.TXT_LEN)
    g = torch.Generator(device="cuda").manual_seed(11)
    P = plan.packed_rows
    q, k, v = (torch.randn(P, heads, 128, generator=g, device=cuda, dtype=torch.float32).bfloat16() for _ in range(3))
    ov, ot = svr2.ops.window_attn(plan, q, k, v, heads)
    ev, et = _sdpa_reference_gpu(plan, q, k, v, heads)
    dv, dt = float((ov.float() - ev).abs().max()), float((ot.float() - et).abs().max())
    print(f"window_attn {grid} {'swin' if shifted else 'win'} heads {heads}: max|d| vid {dv:.5f} txt {dt:.5f} (<= 2e-2)")
    assert dv <= 2e-2 and dt <= 2e-2
    del ev, et
    ones = torch.ones_like(v)
    o1, t1 = svr2.ops.window_attn(plan, q, k, ones, heads)
    assert float((o1.float() - 1).abs().max()) <= 2 ** -7 and float((t1.float() - 1).abs().max()) <= 2 ** -7


# ----------------------------------------------------------------------------------------------- halo blend
@pytest.mark.parametrize("overlap", [1, 2, 3, 4, 7])
def test_halo_blend(svr2, cuda, overlap):
    g = torch.Generator().manual_seed(overlap)
    a, b = torch.rand(overlap, 18, 20, 3, generator=g), torch.rand(overlap, 18, 20, 3, generator=g)
    out = svr2.ops.halo_blend(a.to(cuda), b.to(cuda), overlap)
    ref = partition.blend(a, b, overlap)
    # the blend itself is separately rounded mul, mul, add like the reference's tensor ops; the Hann weight comes from the
    # device cosf, which may differ from torch's CPU cos by an ulp
    assert float((out.cpu() - ref).abs().max()) <= 1e-6
    if overlap < 3:      # linear weights are exact on both sides: bit-equal to blend_overlapping_frames
        assert torch.equal(out.cpu(), ref)
