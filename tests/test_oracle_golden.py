"""CPU: the oracle (oracle/) against the golden vectors generated from the real reference (tests/golden/make_golden.py).
This is what pins the oracle; everything else is checked against the oracle."""
import hashlib
import json
import os

import numpy as np
import pytest
import torch

import cases
from oracle import block as oblock
from oracle import geometry, partition

GOLD = os.path.join(os.path.dirname(__file__), "golden")
GEO = json.load(open(os.path.join(GOLD, "geometry.json")))


def _sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize("key", sorted(GEO))
def test_geometry_matches_reference(key):
    rec = GEO[key]
    grid, method = key.split(":")
    T, H, W = (int(v) for v in grid.split("x"))
    shifted = method == "swin"
    b = geometry.window_bounds(T, H, W, (4, 3, 3), shifted)
    assert [list(x) for x in b] == rec["bounds"]
    tgt, src, shape, cu = geometry.window_index(T, H, W, (4, 3, 3), shifted, cases.TXT_LEN)
    assert shape.tolist() == rec["window_shape"]
    assert len(b) == rec["window_count"]
    assert cu.tolist() == rec["cu_seqlens_txt58"]
    assert _sha(tgt) == rec["tgt_sha256"] and _sha(src) == rec["src_sha256"]
    if "tgt_idx" in rec:
        assert tgt.tolist() == rec["tgt_idx"] and src.tolist() == rec["src_idx"]
    # domain properties: a permutation and its inverse
    assert np.array_equal(np.sort(tgt), np.arange(T * H * W))
    assert np.array_equal(src[tgt], np.arange(T * H * W))


@pytest.mark.parametrize("case", cases.BLOCK_CASES, ids=lambda c: c.name)
def test_block_matches_reference(case):
    gold = torch.load(os.path.join(GOLD, f"block_{case.name}.pt"))
    params = cases.block_params(case)
    vid, txt, emb = cases.block_inputs(case)
    assert gold["in_digest"] == {"vid": cases.tensor_digest(vid), "txt": cases.tensor_digest(txt), "emb": cases.tensor_digest(emb)}
    cfg = oblock.BlockCfg(model=case.model, dim=case.dim, heads=case.heads, head_dim=case.head_dim, shared=case.shared,
                          is_last=case.is_last, shifted=case.shifted, window=case.window)
    with torch.no_grad():
        ov, ot = oblock.block_forward(cfg, params, vid, txt, case.grid, emb)
    # fp32: same ops in the same order -> identical up to the text-mean summation order
    assert torch.equal(ov, gold["vid_out_fp32"])
    assert float((ot - gold["txt_out_fp32"]).abs().max()) < 1e-5
    p16 = {k: (v if k.endswith("rope.freqs") else v.bfloat16()) for k, v in params.items()}
    with torch.no_grad():
        ov16, ot16 = oblock.block_forward(cfg, p16, vid.bfloat16(), txt.bfloat16(), case.grid, emb.bfloat16())
    assert torch.equal(ov16, gold["vid_out_bf16"])
    assert float((ot16.float() - gold["txt_out_bf16"].float()).abs().max()) <= 2 ** -7


def test_partition_split_and_blend():
    # inference_cli.py:1087-1097 — 121 frames, 8 GPUs, overlap 4
    r = partition.segment_ranges(121, 8, 4)
    assert [b - a for a, b in r] == [20, 19, 19, 19, 19, 19, 19, 15]
    assert r[0] == (0, 20) and r[-1] == (106, 121)
    assert partition.segment_ranges(10, 3, 0) == [(0, 4), (4, 7), (7, 10)]
    w = partition.blend_weights(4)
    assert torch.allclose(w, torch.tensor([1.0, 1.0, 0.0, 0.0]), atol=1e-6)      # Hann over the middle third
    assert torch.allclose(partition.blend_weights(2), torch.tensor([1.0, 0.0]))
    a, b = torch.ones(4, 2, 2, 3), torch.zeros(4, 2, 2, 3)
    out = partition.stitch([torch.cat([torch.ones(6, 2, 2, 3), a]), torch.cat([b, torch.zeros(5, 2, 2, 3)])], 4)
    assert out.shape[0] == 15 and float(out[:8].min()) == 1.0 and float(out[8:].max()) == 0.0


def _bf16_round(x):
    """round-to-nearest-even fp32 -> bf16, returned as fp32 (what every bf16 ATen op does on store)"""
    u = np.asarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    return ((u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000).astype(np.uint32).view(np.float32)


def _nudge(x, ulps):
    return (np.asarray(x, dtype=np.float32).view(np.uint32).astype(np.int64) + ulps).astype(np.uint32).view(np.float32)


def test_rmsnorm_division_free_rounding_lemma():
    """The CUDA RMSNorm paths (csrc/common.cuh rcp_approx / sqrt_approx) replace the reference's fp32 division and sqrt
    (normalization.py:88-109: x / sqrt(mean + eps) on bf16 tensors) by x * rcp.approx(rms) and sqrt.approx.  That is only
    legitimate if the bf16 result is identical for EVERY input; the quotient / root of 8-bit significands never sits
    within a few fp32 ulps of a bf16 rounding midpoint.  Checked exhaustively over all significand pairs (exponents only
    shift), with the approximations perturbed by more than the PTX error bounds (rcp.approx: 1 ulp, sqrt.approx: 2^-23)."""
    man = np.arange(128, 256, dtype=np.float32)
    for ex in (-20, -7, -1, 0, 3, 11):
        x = (man * np.float32(2.0) ** ex).astype(np.float32)[:, None]
        for ed in (-9, -3, 0, 2, 5):
            d = (man * np.float32(2.0) ** ed).astype(np.float32)[None, :]
            ref = _bf16_round((x / d).astype(np.float32))
            rcp = (np.float32(1) / d).astype(np.float32)
            for ulps in (-2, -1, 0, 1, 2):
                got = _bf16_round((x * _nudge(rcp, ulps)).astype(np.float32))
                assert np.array_equal(got, ref), (ex, ed, ulps)
    for e in range(-30, 10):
        v = (man * np.float32(2.0) ** e).astype(np.float32)
        root = np.sqrt(v).astype(np.float32)
        ref = _bf16_round(root)
        for ulps in range(-4, 5):
            assert np.array_equal(_bf16_round(_nudge(root, ulps)), ref), (e, ulps)


@pytest.mark.parametrize("case", cases.STACK_CASES, ids=lambda c: c.name)
def test_nadit_forward_matches_reference(case):
    """Whole NaDiT.forward (dit_3b/nadit.py:171-248 incl. quirk Q1, dit_7b/nadit.py:152-191): oracle/nadit.py against the
    output of the real reference model (tests/golden/stack_*.pt)."""
    from oracle import nadit as onadit
    import seedvr2_b200
    gold = torch.load(os.path.join(GOLD, f"stack_{case.name}.pt"))
    vid, txt = cases.stack_inputs(case)
    assert gold["in_digest"] == {"vid": cases.tensor_digest(vid), "txt": cases.tensor_digest(txt)}
    st = seedvr2_b200.stack.NaDiTStack(case.model, num_layers=case.num_layers, vid_dim=case.dim, heads=case.heads,
                                       mm_layers=case.mm_layers, txt_in_dim=case.txt_in_dim)
    params = cases.stack_params(case, {k: tuple(v.shape) for k, v in st.state_dict().items()})
    kw = dict(dim=case.dim, heads=case.heads, num_layers=case.num_layers, mm_layers=case.mm_layers)
    with torch.no_grad():
        o32 = onadit.nadit_forward(case.model, params, vid, txt, case.latent_grid, case.timestep, **kw)
        p16 = {k: (v if k.endswith("rope.freqs") else v.bfloat16()) for k, v in params.items()}
        o16 = onadit.nadit_forward(case.model, p16, vid.bfloat16(), txt.bfloat16(), case.latent_grid, case.timestep, **kw)
    assert float((o32 - gold["out_fp32"]).abs().max()) < 1e-4
    assert cases.psnr(o16.float(), gold["out_bf16"].float()) > 80.0      # same ops, same order: bit-equal where run
    # the north star's decoded-output bar applied to the reference's OWN bf16 path: it must clear 45 dB for the
    # bar to be meaningful at these magnitudes
    assert cases.psnr(gold["out_bf16"].float(), gold["out_fp32"]) > 45.0


# ---- one-step SR call around the DiT forward ---------------------------------------------------------------
@pytest.mark.parametrize("case", cases.SAMPLER_CASES, ids=lambda c: c.name)
def test_sampler_oracle_matches_reference(case):
    """oracle/sampler.py against the real VideoDiffusionInfer.get_condition / timestep_transform / inference + EulerSampler
    (tests/golden/sampler_*.pt): condition, DiT input (channel order of the concat) and x_0, bit for bit."""
    from oracle import sampler as osamp
    gold = torch.load(os.path.join(GOLD, f"sampler_{case.name}.pt"))
    latent, base_noise, aug_noise, mix = cases.sampler_inputs(case)
    blur = osamp.add_noise(latent, aug_noise, case.latent_noise_scale)
    assert blur.dtype == gold["latent_blur"].dtype and torch.equal(blur, gold["latent_blur"])
    cond = osamp.sr_condition(base_noise, blur)
    assert torch.equal(cond, gold["condition"])
    vid, x0 = osamp.one_step(cases.standin_dit(mix), base_noise, cond)
    assert torch.equal(vid, gold["dit_input"])
    assert x0.dtype == gold["x0"].dtype and torch.equal(x0, gold["x0"])
    assert gold["timestep"].tolist() == [1000.0] and gold["vid_shape"].tolist() == [list(case.grid)]


def test_postprocess_oracle_matches_reference_module():
    """oracle/postprocess.py against the real src/utils/color_fix.py:wavelet_reconstruction (staged copy or /root/reference), CPU."""
    import torch
    from oracle import postprocess as opost
    from oracle import refstage
    if not refstage.available():
        pytest.skip("reference not staged")
    refstage.activate()
    from src.utils.color_fix import wavelet_reconstruction
    g = torch.Generator().manual_seed(3)
    for shape, dt in (((2, 3, 40, 56), torch.float32), ((1, 3, 20, 37), torch.float32), ((2, 3, 40, 56), torch.bfloat16)):
        a = (torch.rand(*shape, generator=g) * 2.6 - 1.3).to(dt)
        b = (torch.rand(*shape, generator=g) * 2.2 - 1.1).to(dt)
        ref = wavelet_reconstruction(a.clone(), b.clone(), None).permute(0, 2, 3, 1).contiguous().clamp_(-1, 1).mul_(0.5).add_(0.5)
        assert torch.equal(opost.wavelet_color_fix(a, b), ref)
    # the CLI's default, lab_color_transfer (color_fix.py:249-367)
    from src.utils.color_fix import lab_color_transfer

    class _D:
        def log(self, *a, **k):
            pass
    for dt in (torch.float32, torch.bfloat16):
        a = (torch.rand(2, 3, 40, 56, generator=g) * 2.4 - 1.2).to(dt)
        b = (torch.rand(2, 3, 40, 56, generator=g) * 2.0 - 1.0).to(dt)
        ref = lab_color_transfer(a.clone(), b.clone(), _D(), luminance_weight=0.8).permute(0, 2, 3, 1).contiguous().clamp_(-1, 1).mul_(0.5).add_(0.5)
        assert torch.equal(opost.lab_color_fix(a, b), ref)


def test_vae_conv_oracle_matches_reference_module():
    """oracle/vae_conv.py against the real InflatedCausalConv3d (staged copy) on the CPU: the three memory states, temporal slices of
    several lengths (one frame included), kernels (3,3,3) / (1,3,3) / (1,1,1), and the memory-limited spatially split path."""
    import torch
    from oracle import refstage, vae_conv as ovc
    if not refstage.available():
        pytest.skip("reference not staged")
    refstage.activate()
    from src.models.video_vae_v3.modules.causal_inflation_lib import init_causal_conv3d
    from src.models.video_vae_v3.modules.types import MemoryState
    torch.manual_seed(5)
    for kernel, pad in ((3, 1), ((1, 3, 3), (0, 1, 1)), (1, 0)):
        for limit in (float("inf"), 1.2e-5):      # GiB: two to three spatial splits of these tensors
            conv = init_causal_conv3d(8, 12, kernel_size=kernel, stride=1, padding=pad, inflation_mode="tail").eval()
            conv.set_memory_limit(limit)
            x = torch.randn(1, 8, 9, 10, 12)
            with torch.no_grad():
                ref = conv(x.clone(), memory_state=MemoryState.DISABLED)
                got, _ = ovc.causal_conv3d(x, conv.weight, conv.bias)
                assert ref.shape == got.shape and torch.allclose(ref, got, atol=2e-6, rtol=0), (kernel, limit)
                # streaming: the slices of the VAE's temporal slicing, memory carried between them
                mem = None
                outs_ref, outs_got = [], []
                for i, (a, b) in enumerate(((0, 4), (4, 5), (5, 9))):
                    st = MemoryState.INITIALIZING if i == 0 else MemoryState.ACTIVE
                    outs_ref.append(conv(x[:, :, a:b].clone(), memory_state=st))
                    y, mem = ovc.causal_conv3d(x[:, :, a:b], conv.weight, conv.bias, mem)
                    outs_got.append(y)
                    if conv.kernel_size[0] > 1:
                        assert torch.equal(conv.memory, mem), (kernel, limit, i)
                assert torch.allclose(torch.cat(outs_ref, 2), torch.cat(outs_got, 2), atol=2e-6, rtol=0)
                # and slicing changes nothing: the streamed result is the one-shot result
                assert torch.allclose(torch.cat(outs_got, 2), got, atol=2e-6, rtol=0)


def test_vae_downsample_oracle_matches_reference_module():
    """oracle/vae_conv.causal_conv3d_down against the reference's strided InflatedCausalConv3d behind Downsample3D's padding
    (attn_video_vae.py:239-245), CPU: spatial and spatial + temporal down-sampling, streamed slices of 5, 4 and 2 frames."""
    import torch
    import torch.nn.functional as F
    from oracle import refstage, vae_conv as ovc
    if not refstage.available():
        pytest.skip("reference not staged")
    refstage.activate()
    from src.models.video_vae_v3.modules.causal_inflation_lib import init_causal_conv3d
    from src.models.video_vae_v3.modules.types import MemoryState
    torch.manual_seed(6)
    for temporal in (False, True):
        tk, tr = (3, 2) if temporal else (1, 1)
        conv = init_causal_conv3d(8, 12, kernel_size=(tk, 3, 3), stride=(tr, 2, 2), padding=(1 if temporal else 0, 0, 0),
                                  inflation_mode="tail").eval()
        x = torch.randn(1, 8, 11, 9, 12)
        with torch.no_grad():
            mem, outs_r, outs_o = None, [], []
            for i, (a, b) in enumerate(((0, 5), (5, 9), (9, 11))):
                st = MemoryState.INITIALIZING if i == 0 else MemoryState.ACTIVE
                outs_r.append(conv(F.pad(x[:, :, a:b], (0, 1, 0, 1)), memory_state=st))
                y, mem = ovc.causal_conv3d_down(x[:, :, a:b], conv.weight, conv.bias, mem)
                outs_o.append(y)
                if temporal:
                    assert torch.equal(conv.memory, F.pad(mem, (0, 1, 0, 1)))
            r, o = torch.cat(outs_r, 2), torch.cat(outs_o, 2)
            assert r.shape == o.shape == (1, 12, 6 if temporal else 11, 4, 6)
            assert torch.allclose(r, o, atol=2e-6, rtol=0)


def test_vae_conv_oracle_matches_golden_vectors():
    """oracle/vae_conv.py against tests/golden/vae_conv.pt (outputs and memories of the real InflatedCausalConv3d, fp32 CPU, made by
    tests/golden/make_golden_vae.py): needs no reference checkout."""
    import torch
    from oracle import vae_conv as ovc
    gold = torch.load(os.path.join(os.path.dirname(__file__), "golden", "vae_conv.pt"))
    assert set(gold) == {"conv333", "conv133", "down_spatial", "down_temporal"}
    for name, c in gold.items():
        w, b, x = c["weight"].float(), c["bias"].float(), c["x"].float()
        mem = None
        for (a, e), y_ref, m_ref in zip(c["slices"], c["y"], c["memory"]):
            if tuple(c["stride"]) == (1, 1, 1):
                y, mem = ovc.causal_conv3d(x[:, :, a:e], w, b, mem)
            else:
                y, mem = ovc.causal_conv3d_down(x[:, :, a:e], w, b, mem, c["pad_hw"])
            assert y.shape == y_ref.shape and torch.allclose(y, y_ref, atol=3e-6, rtol=0), name
            if m_ref is not None:
                # the strided module keeps its memory with Downsample3D's padding column / row attached
                assert torch.equal(m_ref[..., :mem.shape[-2], :mem.shape[-1]], mem), name
