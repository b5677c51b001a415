"""GPU: the ACTUAL drop-in — the reference's own ``NaDiT`` (staged unmodified under baseline/_ref by scripts/stage_reference.py)
runs on the B200 with the B200 blocks that ``seedvr2_b200.install()`` put into its block registry.

Everything outside the blocks is the reference's code: ``NaDiT.__init__`` / ``.forward`` (dit_3b/nadit.py:39-248,
dit_7b/nadit.py:39-191) with its real ``Cache`` (src/common/cache.py), ``NaPatchIn/Out``, ``TimeEmbedding``, output norm + ada,
wrapped in the product's ``FP8CompatibleDiT`` (compatibility.py:231-660: RoPE "stabilisation" + attention forward patching) and
called under ``torch.autocast(bf16)`` + ``no_grad`` as generation_phases.py:713-714 / infer.py:293 do.  Construction follows
model_loader.py:437-438,809: meta device, then ``load_state_dict(assign=True)``.

Checked against (1) the golden outputs of the stock reference model (tests/golden/stack_*.pt, >= 45 dB PSNR — the north star's
bar; on the predicted latent), (2) ``NaDiTStack``, the repo's own shell around the same blocks, and (3) for one
full-width block, the stock reference block run on the same GPU.  Skips where the reference is not staged."""
import os

import pytest
import torch

import cases
from oracle import refstage

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not refstage.available(), reason="reference not staged under baseline/_ref")]
GOLD = os.path.join(os.path.dirname(__file__), "golden")


class _Debug:
    """Duck-typed stand-in for the product's Debug logger (src/utils/debug.py) — FP8CompatibleDiT only logs and times."""

    def log(self, *a, **k):
        pass

    def start_timer(self, *a, **k):
        pass

    def end_timer(self, *a, **k):
        return 0.0


def _load_like_model_loader(model, params, dev):
    sd = {k: (v if k.endswith("freqs") else v.bfloat16()).to(dev) for k, v in params.items()}
    missing, unexpected = model.load_state_dict(sd, strict=False, assign=True)
    assert not missing and not unexpected, (missing, unexpected)
    # model_loader.py:763-801: whatever is still on the meta device after the load is a non-persistent buffer -> zero-filled
    for mod in model.modules():
        for name, buf in list(mod._buffers.items()):
            if buf is not None and buf.is_meta:
                mod._buffers[name] = torch.zeros(buf.shape, dtype=buf.dtype, device=dev)
    assert not any(p.is_meta for p in model.parameters())
    return model.eval()


@pytest.mark.parametrize("case", cases.STACK_CASES, ids=lambda c: c.name)
def test_reference_nadit_runs_installed_blocks_on_gpu(svr2, cuda, case):
    NaDiT = refstage.nadit_class(case.model)
    from src.optimization.compatibility import FP8CompatibleDiT
    gold = torch.load(os.path.join(GOLD, f"stack_{case.name}.pt"))
    svr2.install()
    try:
        with torch.device("meta"):
            model = NaDiT(**cases.stack_kwargs(case))
        ours_cls = svr2.NaMMSRTransformerBlockB200 if case.model == "3b" else svr2.NaMMSRTransformerBlockB200_7B
        assert all(type(b) is ours_cls for b in model.blocks)
        shapes = {k: tuple(v.shape) for k, v in model.state_dict().items()}
        params = cases.stack_params(case, shapes)
        _load_like_model_loader(model, params, cuda)
        wrapped = FP8CompatibleDiT(model, _Debug(), compute_dtype=torch.bfloat16)
        vid, txt = cases.stack_inputs(case)
        assert cases.tensor_digest(vid) == gold["in_digest"]["vid"]
        vid, txt = vid.bfloat16().to(cuda), txt.bfloat16().to(cuda)
        vid_shape = torch.tensor([list(case.latent_grid)], device=cuda)          # lives on the GPU in the pipeline (na.py:97)
        txt_shape = torch.tensor([[case.txt_len]], device=cuda)
        ts = torch.tensor([case.timestep], device=cuda)
        kw = dict(disable_cache=False) if case.model == "3b" else {}
        n0 = svr2.ops.launch_count()
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            out = wrapped(vid, txt, vid_shape, txt_shape, ts, **kw).vid_sample
            out2 = wrapped(vid, txt, vid_shape.clone(), txt_shape.clone(), ts, **kw).vid_sample     # second call: memoised plans
        torch.cuda.synchronize()
        assert svr2.ops.launch_count() - n0 >= 2 * 6 * case.num_layers, "the forward did not go through the C-ABI kernels"
        assert torch.equal(out, out2)
    finally:
        svr2.uninstall()
    o = out.float().cpu()
    p32 = cases.psnr(o, gold["out_fp32"])
    p16 = cases.psnr(o, gold["out_bf16"].float())
    ref_own = cases.psnr(gold["out_bf16"].float(), gold["out_fp32"])
    # the repo's own shell around the same blocks, same parameters
    stack = svr2.NaDiTStack(case.model, num_layers=case.num_layers, vid_dim=case.dim, heads=case.heads, mm_layers=case.mm_layers,
                            txt_in_dim=case.txt_in_dim)
    stack.load_state_dict(params, strict=True)
    stack = stack.bfloat16().to(cuda).eval()
    for m in stack.modules():
        if hasattr(m, "freqs") and isinstance(m.freqs, torch.Tensor):
            m.freqs.data = cases.rope_freqs(case.model).to(cuda)
    with torch.no_grad():
        so = stack(vid, txt, vid_shape, torch.tensor([[case.txt_len]]), ts)
    torch.cuda.synchronize()
    d_stack = float((so.float() - out.float()).abs().max())
    print(f"\n{case.name}: reference NaDiT + install() on GPU: PSNR {p32:.1f} dB vs stock fp32, {p16:.1f} dB vs stock bf16 "
          f"(stock bf16-vs-fp32 {ref_own:.1f} dB); max|d| vs NaDiTStack {d_stack:.4f}")
    assert p32 >= 45.0 and p16 >= 45.0
    assert cases.psnr(so.float().cpu(), o) >= 60.0, "NaDiTStack and the reference shell disagree around the same blocks"


@pytest.mark.parametrize("model,grid,shared,shifted", [("3b", (1, 128, 128), False, False), ("3b", (9, 68, 120), True, True),
                                                       ("7b", (6, 68, 120), False, True)],
                         ids=["3b_c1_sep_win", "3b_1080p33f_shared_swin", "7b_c3r_swin"])
def test_fullwidth_block_vs_stock_reference_block_on_gpu(svr2, cuda, model, grid, shared, shifted):
    """One block of the real architecture (configs_3b / configs_7b widths) on a BASELINE grid: the B200 block against the STOCK
    reference block (imported from the staged copy, bf16, attention_mode='sdpa') on the same GPU, same parameters and inputs;
    and the oracle against the stock block, which carries the oracle's pin to full width."""
    from oracle import block as oblock
    from oracle import workload as owork
    from test_gpu_fullwidth import _product_block
    cfg = owork.block_cfg(model, shared=shared, shifted=shifted)
    params = owork.block_params(cfg, seed=3)
    sd = {k: (v.float() if k.endswith("freqs") else v.bfloat16()).to(cuda) for k, v in params.items()}
    with torch.device("meta"):
        ref = refstage.build_block(model, cfg.dim, cfg.heads, shared=cfg.shared, shifted=shifted)
    ref.load_state_dict(sd, strict=True, assign=True)
    ref.eval()
    blk = _product_block(svr2, cfg, params, cuda)
    vid, txt, emb = owork.block_inputs(cfg, grid, 58, seed=5)
    vid, txt, emb = (0.25 * vid).bfloat16().to(cuda), (0.25 * txt).bfloat16().to(cuda), (0.5 * emb).bfloat16().to(cuda)
    vid_shape = torch.tensor([list(grid)], device=cuda)
    txt_shape = torch.tensor([[58]], device=cuda)
    with torch.no_grad():
        rv, rt, _, _ = ref(vid=vid.clone(), txt=txt.clone(), vid_shape=vid_shape, txt_shape=txt_shape, emb=emb.clone(),
                           cache=refstage.new_cache())
        v, t, _, _ = blk(vid=vid, txt=txt, vid_shape=vid_shape, txt_shape=txt_shape, emb=emb, cache=refstage.new_cache())
        ov, ot = oblock.block_forward(cfg, sd, vid, txt, grid, emb)
    torch.cuda.synchronize()
    d_v, d_t = float((v.float() - rv.float()).abs().max()), float((t.float() - rt.float()).abs().max())
    o_v, o_t = float((ov.float() - rv.float()).abs().max()), float((ot.float() - rt.float()).abs().max())
    print(f"\n{model} {grid} shared={shared} shifted={shifted}: B200 block vs stock reference block (bf16, GPU): vid {d_v:.4f} txt {d_t:.4f}"
          f" | oracle vs stock: vid {o_v:.4f} txt {o_t:.4f} | |out|max {float(rv.float().abs().max()):.2f}")
    assert d_v <= 2e-2 and d_t <= 2e-2
    assert o_v <= 2e-2 and o_t <= 2e-2, "the oracle has drifted from the stock reference block at full width"


@pytest.mark.parametrize("model", ["3b", "7b"])
def test_reference_nadit_full_width_installed_vs_stock_on_gpu(svr2, cuda, model):
    """The drop-in at the REAL widths (configs_3b: D = 2560, 20 heads, SwiGLU 6912, separate + shared + last layer; configs_7b:
    D = 3072, 24 heads, GELU 12288; 4 layers of each): the reference's NaDiT built twice from the same kwargs and the same random
    state dict — once with the stock blocks, once after install() — both in bf16 on the same GPU, real Cache, autocast.  The two
    outputs differ only by what the blocks do: PSNR >= 45 dB (north star) and no further from the fp32 stock model than the stock
    bf16 model is."""
    import math
    NaDiT = refstage.nadit_class(model)
    mm = 2 if model == "3b" else None
    kw = refstage.nadit_kwargs(model, num_layers=4, mm_layers=mm, txt_in_dim=512)
    torch.manual_seed(7)
    with torch.device("meta"):
        probe = NaDiT(**kw)
    g = torch.Generator().manual_seed(17)
    sd = {}
    for k, v in probe.state_dict().items():
        if k.endswith("freqs"):
            sd[k] = cases.rope_freqs(model)
        elif k.endswith(("_scale", "norm_q.vid.weight", "norm_k.vid.weight", "norm_q.txt.weight", "norm_k.txt.weight", "norm_q.all.weight",
                         "norm_k.all.weight", "vid_out_norm.weight")):
            sd[k] = (1.0 + 0.05 * torch.randn(v.shape, generator=g)).bfloat16()
        elif v.dim() >= 2:
            sd[k] = (torch.randn(v.shape, generator=g) / math.sqrt(v.shape[-1])).bfloat16()
        else:
            sd[k] = (0.02 * torch.randn(v.shape, generator=g)).bfloat16()
    grid = (2, 48, 80)                       # latent grid -> 2 x 24 x 40 tokens
    vid = torch.randn(grid[0] * grid[1] * grid[2], 33, generator=g).bfloat16().to(cuda)
    txt = (0.25 * torch.randn(58, 512, generator=g)).bfloat16().to(cuda)
    vs = torch.tensor([list(grid)], device=cuda)
    ts = torch.tensor([[58]], device=cuda)
    t = torch.tensor([1000.0], device=cuda)
    fkw = dict(disable_cache=False) if model == "3b" else {}

    def build(dtype):
        with torch.device("meta"):
            m = NaDiT(**kw)
        state = {k: (v.float() if k.endswith("freqs") else v.to(dtype)).to(cuda) for k, v in sd.items()}
        missing, unexpected = m.load_state_dict(state, strict=False, assign=True)
        assert not missing and not unexpected
        for mod in m.modules():
            for name, buf in list(mod._buffers.items()):
                if buf is not None and buf.is_meta:
                    mod._buffers[name] = torch.zeros(buf.shape, dtype=buf.dtype, device=cuda)
        return m.eval()

    with torch.no_grad():
        stock32 = build(torch.float32)(vid.float(), txt.float(), vs, ts, t, **fkw).vid_sample
        stock16 = build(torch.bfloat16)(vid, txt, vs, ts, t, **fkw).vid_sample
        svr2.install()
        try:
            ours_model = build(torch.bfloat16)
            ours_cls = svr2.NaMMSRTransformerBlockB200 if model == "3b" else svr2.NaMMSRTransformerBlockB200_7B
            assert all(type(b) is ours_cls for b in ours_model.blocks)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                ours = ours_model(vid, txt, vs, ts, t, **fkw).vid_sample
        finally:
            svr2.uninstall()
    torch.cuda.synchronize()
    p_vs16 = cases.psnr(ours.float().cpu(), stock16.float().cpu())
    p_vs32 = cases.psnr(ours.float().cpu(), stock32.cpu())
    p_ref = cases.psnr(stock16.float().cpu(), stock32.cpu())
    print(f"\n{model} full width, 4 layers: install()ed vs stock bf16 {p_vs16:.1f} dB | vs stock fp32: ours {p_vs32:.1f} dB, stock bf16 {p_ref:.1f} dB")
    assert p_vs16 >= 45.0 and p_vs32 >= 45.0
    assert p_vs32 >= p_ref - 1.0
