"""CPU: the drop-in blocks keep the reference's constructor kwargs and state-dict contract (names + shapes), are
meta-device safe, and the product refuses to run without CUDA instead of falling back."""
import pytest
import torch

import cases


def make_block(svr2, c: cases.BlockCase, device=None):
    method = "720pswin_by_size_bysize" if c.shifted else "720pwin_by_size_bysize"
    common = dict(vid_dim=c.dim, txt_dim=c.dim, emb_dim=6 * c.dim, heads=c.heads, head_dim=c.head_dim, expand_ratio=4,
                  norm=None, norm_eps=1e-5, ada=None, qk_bias=False, qk_norm=None, window=c.window, window_method=method,
                  attention_mode="sdpa")
    if c.model == "3b":
        blk = svr2.NaMMSRTransformerBlockB200(**common, mlp_type="swiglu", shared_weights=c.shared, rope_type="mmrope3d",
                                              rope_dim=128, is_last_layer=c.is_last, msa_type=None, mca_type=None)
    else:
        blk = svr2.NaMMSRTransformerBlockB200_7B(**common, qk_rope=True, shared_qkv=False, shared_mlp=False,
                                                 mlp_type="normal", temporal_window_size=None, temporal_shifted=None)
    return blk


@pytest.mark.parametrize("case", cases.BLOCK_CASES, ids=lambda c: c.name)
def test_state_dict_contract(svr2, case):
    blk = make_block(svr2, case)
    got = {k: tuple(v.shape) for k, v in blk.state_dict().items()}
    assert got == cases.block_param_shapes(case)       # the same dict make_golden.py asserted against the reference
    assert torch.equal(blk.state_dict()["attn.rope.rope.freqs"], cases.rope_freqs(case.model))
    blk.load_state_dict(cases.block_params(case), strict=True)


def test_meta_device_construction(svr2):
    """model_loader.py:437-438 builds the model under torch.device('meta') and fills it with assign=True."""
    c = cases.BLOCK_CASES[0]
    with torch.device("meta"):
        blk = make_block(svr2, c)
    assert all(p.is_meta for p in blk.parameters())
    blk.load_state_dict({k: v.bfloat16() if not k.endswith("freqs") else v for k, v in cases.block_params(c).items()},
                        strict=True, assign=True)
    assert not any(p.is_meta for p in blk.parameters()) and blk.attn.proj_qkv.vid.weight.dtype == torch.bfloat16


def test_no_cpu_fallback(svr2):
    c = cases.BLOCK_CASES[0]
    blk = make_block(svr2, c).bfloat16()
    vid, txt, emb = (t.bfloat16() for t in cases.block_inputs(c))
    with pytest.raises((RuntimeError, NotImplementedError)):
        blk(vid=vid, txt=txt, vid_shape=torch.tensor([list(c.grid)]), txt_shape=torch.tensor([[c.txt_len]]), emb=emb, cache=None)


def test_unsupported_configs_raise(svr2):
    with pytest.raises(NotImplementedError):
        svr2.NaMMSRTransformerBlockB200(vid_dim=128, txt_dim=128, emb_dim=768, heads=2, head_dim=64, expand_ratio=4,
                                        window=(4, 3, 3), window_method="720pwin_by_size_bysize")
    with pytest.raises(NotImplementedError):
        svr2.NaMMSRTransformerBlockB200(vid_dim=128, txt_dim=128, emb_dim=768, heads=2, head_dim=128, expand_ratio=4,
                                        window=(4, 3, 3), window_method="win")


def test_non_bf16_parameters_are_cast_once(svr2):
    """assign=True keeps the checkpoint's dtype (fp16 / fp32 / fp8 files, model_loader.py:809): the blocks use a bf16 copy made
    once, refreshed when the parameter is replaced, written in place or moved."""
    from seedvr2_b200.block import _bf16
    p = torch.nn.Parameter(torch.randn(8, 8).half())
    a = _bf16(p)
    assert a.dtype == torch.bfloat16 and _bf16(p) is a
    with torch.no_grad():
        p.mul_(2)
    b = _bf16(p)
    assert b is not a and torch.equal(b, p.detach().to(torch.bfloat16))
    q = torch.nn.Parameter(torch.randn(4).bfloat16())
    assert _bf16(q) is q


def test_convert_parameters_in_place(svr2):
    """fp16 checkpoint parameters (assign=True keeps the file's dtype) can be rewritten to bf16 in place: one copy instead of the
    original + the block's lazily cast bf16 copy; the checkpoint's rope frequencies keep their dtype."""
    c = cases.BLOCK_CASES[0]
    with torch.device("meta"):
        blk = make_block(svr2, c)
    sd = {k: (v if k.endswith("freqs") else v.half()) for k, v in cases.block_params(c).items()}
    blk.load_state_dict(sd, strict=True, assign=True)
    holder = torch.nn.ModuleList([blk])
    n = svr2.convert_parameters_(holder)
    assert n == sum(1 for _ in blk.parameters()) and all(p.dtype == torch.bfloat16 for p in blk.parameters())
    assert blk.attn.rope.rope.freqs.dtype == torch.float32
    assert svr2.convert_parameters_(holder) == 0
