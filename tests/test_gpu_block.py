"""GPU parity of the drop-in blocks against the golden outputs of the REAL reference block (tests/golden/block_*.pt):
same parameters (load_state_dict), same inputs, bf16.  Tolerance = the north star's: max-abs <= 2e-2 per block output
against the reference's own bf16 path; the attention core (before proj_out) is held to the same bound."""
import os

import pytest
import torch

import cases
from test_block_contract import make_block

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
TOL = 2e-2


class RefCache:
    """Same call protocol as the reference's src/common/cache.py Cache."""

    def __init__(self):
        self.cache = {}

    def __call__(self, key, fn):
        if key not in self.cache:
            self.cache[key] = fn()
        return self.cache[key]

    def namespace(self, ns):
        return self


@pytest.mark.parametrize("case", cases.BLOCK_CASES, ids=lambda c: c.name)
def test_block_vs_reference_golden(svr2, cuda, case):
    gold = torch.load(os.path.join(GOLD, f"block_{case.name}.pt"))
    blk = make_block(svr2, case)
    blk.load_state_dict(cases.block_params(case), strict=True)
    blk = blk.bfloat16().to(cuda)
    blk.attn.rope.rope.freqs.data = cases.rope_freqs(case.model).to(cuda)     # checkpoint buffer stays fp32
    vid, txt, emb = (t.bfloat16().to(cuda) for t in cases.block_inputs(case))
    vid_shape = torch.tensor([list(case.grid)], device=cuda)
    txt_shape = torch.tensor([[case.txt_len]], device=cuda)
    captured = {}
    orig_core = blk.attn.core
    blk.attn.core = lambda *a, **k: captured.setdefault("core", orig_core(*a, **k))
    cache = RefCache()
    with torch.no_grad():
        v, t, vs, ts = blk(vid=vid, txt=txt, vid_shape=vid_shape, txt_shape=txt_shape, emb=emb, cache=cache)
    torch.cuda.synchronize()
    assert vs is vid_shape and ts is txt_shape
    av, at = captured["core"]
    d_av = float((av.cpu().float() - gold["attn_vid_bf16"].float()).abs().max())
    d_at = float((at.cpu().float() - gold["attn_txt_bf16"].float()).abs().max())
    d_v = float((v.cpu().float() - gold["vid_out_bf16"].float()).abs().max())
    d_t = float((t.cpu().float() - gold["txt_out_bf16"].float()).abs().max())
    e_v = float((v.cpu().float() - gold["vid_out_fp32"]).abs().max())
    r_v = float((gold["vid_out_bf16"].float() - gold["vid_out_fp32"]).abs().max())
    print(f"{case.name}: attn vid {d_av:.4f} txt {d_at:.4f} | block vid {d_v:.4f} txt {d_t:.4f} | vs fp32: ours {e_v:.4f} ref-bf16 {r_v:.4f}")
    assert d_av <= TOL and d_at <= TOL, "attention core exceeds 2e-2 vs the reference bf16 path"
    assert d_v <= TOL and d_t <= TOL, "block output exceeds 2e-2 vs the reference bf16 path"
    # quirk Q1: the blocks leave the ATTN-layer (shift, scale, gate) triple in the cache for vid_out_ada
    if case.model == "3b":
        rep = cache.cache["emb_repeat_0_vid"]
        assert rep.shape == (vid.shape[0], case.dim, 3)
        assert torch.equal(rep[0], emb.reshape(1, -1, 2, 3)[0, :, 0, :])


def test_block_with_fp16_checkpoint_parameters(svr2, cuda):
    """A checkpoint loaded with assign=True keeps its dtype (the shipped files are fp16, model_registry.py:34-49); the pipeline
    then runs the model under bf16 autocast.  The block casts such parameters to bf16 once and must give the same output as
    with bf16 parameters (the test parameters are bf16-representable, so fp16 holds them exactly)."""
    case = cases.BLOCK_CASES[1]
    outs = []
    for dt in (torch.bfloat16, torch.float16):
        with torch.device("meta"):
            blk = make_block(svr2, case)
        sd = {k: (v if k.endswith("freqs") else v.to(dt)).to(cuda) for k, v in cases.block_params(case).items()}
        blk.load_state_dict(sd, strict=True, assign=True)
        vid, txt, emb = (t.bfloat16().to(cuda) for t in cases.block_inputs(case))
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            v, t, _, _ = blk(vid=vid, txt=txt, vid_shape=torch.tensor([list(case.grid)], device=cuda),
                             txt_shape=torch.tensor([[case.txt_len]], device=cuda), emb=emb, cache=None)
        torch.cuda.synchronize()
        outs.append((v.clone(), t.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
