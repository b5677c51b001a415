"""CPU, build container only: the drop-in boundary exercised with the REAL reference.

``seedvr2_b200.install()`` swaps ``nadit_blocks["mmdit_sr"]`` in both reference packages; the reference's own
``NaDiT.__init__`` (dit_3b/nadit.py:135-165, dit_7b/nadit.py:109-136) must then build B200 blocks from its unchanged
kwargs (factory callables for norm / ada / qk_norm included), expose the SAME state dict as the stock model, build under
``torch.device("meta")`` and load with ``assign=True`` like model_loader.py:437-438,809 does.  The reference is mounted
read-only at /root/reference in this container and absent on the GPU box, where these tests skip."""
import os
import sys

import pytest
import torch

import cases

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src", "models")), reason="reference not mounted")


@pytest.fixture()
def ref_models(svr2):
    shim = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "shim")
    added = [p for p in (shim, REF) if p not in sys.path]
    sys.path[0:0] = added
    try:
        yield
    finally:
        svr2.uninstall()
        for p in added:
            sys.path.remove(p)


def _nadit(model):
    if model == "3b":
        from src.models.dit_3b.nadit import NaDiT
    else:
        from src.models.dit_7b.nadit import NaDiT
    return NaDiT


@pytest.mark.parametrize("case", cases.STACK_CASES, ids=lambda c: c.name)
def test_install_swaps_blocks_inside_reference_nadit(svr2, ref_models, case):
    NaDiT = _nadit(case.model)
    stock = NaDiT(**cases.stack_kwargs(case))
    want = {k: (tuple(v.shape), v.dtype) for k, v in stock.state_dict().items()}
    stock_cls = type(stock.blocks[0])
    svr2.install()
    ours_cls = svr2.NaMMSRTransformerBlockB200 if case.model == "3b" else svr2.NaMMSRTransformerBlockB200_7B
    model = NaDiT(**cases.stack_kwargs(case))
    assert all(type(b) is ours_cls for b in model.blocks) and stock_cls is not ours_cls
    got = {k: (tuple(v.shape), v.dtype) for k, v in model.state_dict().items()}
    assert got == want, sorted(set(got.items()) ^ set(want.items()))
    # checkpoints of the stock model load unchanged, including under meta construction + assign=True
    sd = {k: v.clone() for k, v in stock.state_dict().items()}
    model.load_state_dict(sd, strict=True)
    with torch.device("meta"):
        meta = NaDiT(**cases.stack_kwargs(case))
    assert all(p.is_meta for p in meta.parameters())
    meta.load_state_dict({k: (v if k.endswith("freqs") else v.bfloat16()) for k, v in sd.items()}, strict=True, assign=True)
    assert not any(p.is_meta for p in meta.parameters())
    assert not any(b.is_meta for b in meta.buffers())
    # the reference's forward now reaches the B200 blocks: on a CPU-only host they refuse instead of falling back
    vid, txt = cases.stack_inputs(case)
    with torch.no_grad(), pytest.raises((RuntimeError, NotImplementedError)):
        meta.bfloat16()(vid.bfloat16(), txt.bfloat16(), torch.tensor([list(case.latent_grid)]),
                        torch.tensor([[case.txt_len]]), torch.tensor([case.timestep]))
    svr2.uninstall()
    assert type(NaDiT(**cases.stack_kwargs(case)).blocks[0]) is stock_cls


def test_block_tolerates_product_monkey_patching(svr2, ref_models):
    """compatibility.py:437-522 replaces ``forward`` of every module whose name contains 'attn' with a wrapper calling the
    saved original; blockswap.py:415-453 wraps ``block.forward`` and moves blocks with ``.to()``.  Both must leave the
    call chain and the parameters intact."""
    c = cases.BLOCK_CASES[0]
    from test_block_contract import make_block
    blk = make_block(svr2, c)
    calls = []
    for name, mod in blk.named_modules():
        if "attn" in name.lower() and hasattr(mod, "forward") and type(mod).__name__ == "NaSwinAttentionB200":
            mod._original_forward = mod.forward
            mod.forward = (lambda m: (lambda *a, **k: (calls.append("attn"), m._original_forward(*a, **k))[1]))(mod)
    orig = blk.forward
    blk.forward = lambda *a, **k: (calls.append("block"), orig(*a, **k))[1]
    blk = blk.to(torch.bfloat16)
    assert all(isinstance(p, torch.nn.Parameter) for p in blk.parameters())
    vid, txt, emb = (t.bfloat16() for t in cases.block_inputs(c))
    with pytest.raises((RuntimeError, NotImplementedError)):     # CPU tensors: reaches the C-ABI wrapper and is refused there
        blk(vid=vid, txt=txt, vid_shape=torch.tensor([list(c.grid)]), txt_shape=torch.tensor([[c.txt_len]]), emb=emb, cache=None)
    assert calls[0] == "block"
