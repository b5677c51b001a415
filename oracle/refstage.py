"""Access to the REAL reference staged under baseline/_ref (scripts/stage_reference.py).  TEST INFRASTRUCTURE ONLY.

Used by the GPU parity tests and by bench.py's baseline legs (`--impl reference`, `cpu_baseline`, the GPU-vs-GPU
`ref_torch_gpu` figure) — never by the product.  The reference's modules are imported unmodified from the staged copy; the
three third-party packages it needs and the image lacks come from oracle/shim (restatements of their published behaviour;
`rotary_embedding_torch` is the one arithmetic boundary that stays unpinned).
"""
import os
import sys
from typing import Dict, Optional, Tuple

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STAGED = os.path.join(ROOT, "baseline", "_ref")
SHIM = os.path.join(ROOT, "oracle", "shim")


def available() -> bool:
    return os.path.isdir(os.path.join(STAGED, "src", "models", "dit_3b"))


def activate() -> None:
    """Put the staged reference (package ``src``) and the shims on sys.path; idempotent."""
    if not available():
        raise RuntimeError("the reference is not staged under baseline/_ref (run scripts/stage_reference.py in the build container)")
    os.environ.setdefault("SEEDVR2_OPTIMIZATIONS_LOGGED", "1")      # silence the import-time banner of compatibility.py
    for p in (STAGED, SHIM):
        if p not in sys.path:
            sys.path.insert(0, p)


def real_rotary_embedding_torch() -> bool:
    """True when the box has the genuine lucidrains package (then the shim is NOT used and the boundary is pinned)."""
    import importlib.util
    saved = [p for p in sys.path]
    try:
        sys.path[:] = [p for p in sys.path if os.path.abspath(p) != SHIM]
        return importlib.util.find_spec("rotary_embedding_torch") is not None
    finally:
        sys.path[:] = saved


def nadit_class(model: str):
    activate()
    if model == "3b":
        from src.models.dit_3b.nadit import NaDiT
    else:
        from src.models.dit_7b.nadit import NaDiT
    return NaDiT


def nadit_kwargs(model: str, num_layers: Optional[int] = None, dim: Optional[int] = None, heads: Optional[int] = None,
                 mm_layers: Optional[int] = None, txt_in_dim: int = 5120) -> Dict:
    """Constructor arguments of the reference NaDiT as configs_3b/main.yaml:11-36 / configs_7b/main.yaml:11-33 resolve them
    (the YAMLs' ${eval:...} entries written out), optionally narrowed."""
    full = {"3b": dict(dim=2560, heads=20, layers=32, mm=10), "7b": dict(dim=3072, heads=24, layers=36, mm=0)}[model]
    n = num_layers or full["layers"]
    d = dim or full["dim"]
    h = heads or full["heads"]
    methods = ["720pwin_by_size_bysize", "720pswin_by_size_bysize"] * ((n + 1) // 2)
    common = dict(vid_in_channels=33, vid_out_channels=16, vid_dim=d, txt_in_dim=txt_in_dim, txt_dim=d, emb_dim=6 * d, heads=h,
                  head_dim=128, expand_ratio=4, norm="fusedrms", norm_eps=1e-5, ada="single", qk_bias=False, qk_norm="fusedrms",
                  patch_size=[1, 2, 2], num_layers=n, block_type=["mmdit_sr"] * n, window=[(4, 3, 3)] * n,
                  window_method=methods[:n])
    if model == "3b":
        return dict(common, mm_layers=full["mm"] if mm_layers is None else mm_layers, mlp_type="swiglu", patch_type="v1",
                    rope_type="mmrope3d", rope_dim=128, msa_type=None, mca_type=None, vid_out_norm="fusedrms")
    return dict(common, qk_rope=True, shared_mlp=False, shared_qkv=False, mlp_type="normal")


def build_block(model: str, dim: int, heads: int, shared: bool = False, is_last: bool = False, shifted: bool = False,
                window: Tuple[int, int, int] = (4, 3, 3), attention_mode: str = "sdpa"):
    """The reference's own NaMMSRTransformerBlock, built the way dit_3b/nadit.py:135-165 / dit_7b/nadit.py:109-136 build it."""
    activate()
    method = "720pswin_by_size_bysize" if shifted else "720pwin_by_size_bysize"
    if model == "3b":
        from src.models.dit_3b.modulation import get_ada_layer
        from src.models.dit_3b.nablocks.mmsr_block import NaMMSRTransformerBlock
        from src.models.dit_3b.normalization import get_norm_layer
        blk = NaMMSRTransformerBlock(
            vid_dim=dim, txt_dim=dim, emb_dim=6 * dim, heads=heads, head_dim=128, expand_ratio=4,
            norm=get_norm_layer("fusedrms"), norm_eps=1e-5, ada=get_ada_layer("single"), qk_bias=False,
            qk_norm=get_norm_layer("fusedrms"), shared_weights=shared, mlp_type="swiglu", window=window,
            window_method=method, msa_type=None, mca_type=None, rope_type="mmrope3d", rope_dim=128,
            is_last_layer=is_last, attention_mode=attention_mode)
    else:
        from src.models.dit_7b.modulation import get_ada_layer
        from src.models.dit_7b.nablocks.mmsr_block import NaMMSRTransformerBlock
        from src.models.dit_7b.normalization import get_norm_layer
        blk = NaMMSRTransformerBlock(
            vid_dim=dim, txt_dim=dim, emb_dim=6 * dim, heads=heads, head_dim=128, expand_ratio=4,
            norm=get_norm_layer("fusedrms"), norm_eps=1e-5, ada=get_ada_layer("single"), qk_bias=False, qk_rope=True,
            qk_norm=get_norm_layer("fusedrms"), shared_qkv=False, shared_mlp=False, mlp_type="normal", window=window,
            window_method=method, temporal_window_size=None, temporal_shifted=None, attention_mode=attention_mode)
    return blk.eval()


def new_cache(disable: bool = False):
    activate()
    from src.common.cache import Cache
    return Cache(disable=disable)
