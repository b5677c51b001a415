"""seedvr2_b200 — B200-native (sm_100a) hot path of SeedVR2's one-step NaDiT block, as a drop-in behind the
reference's ``src/models`` DiT block interface (disk02/ComfyUI-SeedVR2).

    import seedvr2_b200            # repo-root alias of this package (the directory name has a hyphen)
    seedvr2_b200.install()         # before the reference builds its model: nadit_blocks["mmdit_sr"] -> B200 blocks

The package holds only what the path needs: csrc/ (CUDA kernels + the C ABI of include/seedvr2_b200.h), the ctypes
binding, the window plan, the block modules mirroring the reference interface, a minimal NaDiT driver for benches and
the temporal-segment partitioner.  It never imports ``oracle/`` and raises if libseedvr2_b200.so is missing.
"""
from . import _lib, ops, plan, postprocess, sampler, vae_attn, vae_block, vae_conv
from .block import NaMMSRTransformerBlockB200, NaMMSRTransformerBlockB200_7B, NaSwinAttentionB200
from .plan import WindowPlan, get_plan
from .stack import NaDiTStack

__all__ = ["install", "uninstall", "convert_parameters_", "build_extension", "NaMMSRTransformerBlockB200", "NaMMSRTransformerBlockB200_7B",
           "NaSwinAttentionB200", "NaDiTStack", "WindowPlan", "get_plan", "ops", "plan"]

_saved = {}


def build_extension(verbose: bool = False) -> str:
    return _lib.build(verbose)


def install(models_pkg: str = "src.models") -> None:
    """Swap the reference's block registry entries (dit_3b/nablocks/__init__.py:18-26, dit_7b/nablocks/__init__.py:17-25).

    ``get_nablock`` reads the module-level dict ``nadit_blocks`` at call time, so this must run before
    ``create_object(config.dit.model)`` (src/core/model_loader.py:437-438) — e.g. at ComfyUI custom-node import or ahead
    of ``inference_cli.main()``.  Loading libseedvr2_b200.so is forced here so a missing build fails at install time."""
    import importlib
    _lib.lib()
    for sub, cls in (("dit_3b", NaMMSRTransformerBlockB200), ("dit_7b", NaMMSRTransformerBlockB200_7B)):
        mod = importlib.import_module(f"{models_pkg}.{sub}.nablocks")
        _saved.setdefault((models_pkg, sub), mod.nadit_blocks["mmdit_sr"])
        mod.nadit_blocks["mmdit_sr"] = cls


def convert_parameters_(model) -> int:
    """Optional, after the checkpoint is loaded: rewrite the parameters of every B200 block in ``model`` to bf16 IN PLACE
    (``param.data``), keeping ``attn.rope.rope.freqs`` in its loaded dtype.  Checkpoints loaded with ``assign=True`` keep the file's
    dtype (fp16 for the shipped files, model_loader.py:809); the blocks then cast each such parameter once and keep the bf16 copy
    next to the original (block._bf16), which doubles the weight memory (+7 GB for 3B, +16 GB for 7B) and survives a BlockSwap
    offload of the original.  After this call there is one copy.  Returns the number of tensors converted."""
    import torch
    from .block import _BlockBase, _cast_cache
    n = 0
    for mod in model.modules():
        if isinstance(mod, _BlockBase):
            for name, p in mod.named_parameters():
                if p.dtype != torch.bfloat16 and p.dtype.is_floating_point:
                    p.data = p.data.to(torch.bfloat16)
                    n += 1
    _cast_cache.clear()
    return n


def uninstall(models_pkg: str = "src.models") -> None:
    import importlib
    for sub in ("dit_3b", "dit_7b"):
        key = (models_pkg, sub)
        if key in _saved:
            importlib.import_module(f"{models_pkg}.{sub}.nablocks").nadit_blocks["mmdit_sr"] = _saved.pop(key)
