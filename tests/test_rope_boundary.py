"""The one third-party arithmetic boundary of the path: ``rotary_embedding_torch`` (lucidrains, pinned only as >= 0.5.3,
reference requirements.txt:11).  The image does not have it, so the goldens were generated through the restatement in
oracle/shim; wherever the genuine package IS importable (a user's machine, a future image) this test pins the restatement
to it at the reference's call sites (dit_3b/rope.py:28-32,46,118-126,162-167; dit_7b/rope.py:28-32,46,88-89,109).
Skips when the package is absent — the boundary then stays "parity unpinned", as DESIGN §7 states."""
import importlib
import os
import sys

import pytest
import torch

from oracle import refstage

SHIM = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "shim")


def _load_real():
    saved_path, saved_mod = list(sys.path), sys.modules.pop("rotary_embedding_torch", None)
    try:
        sys.path[:] = [p for p in sys.path if os.path.abspath(p) != SHIM]
        return importlib.import_module("rotary_embedding_torch")
    finally:
        sys.path[:] = saved_path
        real = sys.modules.pop("rotary_embedding_torch", None)
        if saved_mod is not None:
            sys.modules["rotary_embedding_torch"] = saved_mod
        _load_real.mod = real


@pytest.mark.skipif(not refstage.real_rotary_embedding_torch(), reason="the genuine rotary_embedding_torch is not installed: boundary unpinned")
def test_shim_matches_genuine_rotary_embedding_torch():
    real = _load_real()
    sys.path.insert(0, SHIM)
    try:
        sys.modules.pop("rotary_embedding_torch", None)
        shim = importlib.import_module("rotary_embedding_torch")
    finally:
        sys.path.remove(SHIM)
        sys.modules.pop("rotary_embedding_torch", None)
    for kw, dims in ((dict(dim=128 // 3, freqs_for="lang", theta=10000), (61, 5, 7)), (dict(dim=(128 // 2) // 3, freqs_for="pixel", max_freq=256), (3, 15, 27))):
        a, b = real.RotaryEmbedding(**kw), shim.RotaryEmbedding(**kw)
        assert torch.equal(a.freqs.detach(), b.freqs.detach())
        fa, fb = a.get_axial_freqs(*dims), b.get_axial_freqs(*dims)
        assert torch.equal(fa, fb)
        t = torch.randn(dims[0] * dims[1] * dims[2], 2, 128, generator=torch.Generator().manual_seed(1))
        f = fa.reshape(-1, 1, fa.shape[-1])
        assert torch.equal(real.apply_rotary_emb(f, t), shim.apply_rotary_emb(f, t))
