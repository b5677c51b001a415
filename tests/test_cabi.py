"""CPU: the C-ABI shared library loads and exports every symbol include/seedvr2_b200.h declares; host-only entry points
(window geometry, work list) are checked against the reference golden vectors.  No device work here."""
import ctypes
import json
import os
import re

import numpy as np
import pytest

import cases

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GEO = json.load(open(os.path.join(ROOT, "tests", "golden", "geometry.json")))


def test_exports_every_declared_symbol(svr2):
    hdr = open(os.path.join(ROOT, "include", "seedvr2_b200.h")).read()
    declared = set(re.findall(r"\b(svr2_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "header parse failed"
    assert declared == set(svr2._lib.SIGNATURES), declared ^ set(svr2._lib.SIGNATURES)
    handle = ctypes.CDLL(svr2._lib.LIB_PATH)
    for name in declared:
        assert hasattr(handle, name), f"{name} is declared but not exported"
    assert svr2._lib.lib().svr2_abi_version() == svr2._lib.ABI_VERSION


def test_shipped_library_holds_only_dispatched_attention_kernels(svr2):
    """The product library carries the two dispatched attention kernels and the mma.sync test comparator; the earlier
    generations (csrc/experiments/) are compiled only with `make EXPERIMENTS=1`, and no entry point reads an environment
    variable to pick a kernel."""
    blob = open(svr2._lib.LIB_PATH, "rb").read()
    for name in (b"window_attn_tc2_kernel", b"window_attn_tc6_kernel", b"window_attn_mma_kernel"):
        assert name in blob, name
    for name in (b"window_attn_tc3_kernel", b"window_attn_tc4_kernel", b"window_attn_tc5_kernel", b"window_attn_tc7_kernel",
                 b"window_attn_tc_kernel", b"SVR2_ATTN_IMPL", b"SVR2_ATTN_TURNS", b"SVR2_MODNORM_PREFETCH"):
        assert name not in blob, name


def test_errors_are_reported_not_swallowed(svr2):
    L = svr2._lib.lib()
    counts = (ctypes.c_int64 * 8)()
    assert L.svr2_window_plan(0, 4, 4, 4, 3, 3, 0, 58, None, 0, counts) == -1
    assert b"grid" in L.svr2_last_error()
    with pytest.raises(RuntimeError):
        svr2._lib.check(-1, "svr2_window_plan")


def test_new_entry_points_validate_their_arguments_before_launching(svr2):
    """Argument errors of the round's later entry points come back as codes + text, without touching a device."""
    L = svr2._lib.lib()
    assert L.svr2_sr_input_fwd(None, None, None, 1.0, 0.0, 0, None, 4, 16, None) == -1
    assert b"svr2_sr_input_fwd" in L.svr2_last_error()
    assert L.svr2_v_lerp_endpoint_fwd(None, None, 0.0, 1.0, None, 8, None) == -1
    assert b"svr2_v_lerp_endpoint_fwd" in L.svr2_last_error()
    assert L.svr2_segment_stitch_peer(None, None, None, 0, 0, 1, 16, None) == -1
    assert b"svr2_segment_stitch_peer" in L.svr2_last_error()
    n = ctypes.c_int64(0)
    assert L.svr2_window_attn_worklist_tiles(None, 1, 58, 4, None, 0, ctypes.byref(n)) == -1
    assert b"worklist" in L.svr2_last_error()
    # round 2: GEMMs, fused qkv projection and its plan tables, post-processing
    one = ctypes.c_void_p(16)      # a non-null pointer that is never dereferenced: every call below fails validation first
    assert L.svr2_linear_fwd(one, one, None, one, 8, 60, 64, 0, None) == -1 and b"multiple of 64" in L.svr2_last_error()
    assert L.svr2_linear_fwd(one, one, None, one, 8, 64, 64, 7, None) == -1 and b"activation" in L.svr2_last_error()
    assert L.svr2_linear_swiglu_fwd(one, one, None, one, 8, 128, 64, None) == -1
    assert L.svr2_linear_gate_residual_fwd(one, one, None, one, one, 2, one, one, 8, 64, 64, None) == -1 and b"layer" in L.svr2_last_error()
    assert L.svr2_linear_qkv_pack_fwd(one, one, 8, 64, 3, one, None, 0, 0, one, one, one, one, 1e-5, one, one, one, one, None) == -1
    assert b"even" in L.svr2_last_error()
    assert L.svr2_linear_qkv_pack_fwd(one, one, 8, 64, 2, one, one, 4, 7, one, one, one, one, 1e-5, one, one, one, one, None) == -1
    assert b"n_freqs" in L.svr2_last_error()
    assert L.svr2_linear_qkv_pack_fwd(one, one, 8, 64, 2, one, one, 4000, 21, one, one, one, one, 1e-5, one, one, one, one, None) == -1
    assert b"shared memory" in L.svr2_last_error()
    assert L.svr2_rope_plan(None, None, 1, 1, 8, 8, 58, 100, 1, None, 21, None, 70, None, None, None) == -1
    assert L.svr2_txt_replicate_fwd(None, None, 0, 58, 2, None, None, None, None) == -1
    assert L.svr2_wavelet_level_fwd(one, 1, 1, 1, 1, one, None, 1, 1, 3, 8, 8, 0, 0, None) == -1 and b"radius" in L.svr2_last_error()
    assert L.svr2_wavelet_level_fwd(one, 1, 1, 1, 1, one, None, 1, 1, 3, 8, 8, 1, 5, None) == -1 and b"dtype" in L.svr2_last_error()
    assert L.svr2_colorfix_finish_fwd(one, one, None, None, 0, 1, 3, 8, 8, 0, None) == -1
    assert L.svr2_rgb_to_lab_fwd(one, None, one, 1, 8, 8, 9, None) == -1 and b"dtype" in L.svr2_last_error()
    assert L.svr2_lab_to_rgb_finish_fwd(one, None, None, 0, 1, 8, 8, 0, None) == -1


@pytest.mark.parametrize("key", sorted(GEO))
def test_host_plan_matches_reference(svr2, key):
    rec = GEO[key]
    grid, method = key.split(":")
    T, H, W = (int(v) for v in grid.split("x"))
    p = svr2.WindowPlan((T, H, W), (4, 3, 3), method == "swin", cases.TXT_LEN)
    assert p.bounds.tolist() == rec["bounds"]
    assert p.window_shape.tolist() == rec["window_shape"]
    assert p.cu_seqlens.tolist() == rec["cu_seqlens_txt58"]
    assert p.n_windows == rec["window_count"] and p.n_tokens == T * H * W
    assert p.packed_rows == rec["cu_seqlens_txt58"][-1]
    # work list: every 256-row query pair of every window exactly once, longest sequences first
    seqs = p.window_lens + cases.TXT_LEN
    want = sorted((int(w), q) for w in range(p.n_windows) for q in range((int(seqs[w]) + 255) // 256))
    assert sorted(map(tuple, p.work.tolist())) == want
    lens = [int(seqs[w]) for w, _ in p.work.tolist()]
    assert lens == sorted(lens, reverse=True)


def test_product_has_no_oracle_or_fallback():
    """The product package must not import oracle/ or fall back to torch attention."""
    pkg = os.path.join(ROOT, "comfyui-seedvr2_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "import oracle" not in src and "from oracle" not in src, fn
            assert "scaled_dot_product_attention" not in src and "flash_attn" not in src, fn


@pytest.mark.parametrize("grid,shifted", [((1, 128, 128), False), ((1, 128, 128), True), ((9, 90, 160), True), ((31, 68, 120), False),
                                          ((1, 3, 5), True)])
def test_attention_work_lists_cover_every_query_tile_once(svr2, grid, shifted):
    """Host-side work lists of the attention kernels (pure integer, no GPU): the 2-tile list (generations 4 / 8) and the
    4-tile list (generation 9) must each cover every 128-row query tile of every window exactly once, longest sequences
    first — the per-window Python loop of pytorch_varlen_attention (attention.py:24-61) restated as a static schedule."""
    import numpy as np
    plan = svr2.get_plan(grid, (4, 3, 3), shifted, 58)
    seq = plan.table[:, 7].astype(np.int64) + plan.txt_len
    for work, tiles in ((plan.work, 2), (plan.work4, 4)):
        seen = {}
        for wi, grp in work.tolist():
            for t in range(grp * tiles, (grp + 1) * tiles):
                if t * 128 < seq[wi]:
                    assert (wi, t) not in seen
                    seen[(wi, t)] = True
            assert grp * tiles * 128 < seq[wi], "an entry without a single valid tile"
        assert len(seen) == int(sum((s + 127) // 128 for s in seq))
        lens = [int(seq[wi]) for wi, _ in work.tolist()]
        assert lens == sorted(lens, reverse=True)
