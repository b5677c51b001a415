"""ctypes binding of libseedvr2_b200.so (the C ABI declared in include/seedvr2_b200.h).

There is NO fallback: if the shared library is missing or a call returns a non-zero code, a RuntimeError is raised
(north star: "no CPU fallback"; reference error convention = Python exceptions, src/optimization/compatibility.py:652-660).
"""
import ctypes
import os
import subprocess
from ctypes import POINTER, c_char_p, c_float, c_int, c_int32, c_int64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SVR2_LIB") or os.path.join(_HERE, "libseedvr2_b200.so")   # SVR2_LIB: kernel experiments
CSRC = os.path.join(_HERE, "csrc")

# name -> (restype, argtypes); must list EVERY symbol of include/seedvr2_b200.h (tests/test_cabi.py checks this)
SIGNATURES = {
    "svr2_last_error": (c_char_p, []),
    "svr2_abi_version": (c_int, []),
    "svr2_window_plan": (c_int, [c_int] * 8 + [POINTER(c_int32), c_int, POINTER(c_int64)]),
    "svr2_window_index": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "svr2_modnorm_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_float, c_void_p]),
    "svr2_gate_residual_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p]),
    "svr2_modnorm_affine_fwd": (c_int, [c_void_p] * 6 + [c_int64, c_int, c_int, c_float, c_void_p]),
    "svr2_gate_residual_modnorm_fwd": (c_int, [c_void_p] * 8 + [c_int64, c_int, c_int, c_float, c_void_p]),
    "svr2_swiglu_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "svr2_swiglu_rows_fwd": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "svr2_qkv_pack_fwd": (c_int, [c_void_p] * 7 + [c_int, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
                                  c_float, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "svr2_pack_rowmap": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_int64, c_void_p]),
    "svr2_window_attn_worklist": (c_int, [POINTER(c_int32), c_int, c_int, POINTER(c_int32), c_int, POINTER(c_int64)]),
    "svr2_window_attn_worklist_tiles": (c_int, [POINTER(c_int32), c_int, c_int, c_int, POINTER(c_int32), c_int, POINTER(c_int64)]),
    "svr2_window_attn_fwd": (c_int, [c_void_p] * 5 + [c_int, c_void_p] + [c_int] * 7 + [c_int64, c_void_p, c_void_p, c_void_p, c_void_p]),
    "svr2_window_attn_fwd_impl": (c_int, [c_void_p] * 5 + [c_int, c_void_p] + [c_int] * 7 + [c_int64, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "svr2_linear_fwd": (c_int, [c_void_p] * 4 + [c_int64, c_int, c_int, c_int, c_void_p]),
    "svr2_linear_swiglu_fwd": (c_int, [c_void_p] * 4 + [c_int64, c_int, c_int, c_void_p]),
    "svr2_linear_gate_residual_fwd": (c_int, [c_void_p] * 5 + [c_int, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p]),
    "svr2_linear_qkv_pack_fwd": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p, c_int, c_int] + [c_void_p] * 4 +
                                 [c_float] + [c_void_p] * 5),
    "svr2_rope_plan": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int64, c_int, c_void_p, c_int, c_void_p, c_int,
                               c_void_p, c_void_p, c_void_p]),
    "svr2_txt_replicate_fwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "svr2_window_attn_fwd_noalloc": (c_int, [c_void_p] * 5 + [c_int] * 7 + [c_int64, c_void_p, c_void_p, c_void_p]),
    "svr2_debug_attn_trace": (c_int, [c_void_p, c_int]),
    "svr2_txt_pool_fwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "svr2_halo_blend": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int64, c_void_p]),
    "svr2_segment_stitch_peer": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_void_p]),
    "svr2_wavelet_level_fwd": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_int64, c_void_p, c_void_p] + [c_int] * 7 + [c_void_p]),
    "svr2_colorfix_finish_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p] + [c_int] * 6 + [c_void_p]),
    "svr2_rgb_to_lab_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "svr2_lab_to_rgb_finish_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "svr2_conv3d_pack_weight": (c_int, [c_void_p, c_void_p] + [c_int] * 5 + [c_void_p]),
    "svr2_causal_conv3d_fwd": (c_int, [c_void_p] * 7 + [c_int] * 8 + [c_void_p]),
    "svr2_conv3d_gn_slots": (c_int, []),
    "svr2_causal_conv3d_down_fwd": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p] + [c_int] * 8 + [c_void_p]),
    "svr2_groupnorm_workspace_bytes": (c_int64, [c_int, c_int]),
    "svr2_groupnorm_silu_fwd": (c_int, [c_void_p] * 5 + [c_int, c_int, c_int64, c_int, c_int, c_float, c_int, c_int, c_void_p]),
    "svr2_transpose_2b": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_int, c_int64, c_int64, c_void_p]),
    "svr2_pixel_shuffle3d_fwd": (c_int, [c_void_p, c_void_p] + [c_int] * 7 + [c_void_p]),
    "svr2_row_softmax_fwd": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_void_p]),
    "svr2_sr_input_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_float, c_float, c_int, c_void_p, c_int64, c_int, c_void_p]),
    "svr2_v_lerp_endpoint_fwd": (c_int, [c_void_p, c_void_p, c_float, c_float, c_void_p, c_int64, c_void_p]),
}

ABI_VERSION = 14      # svr2_abi_version() of the library these SIGNATURES describe
_lib = None


def build(verbose: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a into libseedvr2_b200.so (nvcc cross-compiles without a GPU)."""
    r = subprocess.run(["make", "-C", CSRC, "-j8"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"building libseedvr2_b200.so failed:\n{r.stdout}\n{r.stderr}")
    if verbose:
        print(r.stdout)
    return LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: the seedvr2_b200 hot path is CUDA-only (sm_100a) and has no fallback. "
                f"Build it with `python -c 'import __graft_entry__ as g; g.build()'` or `make -C {CSRC}`.")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)      # AttributeError if the library does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        got = handle.svr2_abi_version()
        if got != ABI_VERSION:      # a stale in-tree .so would be called with the wrong argument lists
            raise RuntimeError(f"{LIB_PATH} has ABI version {got}, this package binds version {ABI_VERSION}: rebuild it "
                               f"(`make -C {CSRC}`)")
        _lib = handle
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().svr2_last_error().decode("utf-8", "replace")
        if rc == -3:
            raise NotImplementedError(f"{what}: {msg}")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")
