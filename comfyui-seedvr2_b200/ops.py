"""Tensor-level wrappers over the C ABI: validate, fetch raw pointers + the current CUDA stream, call, raise on error.

PyTorch is plumbing here (memory + streams).  Every function runs on the tensor's CUDA device and is stream-ordered on
``torch.cuda.current_stream()``.  Non-CUDA or non-bf16 inputs raise — there is no CPU / eager fallback.
Calls sit behind ``torch._dynamo.disable`` like the reference's own flash-attn call (src/models/dit_3b/attention.py:64),
so an optional ``torch.compile(model)`` (model_configuration.py:1217-1219) treats them as opaque.
"""
from typing import Optional, Tuple

import torch

from . import _lib
from .plan import WindowPlan

ROPE_NONE, ROPE_3B_MM, ROPE_7B_PIX = 0, 1, 2
ATTN_AUTO, ATTN_GEN4, ATTN_GEN8, ATTN_GEN10, ATTN_MMA = 0, 4, 8, 10, 100      # include/seedvr2_b200.h SVR2_ATTN_*


class KernelTimer:
    """Optional per-kernel CUDA-event timing on the launching stream (bench.py's live roofline measurement).
    Install with ``ops.set_timer(KernelTimer())``; every C-ABI launch is then bracketed by two events."""

    def __init__(self, only=None):
        self.only = None if only is None else set(only)
        self.records = {}

    def wants(self, name: str) -> bool:
        return self.only is None or name in self.only

    def start(self, device):
        e = torch.cuda.Event(enable_timing=True)
        e.record(torch.cuda.current_stream(device))
        return e

    def stop(self, name: str, e0, device):
        e1 = torch.cuda.Event(enable_timing=True)
        e1.record(torch.cuda.current_stream(device))
        self.records.setdefault(name, []).append((e0, e1))

    def summary(self):
        """{name: (launches, total_ms)} — call after torch.cuda.synchronize()."""
        return {k: (len(v), sum(a.elapsed_time(b) for a, b in v)) for k, v in self.records.items()}


    def durations(self):
        """{name: [ms of every launch]} — call after torch.cuda.synchronize()."""
        return {k: [a.elapsed_time(b) for a, b in v] for k, v in self.records.items()}


_timer: Optional[KernelTimer] = None
_launches = 0      # kernels of libseedvr2_b200.so launched by this process (bench.py's `gpu_launches`)


def set_timer(t: Optional[KernelTimer]) -> None:
    global _timer
    _timer = t


def launch_count() -> int:
    return _launches


def count_launches(n: int) -> None:
    """A CUDA-graph replay launches the library kernels captured in it without passing through ``_call``."""
    global _launches
    _launches += n


def _call(name: str, n_kernels: int, device, fn, *args) -> None:
    global _launches
    t = _timer
    timed = t is not None and t.wants(name)
    with torch.cuda.device(device):
        e0 = t.start(device) if timed else None
        _lib.check(fn(*args), name)
        if timed:
            t.stop(name, e0, device)
    _launches += n_kernels


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _need(t: torch.Tensor, name: str, dtype=torch.bfloat16) -> torch.Tensor:
    if not t.is_cuda:
        raise RuntimeError(f"seedvr2_b200: `{name}` must live on a CUDA device (got {t.device}); there is no CPU path")
    if t.dtype != dtype:
        raise NotImplementedError(f"seedvr2_b200: `{name}` must be {dtype} (got {t.dtype})")
    if not t.is_contiguous():
        raise RuntimeError(f"seedvr2_b200: `{name}` must be contiguous")
    return t


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


@torch._dynamo.disable
def modnorm(x: torch.Tensor, emb: Optional[torch.Tensor], shift_b: Optional[torch.Tensor],
            scale_b: Optional[torch.Tensor], layer: int, eps: float, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """y = rmsnorm(x) * (scaleA + scaleB) + (shiftA + shiftB);  emb/shift_b/scale_b all None -> norm only."""
    _need(x, "x")
    L, D = x.shape
    y = torch.empty_like(x) if out is None else _need(out, "out")
    for n, t in (("emb", emb), ("shift_b", shift_b), ("scale_b", scale_b)):
        if t is not None:
            _need(t, n)
    if L:
        _call("svr2_modnorm_fwd", 1, x.device, _lib.lib().svr2_modnorm_fwd, x.data_ptr(), _ptr(emb), _ptr(shift_b),
              _ptr(scale_b), y.data_ptr(), L, D, layer, float(eps), _stream(x))
    return y


@torch._dynamo.disable
def modnorm_affine(x: torch.Tensor, weight: torch.Tensor, emb: Optional[torch.Tensor], shift_b: Optional[torch.Tensor],
                   scale_b: Optional[torch.Tensor], layer: int, eps: float) -> torch.Tensor:
    """y = rmsnorm(x) * weight * (scaleA + scaleB) + (shiftA + shiftB): the model's output norm + output modulation."""
    _need(x, "x"); _need(weight, "weight")
    L, D = x.shape
    y = torch.empty_like(x)
    for n, t in (("emb", emb), ("shift_b", shift_b), ("scale_b", scale_b)):
        if t is not None:
            _need(t, n)
    if L:
        _call("svr2_modnorm_affine_fwd", 1, x.device, _lib.lib().svr2_modnorm_affine_fwd, x.data_ptr(), weight.data_ptr(),
              _ptr(emb), _ptr(shift_b), _ptr(scale_b), y.data_ptr(), L, D, layer, float(eps), _stream(x))
    return y


@torch._dynamo.disable
def gate_residual(h: torch.Tensor, resid: torch.Tensor, emb: Optional[torch.Tensor], gate_b: Optional[torch.Tensor],
                  layer: int) -> torch.Tensor:
    """out = h * (gateA + gateB) + resid;  emb/gate_b None -> out = h + resid."""
    _need(h, "h"); _need(resid, "resid")
    if h.shape != resid.shape:
        raise RuntimeError("gate_residual: shape mismatch")
    L, D = h.shape
    out = torch.empty_like(h)
    for n, t in (("emb", emb), ("gate_b", gate_b)):
        if t is not None:
            _need(t, n)
    if L:
        _call("svr2_gate_residual_fwd", 1, h.device, _lib.lib().svr2_gate_residual_fwd, h.data_ptr(), resid.data_ptr(),
              _ptr(emb), _ptr(gate_b), out.data_ptr(), L, D, layer, _stream(h))
    return out


@torch._dynamo.disable
def gate_residual_modnorm(h: torch.Tensor, resid: torch.Tensor, emb: torch.Tensor, gate_b: torch.Tensor,
                          shift_b2: torch.Tensor, scale_b2: torch.Tensor, eps: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """(out, y): out = h*gate(attn) + resid ; y = rmsnorm(out)*scale(mlp) + shift(mlp)."""
    for n, t in (("h", h), ("resid", resid), ("emb", emb), ("gate_b", gate_b), ("shift_b2", shift_b2), ("scale_b2", scale_b2)):
        _need(t, n)
    L, D = h.shape
    out, y = torch.empty_like(h), torch.empty_like(h)
    if L:
        _call("svr2_gate_residual_modnorm_fwd", 1, h.device, _lib.lib().svr2_gate_residual_modnorm_fwd,
              h.data_ptr(), resid.data_ptr(), emb.data_ptr(), gate_b.data_ptr(), shift_b2.data_ptr(), scale_b2.data_ptr(),
              out.data_ptr(), y.data_ptr(), L, D, 0, float(eps), _stream(h))
    return out, y


@torch._dynamo.disable
def swiglu(gate: torch.Tensor, up: torch.Tensor) -> torch.Tensor:
    """silu(gate) * up (reference mlp.py:60-62), one pass."""
    _need(gate, "gate"); _need(up, "up")
    if gate.shape != up.shape:
        raise RuntimeError("swiglu: shape mismatch")
    out = torch.empty_like(gate)
    if gate.numel():
        _call("svr2_swiglu_fwd", 1, gate.device, _lib.lib().svr2_swiglu_fwd, gate.data_ptr(), up.data_ptr(), out.data_ptr(),
              gate.numel(), _stream(gate))
    return out


@torch._dynamo.disable
def swiglu_rows(gate_up: torch.Tensor) -> torch.Tensor:
    """silu(gate) * up on a [rows, 2*hidden] buffer holding gate | up side by side (one stacked-weight GEMM's output)."""
    _need(gate_up, "gate_up")
    rows, h2 = gate_up.shape
    out = torch.empty(rows, h2 // 2, dtype=torch.bfloat16, device=gate_up.device)
    if rows:
        _call("svr2_swiglu_rows_fwd", 1, gate_up.device, _lib.lib().svr2_swiglu_rows_fwd, gate_up.data_ptr(), out.data_ptr(),
              rows, h2 // 2, _stream(gate_up))
    return out


@torch._dynamo.disable
def qkv_pack(plan: WindowPlan, qkv_vid: torch.Tensor, qkv_txt: torch.Tensor, wq_vid, wk_vid, wq_txt, wk_txt,
             freqs: Optional[torch.Tensor], rope_kind: int, heads: int, eps: float):
    """-> (q, k, v) packed [P, heads, 128] bf16: q/k RMSNorm + RoPE + window partition + text interleave."""
    _need(qkv_vid, "qkv_vid"); _need(qkv_txt, "qkv_txt")
    for n, t in (("wq_vid", wq_vid), ("wk_vid", wk_vid), ("wq_txt", wq_txt), ("wk_txt", wk_txt)):
        _need(t, n)
    T, H, W = plan.grid
    if qkv_vid.shape != (plan.n_tokens, 3 * heads * 128) or qkv_txt.shape != (plan.txt_len, 3 * heads * 128):
        raise RuntimeError(f"qkv_pack: bad shapes {tuple(qkv_vid.shape)} / {tuple(qkv_txt.shape)} for plan {plan.grid}")
    n_freqs = 0
    if rope_kind != ROPE_NONE:
        _need(freqs, "freqs", torch.float32)
        n_freqs = freqs.numel()
    dev = qkv_vid.device
    P = plan.packed_rows
    q = torch.empty(P, heads, 128, dtype=torch.bfloat16, device=dev)
    k = torch.empty_like(q)
    v = torch.empty_like(q)
    d = plan.on(dev)
    _call("svr2_qkv_pack_fwd", 1, dev, _lib.lib().svr2_qkv_pack_fwd,
          qkv_vid.data_ptr(), qkv_txt.data_ptr(), wq_vid.data_ptr(), wk_vid.data_ptr(), wq_txt.data_ptr(),
          wk_txt.data_ptr(), _ptr(freqs), n_freqs, rope_kind, d["table"].data_ptr(), plan.row_map(dev).data_ptr(),
          plan.n_windows, T, H, W, plan.txt_len, heads, float(eps), q.data_ptr(), k.data_ptr(), v.data_ptr(), P,
          _stream(qkv_vid))
    return q, k, v


QKV_FUSED_MAX_ROPE_ENTRIES = 2048      # csrc/gemm.cu kMaxRopeEntries: (cos, sin) pairs of the rope table held in shared memory


def qkv_fused_supported(plan: WindowPlan, rope_kind: int, n_freqs: int, heads: int) -> bool:
    """Whether svr2_linear_qkv_pack_fwd covers this configuration (else: svr2_linear_fwd + svr2_qkv_pack_fwd, same results)."""
    if heads % 2 or (rope_kind != ROPE_NONE and n_freqs not in (21, 10)):
        return False
    if rope_kind == ROPE_3B_MM:
        t = plan.table
        rows = max(plan.txt_len + int(t[:, 3].max()), int(t[:, 4].max()), int(t[:, 5].max()), plan.txt_len, 1)
    elif rope_kind == ROPE_7B_PIX:
        ns = set(int(v) for v in plan.table[:, 3:6].reshape(-1))
        if max(ns) > 64:
            return False
        rows = sum(ns)
    else:
        rows = 0
    return rows * max(n_freqs, 1) <= QKV_FUSED_MAX_ROPE_ENTRIES


@torch._dynamo.disable
def linear_qkv_pack(plan: WindowPlan, x: torch.Tensor, weight: torch.Tensor, row_offset: int, wq_vid, wk_vid, wq_txt, wk_txt,
                    freqs: Optional[torch.Tensor], freqs_key: Tuple[float, ...], rope_kind: int, heads: int, eps: float,
                    out: Optional[Tuple[torch.Tensor, ...]] = None):
    """x [M, K] = rows [row_offset, row_offset + M) of the order (video tokens, then text tokens)  ->  (q, k, v, txt_qkv):
    q, k, v packed [P, heads, 128] with the video rows of this call written, txt_qkv [Lt, 3, heads, 128] with its text rows
    written.  Pass ``out`` to add a second call's rows (separate-weight layers: one call per branch) to the same buffers.
    ``replicate_text`` must follow once all text rows are in."""
    _need(x, "x")
    M, K = x.shape
    _need_weight(weight, "weight", K)
    if weight.shape[0] != 3 * heads * 128:
        raise RuntimeError(f"linear_qkv_pack: weight must be [{3 * heads * 128}, {K}]")
    for n, t in (("wq_vid", wq_vid), ("wk_vid", wk_vid), ("wq_txt", wq_txt), ("wk_txt", wk_txt)):
        _need(t, n)
    if row_offset < 0 or row_offset + M > plan.n_tokens + plan.txt_len:
        raise RuntimeError("linear_qkv_pack: rows out of range for the plan")
    dev = x.device
    if rope_kind != ROPE_NONE:
        _need(freqs, "freqs", torch.float32)
    tok_map, cs, rows = plan.rope_plan(dev, rope_kind, freqs_key if rope_kind != ROPE_NONE else (), freqs)
    if out is None:
        P = plan.packed_rows
        q = torch.empty(P, heads, 128, dtype=torch.bfloat16, device=dev)
        out = (q, torch.empty_like(q), torch.empty_like(q), torch.empty(plan.txt_len, 3, heads, 128, dtype=torch.bfloat16, device=dev))
    q, k, v, tq = out
    if M:
        _call("svr2_linear_qkv_pack_fwd", 1, dev, _lib.lib().svr2_linear_qkv_pack_fwd, x.data_ptr(), weight.data_ptr(), M, K, heads,
              tok_map.data_ptr() + 16 * row_offset, None if cs is None else cs.data_ptr(), rows,
              len(freqs_key) if rope_kind != ROPE_NONE else 0, wq_vid.data_ptr(), wk_vid.data_ptr(), wq_txt.data_ptr(),
              wk_txt.data_ptr(), float(eps), q.data_ptr(), k.data_ptr(), v.data_ptr(), tq.data_ptr(), _stream(x))
    return out


@torch._dynamo.disable
def replicate_text(plan: WindowPlan, txt_qkv: torch.Tensor, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, heads: int) -> None:
    """Copies the text rows of q, k, v into every window's sequence (mmattn.py:223-236)."""
    if plan.txt_len:
        d = plan.on(q.device)
        _call("svr2_txt_replicate_fwd", 1, q.device, _lib.lib().svr2_txt_replicate_fwd, txt_qkv.data_ptr(), d["table"].data_ptr(),
              plan.n_windows, plan.txt_len, heads, q.data_ptr(), k.data_ptr(), v.data_ptr(), _stream(q))


@torch._dynamo.disable
def window_attn(plan: WindowPlan, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, heads: int, impl: int = ATTN_AUTO):
    """-> (out_vid [L, heads*128] in original token order, out_txt [Lt, heads*128] mean over windows).
    ``impl`` names a kernel explicitly (tests / A-B runs); the product path leaves it at ATTN_AUTO."""
    for n, t in (("q", q), ("k", k), ("v", v)):
        _need(t, n)
        if t.shape != (plan.packed_rows, heads, 128):
            raise RuntimeError(f"window_attn: `{n}` has shape {tuple(t.shape)}, expected {(plan.packed_rows, heads, 128)}")
    dev = q.device
    T, H, W = plan.grid
    # one buffer, video rows then text rows: shared-weight layers run proj_out on it as ONE GEMM (block._joined)
    out = torch.empty(plan.n_tokens + plan.txt_len, heads * 128, dtype=torch.bfloat16, device=dev)
    out_vid, out_txt = out[:plan.n_tokens], out[plan.n_tokens:]
    txt_part = torch.empty(plan.n_windows, plan.txt_len, heads * 128, dtype=torch.bfloat16, device=dev)
    d = plan.on(dev)
    # attention and text pooling are two kernels; they are launched separately so the attention kernel can be timed alone
    if impl == ATTN_AUTO:
        _call("svr2_window_attn_fwd", 1, dev, _lib.lib().svr2_window_attn_fwd,
              q.data_ptr(), k.data_ptr(), v.data_ptr(), d["table"].data_ptr(), d["work"].data_ptr(), plan.work.shape[0],
              d["work4"].data_ptr(), plan.work4.shape[0], plan.n_windows, T, H, W, plan.txt_len, heads, plan.packed_rows,
              out_vid.data_ptr(), txt_part.data_ptr(), None, _stream(q))
    else:
        _call("svr2_window_attn_fwd", 1, dev, _lib.lib().svr2_window_attn_fwd_impl,
              q.data_ptr(), k.data_ptr(), v.data_ptr(), d["table"].data_ptr(), d["work"].data_ptr(), plan.work.shape[0],
              d["work4"].data_ptr(), plan.work4.shape[0], plan.n_windows, T, H, W, plan.txt_len, heads, plan.packed_rows,
              out_vid.data_ptr(), txt_part.data_ptr(), None, int(impl), _stream(q))
    if plan.txt_len:
        _call("svr2_txt_pool_fwd", 1, dev, _lib.lib().svr2_txt_pool_fwd, txt_part.data_ptr(), out_txt.data_ptr(),
              plan.n_windows, plan.txt_len, heads, _stream(q))
    return out_vid, out_txt


@torch._dynamo.disable
def halo_blend(prev_tail: torch.Tensor, cur_head: torch.Tensor, overlap: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Crossfade of `overlap` frames [overlap, ...] fp32: prev*w + cur*(1-w), Hann for overlap >= 3 else linear."""
    _need(prev_tail, "prev_tail", torch.float32); _need(cur_head, "cur_head", torch.float32)
    if prev_tail.shape != cur_head.shape or prev_tail.shape[0] != overlap:
        raise RuntimeError("halo_blend: inputs must both be [overlap, ...]")
    out = torch.empty_like(prev_tail) if out is None else _need(out, "out", torch.float32)
    frame = prev_tail[0].numel()
    _call("svr2_halo_blend", 1, prev_tail.device, _lib.lib().svr2_halo_blend, prev_tail.data_ptr(), cur_head.data_ptr(),
          out.data_ptr(), int(overlap), frame, _stream(prev_tail))
    return out


@torch._dynamo.disable
def sr_input(noise: torch.Tensor, latent: torch.Tensor, aug_noise: Optional[torch.Tensor] = None,
             A: float = 1.0, B: float = 0.0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """[N, 2C+1] bf16 DiT input = [noise | A*latent + B*aug_noise (or latent when aug_noise is None) | 1]
    (infer.py:54-78,341; generation_phases.py:685-693).  noise / latent / aug_noise: [..., C] with N = prod(leading dims)."""
    _need(noise, "noise"); _need(latent, "latent")
    if aug_noise is not None:
        _need(aug_noise, "aug_noise")
        if aug_noise.shape != latent.shape:
            raise RuntimeError("sr_input: aug_noise must have the latent's shape")
    if noise.shape != latent.shape:
        raise RuntimeError(f"sr_input: noise {tuple(noise.shape)} and latent {tuple(latent.shape)} differ")
    C = noise.shape[-1]
    N = noise.numel() // C
    if out is None:
        out = torch.empty(N, 2 * C + 1, dtype=torch.bfloat16, device=noise.device)
    else:
        _need(out, "out")
        if out.shape != (N, 2 * C + 1):
            raise RuntimeError(f"sr_input: out must be [{N}, {2 * C + 1}]")
    _call("svr2_sr_input_fwd", 1, noise.device, _lib.lib().svr2_sr_input_fwd, noise.data_ptr(), latent.data_ptr(), _ptr(aug_noise),
          float(A), float(B), int(aug_noise is not None), out.data_ptr(), N, C, _stream(noise))
    return out


@torch._dynamo.disable
def v_lerp_endpoint(x_t: torch.Tensor, pred: torch.Tensor, A: float = 0.0, B: float = 1.0,
                    out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """bf16 (x_t - B*pred) / (A + B), every op rounded to bf16 (samplers/base.py:88-98, schedules/base.py:107-109)."""
    _need(x_t, "x_t"); _need(pred, "pred")
    if x_t.numel() != pred.numel():
        raise RuntimeError("v_lerp_endpoint: x_t and pred differ in size")
    out = torch.empty_like(x_t) if out is None else _need(out, "out")
    _call("svr2_v_lerp_endpoint_fwd", 1, x_t.device, _lib.lib().svr2_v_lerp_endpoint_fwd, x_t.data_ptr(), pred.data_ptr(),
          float(A), float(B), out.data_ptr(), x_t.numel(), _stream(x_t))
    return out


# ----------------------------------------------------------------------------------------------- linear layers (csrc/gemm.cu)
ACT_NONE, ACT_GELU_TANH = 0, 1


def _need_weight(w: torch.Tensor, name: str, K: int) -> torch.Tensor:
    _need(w, name)
    if w.dim() != 2 or w.shape[1] != K:
        raise RuntimeError(f"seedvr2_b200: `{name}` must be [N, {K}] (nn.Linear layout), got {tuple(w.shape)}")
    return w


@torch._dynamo.disable
def linear(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None, activation: int = ACT_NONE,
           out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """y = x @ weight.T (+ bias), optionally GELU-tanh in the epilogue: the tcgen05 GEMM of csrc/gemm.cu."""
    _need(x, "x")
    M, K = x.shape
    _need_weight(weight, "weight", K)
    N = weight.shape[0]
    if bias is not None:
        _need(bias, "bias")
    y = torch.empty(M, N, dtype=torch.bfloat16, device=x.device) if out is None else _need(out, "out")
    if M:
        _call("svr2_linear_fwd", 1, x.device, _lib.lib().svr2_linear_fwd, x.data_ptr(), weight.data_ptr(), _ptr(bias),
              y.data_ptr(), M, N, K, int(activation), _stream(x))
    return y


@torch._dynamo.disable
def linear_swiglu(x: torch.Tensor, w_gate: torch.Tensor, w_up: torch.Tensor) -> torch.Tensor:
    """silu(x @ w_gate.T) * (x @ w_up.T) in one GEMM (reference mlp.py:60-62)."""
    _need(x, "x")
    M, K = x.shape
    _need_weight(w_gate, "w_gate", K)
    _need_weight(w_up, "w_up", K)
    if w_gate.shape != w_up.shape:
        raise RuntimeError("linear_swiglu: w_gate and w_up differ in shape")
    N = w_gate.shape[0]
    y = torch.empty(M, N, dtype=torch.bfloat16, device=x.device)
    if M:
        _call("svr2_linear_swiglu_fwd", 1, x.device, _lib.lib().svr2_linear_swiglu_fwd, x.data_ptr(), w_gate.data_ptr(),
              w_up.data_ptr(), y.data_ptr(), M, N, K, _stream(x))
    return y


@torch._dynamo.disable
def linear_gate_residual(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], emb: torch.Tensor,
                         gate_b: torch.Tensor, layer: int, resid: torch.Tensor) -> torch.Tensor:
    """(x @ weight.T + bias) * (gateA[layer] + gateB) + resid in one GEMM (proj_out + AdaSingle "out" + residual)."""
    _need(x, "x"); _need(emb, "emb"); _need(gate_b, "gate_b"); _need(resid, "resid")
    M, K = x.shape
    _need_weight(weight, "weight", K)
    N = weight.shape[0]
    if bias is not None:
        _need(bias, "bias")
    if resid.shape != (M, N):
        raise RuntimeError(f"linear_gate_residual: resid must be [{M}, {N}]")
    out = torch.empty(M, N, dtype=torch.bfloat16, device=x.device)
    if M:
        _call("svr2_linear_gate_residual_fwd", 1, x.device, _lib.lib().svr2_linear_gate_residual_fwd, x.data_ptr(),
              weight.data_ptr(), _ptr(bias), emb.data_ptr(), gate_b.data_ptr(), int(layer), resid.data_ptr(), out.data_ptr(),
              M, N, K, _stream(x))
    return out
