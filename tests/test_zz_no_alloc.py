"""svr2_window_attn_fwd_noalloc (static device scratch, no allocation inside the call) against svr2_window_attn_fwd.

there is no reference counterpart.  Collected last on purpose (file name).
Every buffer is allocated BEFORE the calls under test; the calls go straight through the C ABI
with raw pointers.  Two calls in a row on different inputs: the second must not see anything of the first in the shared array."""
import ctypes

import pytest
import torch

import cases

pytestmark = pytest.mark.gpu


def _call_pair(svr2, plan, q, k, v, heads, dev):
    L = svr2._lib.lib()
    T, H, W = plan.grid
    d = plan.on(dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    D = heads * 128
    ref_v = torch.empty(plan.n_tokens, D, dtype=torch.bfloat16, device=dev)
    ref_t = torch.empty(plan.txt_len, D, dtype=torch.bfloat16, device=dev)
    part = torch.empty(plan.n_windows, plan.txt_len, D, dtype=torch.bfloat16, device=dev)
    out_v = torch.full_like(ref_v, float("nan"))
    out_t = torch.full_like(ref_t, float("nan"))
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(dev)
    rc = L.svr2_window_attn_fwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), d["table"].data_ptr(), d["work"].data_ptr(),
                                plan.work.shape[0], None, 0, plan.n_windows, T, H, W, plan.txt_len, heads, plan.packed_rows,
                                ref_v.data_ptr(), part.data_ptr(), ref_t.data_ptr(), stream)
    assert rc == 0, L.svr2_last_error()
    rc = L.svr2_window_attn_fwd_noalloc(q.data_ptr(), k.data_ptr(), v.data_ptr(), d["table"].data_ptr(), d["work"].data_ptr(),
                                        plan.work.shape[0], plan.n_windows, T, H, W, plan.txt_len, heads, plan.packed_rows,
                                        out_v.data_ptr(), out_t.data_ptr(), stream)
    assert rc == 0, L.svr2_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated(dev) == before          # nothing went through torch's allocator either
    return (ref_v, ref_t), (out_v, out_t)


def test_noalloc_matches_workspace_entry_point_twice(svr2, cuda):
    heads = 3
    # call 1: shifted windows (many ragged windows); call 2: another grid and fewer windows — stale slots of call 1 remain
    # in the static array and must not leak into call 2's pooled text rows
    for n, (grid, shifted) in enumerate([((2, 20, 36), True), ((1, 32, 32), False)]):
        plan = svr2.get_plan(grid, (4, 3, 3), shifted, cases.TXT_LEN)
        g = torch.Generator().manual_seed(100 + n)
        q, k, v = (torch.randn(plan.packed_rows, heads, 128, generator=g).bfloat16().to(cuda) for _ in range(3))
        (ref_v, ref_t), (out_v, out_t) = _call_pair(svr2, plan, q, k, v, heads, cuda)
        # same kernels, same inputs: the tolerance of the attention tests is 2e-2; these are in fact bit-equal
        assert float((out_v.float() - ref_v.float()).abs().max()) <= 2e-2
        assert float((out_t.float() - ref_t.float()).abs().max()) <= 2e-2
        assert torch.equal(out_v, ref_v) and torch.equal(out_t, ref_t), f"call {n + 1} differs"


def test_noalloc_rejects_sizes_beyond_the_static_array(svr2, cuda):
    L = svr2._lib.lib()
    plan = svr2.get_plan((9, 90, 160), (4, 3, 3), True, cases.TXT_LEN)          # 196 windows x 58 rows x 20 heads = 58 MB > 32 MiB
    d = plan.on(cuda)
    dummy = torch.zeros(8, dtype=torch.bfloat16, device=cuda)
    rc = L.svr2_window_attn_fwd_noalloc(dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), d["table"].data_ptr(),
                                        d["work"].data_ptr(), plan.work.shape[0], plan.n_windows, 9, 90, 160, plan.txt_len, 20,
                                        plan.packed_rows, dummy.data_ptr(), dummy.data_ptr(), None)
    assert rc == -1
    msg = L.svr2_last_error().decode()
    assert "static array" in msg and "svr2_window_attn_fwd" in msg
