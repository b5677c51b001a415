"""The NaDiT driver around the blocks (comfyui-seedvr2_b200/stack.py): the CUDA-graph replay must return exactly what the
kernel-by-kernel forward returns, for both architectures, and keep doing so when the static inputs change."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _stack(svr2, cuda, model, **kw):
    torch.manual_seed(3)
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        with torch.device(cuda):
            st = svr2.NaDiTStack(model, **kw)
    finally:
        torch.set_default_dtype(prev)
    return st.eval()


@pytest.mark.parametrize("model,kw,grid", [
    ("3b", dict(num_layers=3, vid_dim=256, heads=2, mm_layers=1, txt_in_dim=64), (2, 40, 72)),
    ("7b", dict(num_layers=2, vid_dim=256, heads=2, txt_in_dim=64), (1, 34, 60)),
])
def test_graph_replay_equals_eager(svr2, cuda, model, kw, grid):
    st = _stack(svr2, cuda, model, **kw)
    T, H2, W2 = grid
    txt_len = 58
    g = torch.Generator(device="cpu").manual_seed(5)
    ts = torch.tensor([[txt_len]], dtype=torch.long)
    gf = st.graph_for(grid, txt_len, cuda)
    assert gf.launches_per_replay > 0, "no library kernel was captured: the graph would replay a fallback"
    for trial in range(2):          # second trial: new inputs through the same captured graph
        vid = torch.randn(T * H2 * W2, 33, generator=g).bfloat16().to(cuda)
        txt = torch.randn(txt_len, kw["txt_in_dim"], generator=g).bfloat16().to(cuda)
        t = torch.full((1,), 1000.0 - 300.0 * trial, device=cuda)
        with torch.no_grad():
            eager = st(vid, txt, None, ts, t, grid=grid).clone()
        gf.vid.copy_(vid); gf.txt.copy_(txt); gf.timestep.copy_(t)
        out = gf.replay().clone()
        torch.cuda.synchronize()
        assert torch.isfinite(eager.float()).all()
        assert torch.equal(out, eager), f"trial {trial}: graph replay differs from the kernel-by-kernel forward"


def test_forward_host_round_trip(svr2, cuda):
    """forward_host: pinned host buffers in, pinned host buffer out, graphed and not, same bits."""
    st = _stack(svr2, cuda, "3b", num_layers=2, vid_dim=256, heads=2, mm_layers=1, txt_in_dim=64)
    grid, txt_len = (1, 32, 48), 58
    g = torch.Generator(device="cpu").manual_seed(9)
    vid = torch.randn(grid[0] * grid[1] * grid[2], 33, generator=g).bfloat16().pin_memory()
    txt = torch.randn(txt_len, 64, generator=g).bfloat16().pin_memory()
    o1 = torch.empty(vid.shape[0], 16, dtype=torch.bfloat16).pin_memory()
    o2 = torch.empty_like(o1).pin_memory()
    st.forward_host(vid, txt, grid, 500.0, o1, cuda, graphed=False)
    st.forward_host(vid, txt, grid, 500.0, o2, cuda, graphed=True)
    assert torch.equal(o1, o2)
