"""Minimal NaDiT driver around the B200 blocks — for bench.py / smoke() / tests on a box WITHOUT the reference.

In production the reference's own ``NaDiT`` (src/models/dit_3b/nadit.py:39-248, dit_7b/nadit.py:39-191) drives the
blocks after ``install()`` swapped the block class.  The reference is not present on the GPU box, so this file restates
the thin glue AROUND the block loop with plain torch (SURVEY §8f row N2 — "next", not hot path): text projection,
patchify + Linear, timestep embedding, output norm + modulation (incl. cache-key quirk Q1), un-patchify.
Parameter names equal the reference's (``vid_in.proj.weight``, ``emb_in.proj_in.weight``, ``blocks.N. ...``,
``vid_out_norm.weight``, ``vid_out_ada.out_shift`` ...), checked against the real NaDiT by tests/golden/make_golden.py.
"""
import math
from typing import Optional, Tuple

import torch
import torch.nn.functional as F
from torch import nn

from .block import NaMMSRTransformerBlockB200, NaMMSRTransformerBlockB200_7B, _Weight

WIN_METHODS = ("720pwin_by_size_bysize", "720pswin_by_size_bysize")

CONFIGS = {
    # configs_3b/main.yaml:11-36
    "3b": dict(vid_in_channels=33, vid_out_channels=16, vid_dim=2560, txt_in_dim=5120, heads=20, head_dim=128,
               expand_ratio=4, num_layers=32, mm_layers=10, patch_size=(1, 2, 2), window=(4, 3, 3), out_norm=True),
    # configs_7b/main.yaml:11-33
    "7b": dict(vid_in_channels=33, vid_out_channels=16, vid_dim=3072, txt_in_dim=5120, heads=24, head_dim=128,
               expand_ratio=4, num_layers=36, mm_layers=None, patch_size=(1, 2, 2), window=(4, 3, 3), out_norm=False),
}


class _Proj(nn.Module):
    def __init__(self, i: int, o: int):
        super().__init__()
        self.proj = nn.Linear(i, o)


class _TimeEmbedding(nn.Module):
    """src/models/dit_3b/embedding.py:25-62 (+ diffusers get_timestep_embedding, flip_sin_to_cos=False, shift 0)."""

    def __init__(self, hidden: int, out: int, sinusoidal_dim: int = 256):
        super().__init__()
        self.sinusoidal_dim = sinusoidal_dim
        self.proj_in = nn.Linear(sinusoidal_dim, hidden)
        self.proj_hid = nn.Linear(hidden, hidden)
        self.proj_out = nn.Linear(hidden, out)

    def forward(self, timestep: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
        half = self.sinusoidal_dim // 2
        exponent = -math.log(10000) * torch.arange(half, dtype=torch.float32, device=timestep.device) / half
        ang = timestep[:, None].float() * torch.exp(exponent)[None, :]
        emb = torch.cat([torch.sin(ang), torch.cos(ang)], dim=-1).to(dtype)
        return self.proj_out(F.silu(self.proj_hid(F.silu(self.proj_in(emb)))))


class _OutAda(nn.Module):
    def __init__(self, dim: int):
        super().__init__()
        self.out_shift = nn.Parameter(torch.randn(dim) / dim ** 0.5)
        self.out_scale = nn.Parameter(torch.randn(dim) / dim ** 0.5 + 1)


class NaDiTStack(nn.Module):
    def __init__(self, model: str = "3b", num_layers: Optional[int] = None, vid_dim: Optional[int] = None,
                 heads: Optional[int] = None, mm_layers: Optional[int] = None, txt_in_dim: Optional[int] = None):
        super().__init__()
        c = dict(CONFIGS[model])
        for k, v in (("num_layers", num_layers), ("vid_dim", vid_dim), ("heads", heads), ("mm_layers", mm_layers),
                     ("txt_in_dim", txt_in_dim)):
            if v is not None:
                c[k] = v
        self.model, self.cfg = model, c
        D, pt, ph, pw = c["vid_dim"], *c["patch_size"]
        self.patch = (pt, ph, pw)
        self.vid_in = _Proj(c["vid_in_channels"] * pt * ph * pw, D)
        self.txt_in = nn.Linear(c["txt_in_dim"], D)
        self.emb_in = _TimeEmbedding(D, 6 * D)
        n = c["num_layers"]
        blocks = []
        for i in range(n):
            common = dict(vid_dim=D, txt_dim=D, emb_dim=6 * D, heads=c["heads"], head_dim=c["head_dim"],
                          expand_ratio=c["expand_ratio"], norm_eps=1e-5, qk_bias=False, window=c["window"],
                          window_method=WIN_METHODS[i % 2])
            if model == "3b":
                blocks.append(NaMMSRTransformerBlockB200(
                    **common, mlp_type="swiglu", shared_weights=not (i < c["mm_layers"]), rope_type="mmrope3d",
                    rope_dim=128, is_last_layer=(i == n - 1)))
            else:
                blocks.append(NaMMSRTransformerBlockB200_7B(**common, qk_rope=True, shared_qkv=False, shared_mlp=False,
                                                            mlp_type="normal"))
        self.blocks = nn.ModuleList(blocks)
        if c["out_norm"]:
            self.vid_out_norm = _Weight(D)
            self.vid_out_ada = _OutAda(D)
        else:
            self.vid_out_norm = None
        self.vid_out = _Proj(D, c["vid_out_channels"] * pt * ph * pw)

    @torch.no_grad()
    def forward(self, vid: torch.Tensor, txt: torch.Tensor, vid_shape: torch.Tensor, txt_shape: torch.Tensor,
                timestep: torch.Tensor, grid: Optional[Tuple[int, int, int]] = None) -> torch.Tensor:
        """vid [T*H2*W2, 33] (pre-patch latent rows), txt [Lt, 5120], vid_shape [[T, H2, W2]] -> [T*H2*W2, 16].
        `grid` = (T, H2, W2) as Python ints avoids a device sync when vid_shape is on the GPU."""
        pt, ph, pw = self.patch
        T2, H2, W2 = grid if grid is not None else (int(v) for v in vid_shape[0].tolist())
        if pt != 1 or H2 % ph or W2 % pw:
            raise NotImplementedError("NaDiTStack: patch (1,2,2) on even latent grids only")
        C = vid.shape[-1]
        T, H, W = T2, H2 // ph, W2 // pw
        txt = self.txt_in(txt)
        # "(T t)(H h)(W w) c -> T H W (t h w c)"  (patch/patch_v1.py:91)
        x = vid.view(T, H, ph, W, pw, C).permute(0, 1, 3, 2, 4, 5).reshape(T * H * W, ph * pw * C)
        x = self.vid_in.proj(x)
        tok_shape = torch.tensor([[T, H, W]], dtype=torch.long, device=vid.device) if grid is None else \
            torch.tensor([[T, H, W]], dtype=torch.long)       # host tensor: the blocks read it without a sync
        emb = self.emb_in(timestep.to(vid.device), x.dtype)
        for blk in self.blocks:
            x, txt, tok_shape, txt_shape = blk(vid=x, txt=txt, vid_shape=tok_shape, txt_shape=txt_shape, emb=emb, cache=None)
        if self.vid_out_norm is not None:
            # CustomRMSNorm affine (normalization.py:88-109) then vid_out_ada; quirk Q1: the cache hands it the ATTN
            # (shift, scale) of `emb` (dit_3b/nadit.py:236-244) — reproduced, not fixed.
            from . import ops
            from .block import _bf16
            x = ops.modnorm_affine(x.contiguous(), _bf16(self.vid_out_norm.weight), _bf16(emb).reshape(-1).contiguous(),
                                   _bf16(self.vid_out_ada.out_shift), _bf16(self.vid_out_ada.out_scale), 0, 1e-5)
        x = self.vid_out.proj(x)
        # "T H W (t h w c) -> (T t)(H h)(W w) c"  (patch/patch_v1.py:121)
        co = x.shape[-1] // (ph * pw)
        return x.view(T, H, W, ph, pw, co).permute(0, 1, 3, 2, 4, 5).reshape(T * H2 * W2, co)


    @torch.no_grad()
    def forward_host(self, vid_host: torch.Tensor, txt_host: torch.Tensor, grid: Tuple[int, int, int], timestep: float,
                     out_host: torch.Tensor, device: torch.device, graphed: bool = True) -> torch.Tensor:
        """End-to-end call with HOST buffers (what a pipeline stage hands over): pinned host latents/text are copied to
        the device, one NaDiT forward runs, the predicted latent is copied back into the pinned ``out_host``.
        Stream-ordered on the current stream; returns ``out_host`` after the copy-back completed.
        ``graphed`` replays the CUDA graph of this (grid, text length) — captured on first use — instead of launching
        the ~1,500 kernels of a forward one by one."""
        if graphed:
            g = self.graph_for(grid, txt_host.shape[0], device, vid_host.dtype)
            g.vid.copy_(vid_host, non_blocking=True)
            g.txt.copy_(txt_host, non_blocking=True)
            g.timestep.fill_(float(timestep))
            out = g.replay()
        else:
            vid = vid_host.to(device, non_blocking=True)
            txt = txt_host.to(device, non_blocking=True)
            ts = torch.tensor([[txt.shape[0]]], dtype=torch.long)
            t = torch.full((1,), float(timestep), dtype=torch.float32, device=device)
            out = self.forward(vid, txt, None, ts, t, grid=grid)
        out_host.copy_(out, non_blocking=True)
        torch.cuda.current_stream(device).synchronize()
        return out_host

    def graph_for(self, grid: Tuple[int, int, int], txt_len: int, device: torch.device,
                  dtype: torch.dtype = torch.bfloat16) -> "GraphedForward":
        """The captured forward for one (latent grid, text length); cached — a video pipeline sees a handful of shapes."""
        cache = self.__dict__.setdefault("_graphs", {})
        key = (tuple(grid), int(txt_len), str(device), dtype)
        if key not in cache:
            cache[key] = GraphedForward(self, grid, txt_len, device, dtype)
        return cache[key]


class GraphedForward:
    """One NaDiT forward captured as a CUDA graph on static input / output buffers.

    A forward at 2048x2048 is ~1,500 launches, many of them the text branch's few-microsecond kernels; launched one by
    one from Python the GPU idles ~13 % of the step between them (profiles/r01_launch_list_*).  The reference has the
    same shape of problem and answers it with ``torch.compile`` (model_configuration.py:1217-1219); here the launches
    are already hand-written kernels, so the answer is stream capture: no tracing, no recompilation, one launch.
    Everything the forward touches is stream-ordered on the capture stream (the C-ABI takes the stream explicitly) and
    the window plan is resolved at capture time, so the graph holds no host work.

    ``vid`` / ``txt`` / ``timestep`` are the static inputs (write them, then ``replay()``); the returned tensor is the
    static output, overwritten by the next replay."""

    def __init__(self, stack: "NaDiTStack", grid: Tuple[int, int, int], txt_len: int, device: torch.device,
                 dtype: torch.dtype = torch.bfloat16, warmup: int = 2):
        from . import ops
        T2, H2, W2 = grid
        c = stack.cfg
        self.grid = tuple(grid)
        self.vid = torch.zeros(T2 * H2 * W2, c["vid_in_channels"], dtype=dtype, device=device)
        self.txt = torch.zeros(txt_len, c["txt_in_dim"], dtype=dtype, device=device)
        self.timestep = torch.zeros(1, dtype=torch.float32, device=device)
        self.txt_shape = torch.tensor([[txt_len]], dtype=torch.long)
        cur = torch.cuda.current_stream(device)
        side = torch.cuda.Stream(device)
        side.wait_stream(cur)
        with torch.cuda.stream(side), torch.no_grad():      # plans, cuBLAS workspaces, func attributes: all settled here
            for _ in range(warmup):
                stack.forward(self.vid, self.txt, None, self.txt_shape, self.timestep, grid=self.grid)
        cur.wait_stream(side)
        torch.cuda.synchronize(device)
        n0 = ops.launch_count()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph), torch.no_grad():
            self.out = stack.forward(self.vid, self.txt, None, self.txt_shape, self.timestep, grid=self.grid)
        self.launches_per_replay = ops.launch_count() - n0      # kernels of libseedvr2_b200.so inside the graph
        # The graph holds RAW device pointers of everything the forward resolved at capture time.  The window plans sit in an
        # LRU (plan._PLANS) and the bf16 copies of non-bf16 parameters in block._cast_cache; either can drop its entry while
        # this graph is still replayed, so the graph keeps its own strong references (a replay after ~32 other
        # grids read recycled memory).
        from .block import _cast_cache
        from .plan import get_plan
        tok = (T2, H2 // stack.patch[1], W2 // stack.patch[2])
        self._plans = [get_plan(tok, c["window"], shifted, txt_len) for shifted in (False, True)]
        self._casts = [entry[4] for entry in _cast_cache.values()]

    def replay(self) -> torch.Tensor:
        from . import ops
        self.graph.replay()
        ops.count_launches(self.launches_per_replay)
        return self.out
