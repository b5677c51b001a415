#!/usr/bin/env python
"""bench.py — DiT frames/sec of the SeedVR2 one-step NaDiT hot path on B200 (BASELINE.json `metric`: "DiT frames/sec,
SeedVR2-3B 1080p, 1/2/4/8 B200; window-attn TFLOP/s vs peak").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload v1080|v1080_121|v1080_b|c0|c1|c2|c3|c3r|c4] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

Default workload `v1080`: a 232-frame 1920x1080 stream through SeedVR2-3B with the reference's streaming settings
(batch_size 33, temporal_overlap 4).  One "step" = the WHOLE job: the frames are split across the N ranks by the reference's
own segment arithmetic (inference_cli.py:1087-1114), every rank cuts its segment into 4n+1 latent batches
(generation_phases.py:267-282) and runs one NaDiT forward per batch (all 32 blocks + the patch / embedding glue), then the
stitch runs INSIDE the timed region: crossfade of the overlapping frames between a rank's batches, halo exchange + crossfade
between neighbouring ranks and the gather to rank 0 (inference_cli.py:1163-1198) — one peer-memory kernel per rank
(`svr2_segment_stitch_peer`), NCCL where symmetric memory is unavailable.  Total work is fixed as N grows: `scaling` = "strong",
`value` = 232 frames / (max over ranks of the step time).  232 = 8 x (33 - 4): at 1, 2, 4 and 8 ranks every rank runs 8 / N
forwards of 33 frames (the stream's last batch has 29).  The 121-frame 1080p video of BASELINE config C3 (one forward per rank,
4n+1 padding and overlap recompute limit its scaling) is measured beside it under `also.v1080_121`.

`value` = frames / step time with the latents already resident in HBM (CUDA-graph replay per batch); `e2e` = the same through
the public host-buffer call (`NaDiTStack.forward_host`: pinned host latents -> H2D -> forward -> D2H of the prediction, per
batch) plus the same stitch.  `roofline` = the window-attention kernel, `gemm` = the tcgen05 GEMMs, both timed live with CUDA
events around every launch.  The pixel frames the stitch moves are synthetic stand-ins of the decoded segment (the VAE is §8f N3,
outside this path): same shapes, dtype (fp32) and bytes as the reference's stitch.

`--impl reference`: the reference's CPU implementation of the path — the UNMODIFIED reference block from baseline/_ref (staged by
scripts/stage_reference.py; `kind: "reference"`), else the golden-pinned oracle port — fp32 on all host cores.  One step = ONE
transformer block at the workload's first batch grid, cycling through the model's layer kinds (separate / shared weights x
regular / shifted windows); frames/s extrapolated over layers x batches.  Rank 0 only.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# name: model, latent frame (2H, 2W) before the (1,2,2) patching, total pixel frames, batch_size, temporal_overlap, pixel frame, split
WORKLOADS = {
    "v1080": dict(model="3b", hw=(136, 240), frames=232, batch=33, overlap=4, px=(1080, 1920), split="temporal",
                  desc="SeedVR2-3B 1080p stream, 232 frames, batch_size 33, temporal_overlap 4"),
    "v1080_121": dict(model="3b", hw=(136, 240), frames=121, batch=121, overlap=4, px=(1080, 1920), split="temporal",
                      desc="SeedVR2-3B 121 frames 1920x1080, one batch per rank"),
    "v1080_b": dict(model="3b", hw=(136, 240), frames=33, batch=33, overlap=0, px=(1080, 1920), split="replica",
                    desc="SeedVR2-3B ONE 33-frame 1920x1080 batch (one forward of the v1080 stream; profiling size)"),
    "c0": dict(model="3b", hw=(64, 64), frames=1, batch=1, overlap=0, px=(512, 512), split="replica", desc="SeedVR2-3B 1 frame 512x512"),
    "c1": dict(model="3b", hw=(256, 256), frames=1, batch=1, overlap=0, px=(2048, 2048), split="replica",
               desc="SeedVR2-3B 1 frame 2048x2048 image upscale"),
    "c2": dict(model="3b", hw=(180, 320), frames=33, batch=33, overlap=0, px=(1440, 2560), split="replica",
               desc="SeedVR2-3B 33 frames 2560x1440"),
    "c3": dict(model="7b", hw=(136, 240), frames=121, batch=121, overlap=4, px=(1080, 1920), split="temporal",
               desc="SeedVR2-7B 121 frames 1920x1080"),
    "c4": dict(model="7b", hw=(270, 480), frames=37, batch=37, overlap=0, px=(2160, 3840), split="replica",
               desc="SeedVR2-7B 33+4-frame streaming chunk at 3840x2160"),
    "c3r": dict(model="7b", hw=(136, 240), frames=21, batch=21, overlap=0, px=(1080, 1920), split="replica",
                desc="SeedVR2-7B 21-frame per-rank segment of the 121-frame 1080p video at 8 GPUs"),
}
TXT_LEN, TXT_DIM = 58, 5120
LAYERS = {"3b": 32, "7b": 36}
FALLBACK_PEAKS = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return {k: float(p[k]) for k in FALLBACK_PEAKS}, "measured (MEASURED_PEAKS.json)"
    except Exception:
        return dict(FALLBACK_PEAKS), "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 7:
                continue
            try:
                sm.append(float(c[0])); mx.append(float(c[1])); pw.append(float(c[2]))
            except ValueError:
                continue
            for n, v in zip(names, c[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        self.f.close()
        os.unlink(self.f.name)
        if sm:
            s = sorted(sm)
            out.update(sm_mhz=s[len(s) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm), power_w_max=max(pw))
        return out


# ------------------------------------------------------------------------------------------------------------------
# the job: who runs which batches (pure integer arithmetic, shared by both arms)
# ------------------------------------------------------------------------------------------------------------------
def job_layout(name, world):
    """Per rank: the frames it reads and its latent batches [(first frame in the segment, frames, latent grid (T, 2H, 2W))]."""
    from seedvr2_b200 import segments
    w = WORKLOADS[name]
    H2, W2 = w["hw"]
    if w["split"] == "temporal":
        segs = segments.segment_ranges(w["frames"], world, w["overlap"])
    else:       # a single batch cannot be split in time: every rank runs the whole workload (replicas)
        segs = [(0, w["frames"])] * world
    ranks = []
    for (a, b) in segs:
        bt = segments.batch_ranges(b - a, w["batch"], w["overlap"])
        ranks.append(dict(frames=(a, b), batches=[(s, e - s, (segments.latent_frames(e - s), H2, W2)) for s, e in bt]))
    total = w["frames"] if w["split"] == "temporal" else w["frames"] * world
    return dict(ranks=ranks, total_frames=total, temporal=w["split"] == "temporal")


def tokens_of(grid):
    return grid[0] * (grid[1] // 2) * (grid[2] // 2)


def tok_grid(grid):
    return (grid[0], grid[1] // 2, grid[2] // 2)


def workload_config(name, n):
    w = WORKLOADS[name]
    lay = job_layout(name, n)
    per_rank = [{"frames_read": list(r["frames"]), "forwards": len(r["batches"]),
                 "batches": [{"frames": f, "latent_grid": list(g), "video_tokens": tokens_of(g)} for _, f, g in r["batches"]]}
                for r in lay["ranks"]]
    read = sum(f for r in lay["ranks"] for _, f, _ in r["batches"])
    from seedvr2_b200 import segments
    computed = sum(segments.padded_batch_frames(f) for r in lay["ranks"] for _, f, _ in r["batches"])
    return {
        "workload": f"{w['desc']}: one-step NaDiT forward per latent batch, all {LAYERS[w['model']]} blocks, random-init weights, "
                    f"{TXT_LEN} text tokens; job = {lay['total_frames']} frames",
        "name": name, "frames_per_step": lay["total_frames"], "batch_size": w["batch"], "temporal_overlap": w["overlap"],
        "per_rank": per_rank if n <= 2 else [per_rank[0], "...", per_rank[-1]],
        "padded_work_efficiency": lay["total_frames"] / computed if lay["temporal"] else 1.0,
        "frames_read_incl_overlap": read, "frames_computed_incl_4n1_padding": computed,
        "sharding": ("temporal segments with overlap across ranks (reference split arithmetic); halo exchange + crossfade + gather to "
                     "rank 0 inside the timed region" if lay["temporal"] and n > 1 else
                     "replicas: every rank runs the whole single-batch workload" if n > 1 else "single GPU"),
        "l2": "no explicit flush: every forward streams 6.8-16 GB of bf16 weights plus > 300 MB of activations per layer, far beyond the 126 MB L2",
    }


# ------------------------------------------------------------------------------------------------------------------
# baseline legs: the reference's own block (staged copy) or the oracle port; CPU fp32 and GPU bf16
# ------------------------------------------------------------------------------------------------------------------
def layer_kinds(model):
    """(shared, shifted, number of layers of that kind) — 3B: layers 0-9 separate, 10-31 shared, odd layers shifted."""
    if model == "3b":
        return [(False, False, 5), (False, True, 5), (True, False, 11), (True, True, 11)]
    return [(False, False, 18), (False, True, 18)]


class RefBlockRunner:
    """One transformer block of the reference path at full width: the stock NaMMSRTransformerBlock from baseline/_ref when it
    is staged (`kind` "reference"), else oracle.block.block_forward (`kind` "port")."""

    def __init__(self, model, device, dtype, attention_mode="sdpa"):
        import torch
        from oracle import refstage
        from oracle import workload as owork
        self.torch, self.model, self.device, self.dtype = torch, model, device, dtype
        self.attention_mode = attention_mode
        self.owork = owork
        self.kind = "reference" if refstage.available() else "port"
        self.refstage = refstage
        self.cache = {}

    def _block(self, shared, shifted):
        torch = self.torch
        key = (shared, shifted)
        if key in self.cache:
            return self.cache[key]
        cfg = self.owork.block_cfg(self.model, shared=shared, shifted=shifted)
        wkey = ("w", cfg.shared)
        if wkey not in self.cache:       # regular / shifted layers of one kind share synthetic weights (timing only)
            params = self.owork.block_params(cfg, seed=0)
            self.cache[wkey] = {k: (v.float() if k.endswith("freqs") else v.to(self.dtype)).to(self.device) for k, v in params.items()}
        sd = self.cache[wkey]
        if self.kind == "reference":
            with torch.device("meta"):
                blk = self.refstage.build_block(self.model, cfg.dim, cfg.heads, shared=cfg.shared, shifted=shifted,
                                                attention_mode=self.attention_mode)
            blk.load_state_dict(sd, strict=True, assign=True)
            blk.eval()
        else:
            blk = None
        self.cache[key] = (cfg, sd, blk)
        return self.cache[key]

    def run(self, shared, shifted, grid_tok, vid, txt, emb):
        torch = self.torch
        cfg, sd, blk = self._block(shared, shifted)
        with torch.no_grad():
            if blk is not None:
                vs = torch.tensor([list(grid_tok)], device=self.device)
                ts = torch.tensor([[txt.shape[0]]], device=self.device)
                return blk(vid=vid, txt=txt, vid_shape=vs, txt_shape=ts, emb=emb, cache=self.refstage.new_cache())[:2]
            from oracle import block as oblock
            return oblock.block_forward(cfg, sd, vid, txt, grid_tok, emb)

    def inputs(self, grid_tok):
        cfg = self.owork.block_cfg(self.model)
        vid, txt, emb = self.owork.block_inputs(cfg, grid_tok, TXT_LEN)
        return tuple(t.to(self.dtype).to(self.device) for t in (vid, txt, emb))


def extrapolate_fps(name, block_seconds_by_kind):
    """frames/s of the whole job from per-kind block times measured at the first batch's token grid: every batch costs
    sum_kinds(n_layers(kind) x t(kind)) scaled by its token count (the blocks are linear in tokens at fixed window shapes)."""
    w = WORKLOADS[name]
    lay = job_layout(name, 1)
    b0 = lay["ranks"][0]["batches"][0][2]
    per_forward = sum(n * block_seconds_by_kind[(sh, sf)] for sh, sf, n in layer_kinds(w["model"]))
    total_s = sum(per_forward * tokens_of(g) / tokens_of(b0) for _, _, g in lay["ranks"][0]["batches"])
    return lay["total_frames"] / total_s, per_forward


def cpu_block_times(name, steps, warmup, threads=None):
    """`steps` timed CPU runs of one block, cycling through the layer kinds; returns {kind: mean seconds}, runner kind, threads."""
    import torch
    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    w = WORKLOADS[name]
    g0 = tok_grid(job_layout(name, 1)["ranks"][0]["batches"][0][2])
    runner = RefBlockRunner(w["model"], torch.device("cpu"), torch.float32)
    vid, txt, emb = runner.inputs(g0)
    kinds = [(sh, sf) for sh, sf, _ in layer_kinds(w["model"])]
    times = {k: [] for k in kinds}
    for i in range(warmup + steps):
        k = kinds[i % len(kinds)]
        t0 = time.perf_counter()
        runner.run(k[0], k[1], g0, vid.clone(), txt.clone(), emb)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times[k].append(dt)
    measured = {k: sum(v) / len(v) for k, v in times.items() if v}
    mean = sum(measured.values()) / len(measured)
    return {k: measured.get(k, mean) for k in kinds}, runner.kind, threads, g0, sum(len(v) for v in times.values())


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    name = args.workload
    w = WORKLOADS[name]
    by_kind, kind, threads, g0, n_timed = cpu_block_times(name, args.steps, args.warmup)
    fps, per_forward = extrapolate_fps(name, by_kind)
    ms_step = 1e3 * sum(by_kind.values()) / len(by_kind)
    layers = LAYERS[w["model"]]
    sample = (f"{n_timed} timed runs of ONE transformer block of the {layers}-block forward (the "
              f"{'unmodified reference NaMMSRTransformerBlock from baseline/_ref' if kind == 'reference' else 'oracle port of the reference block'}, "
              f"fp32, torch CPU, {threads} threads) at token grid {g0[0]}x{g0[1]}x{g0[2]}, cycling separate/shared x regular/shifted layer kinds; "
              f"frames/s = job frames / sum over batches and layers of the block time ({per_forward:.1f} s per forward)")
    line = {
        "impl": "reference", "metric": "DiT frames/sec", "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong" if w["split"] == "temporal" else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(name, 1),
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ------------------------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------------------------
class Job:
    """Everything one rank needs to run its share of a workload: the per-batch forwards (CUDA-graph replay or eager), pinned host
    buffers for the end-to-end call, and the stitch of the (synthetic) decoded frames."""

    def __init__(self, name, stack, dev, rank, world, use_graph=True):
        import torch
        import torch.distributed as dist
        from seedvr2_b200 import segments
        self.torch, self.dist, self.segments = torch, dist, segments
        self.name, self.stack, self.dev, self.rank, self.world, self.use_graph = name, stack, dev, rank, world, use_graph
        self.w = WORKLOADS[name]
        self.lay = job_layout(name, world)
        self.mine = self.lay["ranks"][rank]
        self.grids = [g for _, _, g in self.mine["batches"]]
        self.txt_shape = torch.tensor([[TXT_LEN]], dtype=torch.long)
        self.tstep = torch.full((1,), 1000.0, device=dev)
        g = torch.Generator(device="cpu").manual_seed(1234 + rank)
        self.txt_host = torch.randn(TXT_LEN, TXT_DIM, generator=g).bfloat16().pin_memory()
        self.txt_dev = self.txt_host.to(dev)
        self.host, self.devin, self.graphs = {}, {}, {}
        for grid in sorted(set(self.grids)):
            n = grid[0] * grid[1] * grid[2]
            vh = torch.randn(n, 33, generator=g).bfloat16().pin_memory()
            self.host[grid] = (vh, torch.empty(n, 16, dtype=torch.bfloat16).pin_memory())
            self.devin[grid] = vh.to(dev)
            if use_graph:
                gf = stack.graph_for(grid, TXT_LEN, dev)
                gf.vid.copy_(self.devin[grid]); gf.txt.copy_(self.txt_dev); gf.timestep.fill_(1000.0)
                self.graphs[grid] = gf
        self.h2d = sum(self.host[g][0].numel() * 2 + self.txt_host.numel() * 2 for g in self.grids)
        self.d2h = sum(self.host[g][1].numel() * 2 for g in self.grids)
        self._setup_stitch()

    # ---- stitch: in-rank crossfade between batches, then halo exchange + crossfade + gather across ranks
    def _setup_stitch(self):
        torch, dist = self.torch, self.dist
        self.stitch_mode, self.stitcher, self.gather_buf = "none", None, None
        if not self.lay["temporal"]:
            if self.world > 1:       # replicas: the reference's mp.Queue hand-off of every rank's result to the parent
                n = self.grids[0][0] * self.grids[0][1] * self.grids[0][2]
                self.gather_buf = [torch.empty(n, 16, dtype=torch.bfloat16, device=self.dev) for _ in range(self.world)] if self.rank == 0 else None
                self.stitch_mode = "replica gather of the predictions to rank 0 (NCCL)"
            return
        Hp, Wp = self.w["px"]
        self.frame_shape = (Hp, Wp, 3)
        ov = self.w["overlap"]
        self.seg_lens = [b - a for a, b in (r["frames"] for r in self.lay["ranks"])]
        max_seg = max(self.seg_lens)
        max_batch = max(f for _, f, _ in self.mine["batches"])
        gen = torch.Generator(device="cpu").manual_seed(99 + self.rank)
        # the "decoded" frames of a batch: synthetic fp32 stand-in of the VAE output, shared by this rank's batches
        self.decoded = torch.rand((max_batch,) + self.frame_shape, generator=gen).to(self.dev)
        if self.world > 1:
            try:
                self.stitcher = self.segments.PeerStitcher(max_seg, self.lay["total_frames"], self.frame_shape, self.dev)
                self.segbuf = self.stitcher.segment_buffer()
                self.segments.stitch_layout(self.seg_lens, ov)      # raises when a segment is too short for the one-kernel form
                self.stitch_mode = "in-rank crossfade (svr2_halo_blend) + svr2_segment_stitch_peer: halo, crossfade and gather as one kernel per rank over NVLink peer memory"
            except Exception as e:      # no symmetric memory / order-dependent split: the NCCL form of the same exchange
                self.stitcher = None
                self.segbuf = torch.empty((max_seg,) + self.frame_shape, dtype=torch.float32, device=self.dev)
                self.stitch_mode = f"in-rank crossfade (svr2_halo_blend) + NCCL halo send/recv + ragged gather ({type(e).__name__}: peer path unavailable)"
        else:
            self.segbuf = torch.empty((max_seg,) + self.frame_shape, dtype=torch.float32, device=self.dev)
            self.stitch_mode = "in-rank crossfade between overlapping batches (svr2_halo_blend)"

    def stitch(self):
        if not self.lay["temporal"]:
            return
        from seedvr2_b200 import ops
        ov = self.w["overlap"]
        seg = self.segbuf
        for i, (s, f, _) in enumerate(self.mine["batches"]):
            dec = self.decoded[:f]
            if i == 0 or ov == 0:
                seg[s:s + f].copy_(dec)
            else:       # generation_phases.py:960-990: blend the overlap with the previous batch's tail, append the rest
                ops.halo_blend(seg[s:s + ov], dec[:ov], ov, out=seg[s:s + ov])
                seg[s + ov:s + f].copy_(dec[ov:])
        if self.world > 1:
            if self.stitcher is not None:
                self.stitcher.stitch(self.seg_lens, ov)
            else:
                self.segments.halo_exchange_and_gather(seg[:self.seg_lens[self.rank]], ov)

    def _replica_gather(self, out):
        if self.world > 1 and not self.lay["temporal"]:
            self.dist.gather(out, self.gather_buf, dst=0)

    # ---- the three ways a step is run
    def step_eager(self):
        out = None
        for grid in self.grids:
            out = self.stack(self.devin[grid], self.txt_dev, None, self.txt_shape, self.tstep, grid=grid)
        return out

    def step_device(self):
        if self.use_graph:
            out = None
            for grid in self.grids:
                out = self.graphs[grid].replay()
        else:
            out = self.step_eager()
        self.stitch()
        self._replica_gather(out)

    def step_host(self):
        out = None
        for grid in self.grids:
            vh, oh = self.host[grid]
            out = self.stack.forward_host(vh, self.txt_host, grid, 1000.0, oh, self.dev, graphed=self.use_graph)
        self.stitch()
        if self.world > 1 and not self.lay["temporal"]:
            self._replica_gather(out.to(self.dev, non_blocking=True))

    def attention_flops(self, svr2, heads):
        """Algorithmic FLOPs of all attention launches of one step on this rank: heads*4*d*sum_w (Lw+Lt)^2 (attention.py:130-136)."""
        layers = LAYERS[self.w["model"]]
        n_win = (layers + 1) // 2
        tot = 0.0
        for grid in self.grids:
            tg = tok_grid(grid)
            tot += n_win * svr2.get_plan(tg, (4, 3, 3), False, TXT_LEN).attention_flops(heads)
            tot += (layers - n_win) * svr2.get_plan(tg, (4, 3, 3), True, TXT_LEN).attention_flops(heads)
        return tot


def vae_block_sample(svr2, dev):
    """ms and TFLOP/s of one VAE residual block (ResnetBlock3D, attn_video_vae.py:253-361) through vae_block.ResnetBlock3DB200."""
    import torch
    ci = co = 128
    T, H, W = 5, 1080, 1920

    class _N:
        def __init__(self, **kw):
            self.__dict__.update(kw)
    g = torch.Generator(device=dev).manual_seed(0)
    rnd = lambda *s, k=1.0: (torch.randn(*s, generator=g, device=dev) * k).half()      # noqa: E731
    norm = lambda c: _N(num_groups=32, eps=1e-6, weight=1 + rnd(c, k=0.1), bias=rnd(c, k=0.1))      # noqa: E731
    conv = lambda a, b: _N(weight=rnd(b, a, 3, 3, 3, k=(27 * a) ** -0.5), bias=rnd(b, k=0.1), stride=(1, 1, 1), dilation=(1, 1, 1), groups=1)      # noqa: E731
    blk = svr2.vae_block.ResnetBlock3DB200(norm(ci), conv(ci, co), norm(co), conv(co, co))
    x = rnd(1, T, H, W, ci).permute(0, 4, 1, 2, 3)          # channels_last_3d
    with torch.no_grad():
        for _ in range(2):
            blk(x, None, memory_state="DISABLED")
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 5
        e0.record()
        for _ in range(n):
            blk(x, None, memory_state="DISABLED")
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    flops = 2.0 * T * H * W * 27 * (ci * co + co * co)
    return {"ms": ms, "tflops": flops / ms / 1e9, "dtype": "f16", "shape": [ci, co, T, H, W], "frames_per_s": T / (ms * 1e-3),
            "inputs": "5.3 GB activations per call, larger than L2"}


def vae_decoder_sample(svr2, dev):
    """The VAE decoder at 1080p through install_vae_blocks(): Decoder3D.forward composed from the reference's real modules (staged copy,
    tests/vae_standins.py, mid-block attention included), product widths, latent 135 x 240, two temporal slices (1 + 4 output frames)."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle import refstage
    if not refstage.available():
        return {"unavailable": "reference modules not staged (baseline/_ref)"}
    import contextlib
    with contextlib.redirect_stdout(sys.stderr):      # the reference prints a memory banner on import; stdout carries ONE JSON line
        from vae_standins import Decoder3D, reference_parts
        _, _, _, MemoryState = reference_parts()
        torch.manual_seed(0)
        dec = Decoder3D(attention=True).eval().to(dev).half()
    installed = svr2.vae_block.install_vae_blocks(dec)
    dec.use_rebound_norm = True
    zs = [torch.randn(1, 16, 1, 135, 240, device=dev).half() for _ in range(2)]

    def run():
        return [dec(z, memory_state=MemoryState.INITIALIZING if i == 0 else MemoryState.ACTIVE) for i, z in enumerate(zs)]
    with torch.no_grad():
        outs = run()
        frames = sum(o.shape[2] for o in outs)
        del outs
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 3
        e0.record()
        for _ in range(n):
            run()
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    return {"ms_per_frame": ms / frames, "frames_per_s": frames / (ms * 1e-3), "output_frames": frames, "dtype": "f16",
            "modules_replaced": list(installed), "reference_same_gpu": "profiles/r02_run37-59_vae_decoder_encoder_1080p.jsonl (153-274 ms per frame)"}


def vae_encoder_sample(svr2, dev):
    """The VAE encoder at 1080p through install_vae_blocks(): Encoder3D.forward composed from the reference's real modules (staged copy,
    tests/vae_standins.py, mid-block attention included), product widths, two temporal slices of 5 + 4 input frames."""
    import contextlib
    import torch
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle import refstage
    if not refstage.available():
        return {"unavailable": "reference modules not staged (baseline/_ref)"}
    with contextlib.redirect_stdout(sys.stderr):
        from vae_standins import Encoder3D, reference_parts
        _, _, _, MemoryState = reference_parts()
        torch.manual_seed(0)
        enc = Encoder3D(attention=True).eval().to(dev).half()
    installed = svr2.vae_block.install_vae_blocks(enc)
    enc.use_rebound_norm = True
    xs = [torch.randn(1, 3, t, 1080, 1920, device=dev).half() for t in (5, 4)]

    def run():
        return [enc(x, memory_state=MemoryState.INITIALIZING if i == 0 else MemoryState.ACTIVE) for i, x in enumerate(xs)]
    with torch.no_grad():
        run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 3
        e0.record()
        for _ in range(n):
            run()
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    return {"ms_per_frame": ms / 9, "frames_per_s": 9 / (ms * 1e-3), "input_frames": 9, "dtype": "f16", "modules_replaced": list(installed),
            "reference_same_gpu": "profiles/r02_run37-59_vae_decoder_encoder_1080p.jsonl (58.8-62.8 ms per frame)"}


def run_ours(args):
    import torch
    import torch.distributed as dist
    import seedvr2_b200 as svr2
    from seedvr2_b200 import ops

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the seedvr2_b200 hot path is CUDA-only (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    svr2._lib.lib()          # fail loudly if libseedvr2_b200.so is missing

    name = args.workload
    w = WORKLOADS[name]
    model = w["model"]
    # random-init weights of the named architecture, created on the device in bf16 (no checkpoints / network here)
    torch.manual_seed(0)
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    with torch.device(dev):
        stack = svr2.NaDiTStack(model)
    torch.set_default_dtype(prev)
    stack.eval()
    layers, heads, D = stack.cfg["num_layers"], stack.cfg["heads"], stack.cfg["vid_dim"]
    use_graph = not args.no_graph
    job = Job(name, stack, dev, rank, world, use_graph)
    total_frames = job.lay["total_frames"]

    def timed(fn, steps):
        """barrier + sync, K steps between two CUDA events on the current stream, sync + barrier; max over ranks."""
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    warm = max(args.warmup, 3)
    for _ in range(warm):
        job.step_device()
    torch.cuda.synchronize()

    # ---- timed region 1: device-resident inputs (value)
    clocks = ClockSampler(local_rank) if rank == 0 else None
    n0 = ops.launch_count()
    ms_total = timed(job.step_device, args.steps)
    launches = ops.launch_count() - n0
    ms_per_step = ms_total / args.steps
    value = total_frames / (ms_per_step * 1e-3)

    # ---- timed region 1b: this rank's forwards launched kernel by kernel with CUDA events around every attention launch
    # (events cannot bracket a kernel inside a graph replay); the clock sampler keeps running over both regions
    steps_1b = max(1, min(args.steps, 3))
    timer = ops.KernelTimer(only={"svr2_window_attn_fwd"})
    job.step_eager()
    ops.set_timer(timer)
    ms_eager_total = timed(job.step_eager, steps_1b)
    ops.set_timer(None)
    clk = clocks.stop() if clocks else None
    attn_n, attn_ms = timer.summary().get("svr2_window_attn_fwd", (0, 0.0))

    # ---- timed region 2: host buffers through the public call (e2e)
    for _ in range(2):
        job.step_host()
    ms_e2e = timed(job.step_host, args.steps) / args.steps
    e2e_value = total_frames / (ms_e2e * 1e-3)

    # ---- per-kernel pass (outside the headline regions): every C-ABI kernel bracketed by events, one step
    timer_all = ops.KernelTimer()
    torch.cuda.synchronize()
    job.step_eager()
    ops.set_timer(timer_all)
    job.step_eager()
    torch.cuda.synchronize()
    ops.set_timer(None)
    per_kernel = timer_all.summary()
    per_launch = timer_all.durations()

    # ---- the 121-frame video of BASELINE config C3 at the same ranks (device-resident region only)
    also = None
    if name == "v1080" and not args.no_also:
        job2 = Job("v1080_121", stack, dev, rank, world, use_graph)
        for _ in range(3):
            job2.step_device()
        k2 = max(2, min(args.steps, 5))
        ms2 = timed(job2.step_device, k2) / k2
        cfg2 = workload_config("v1080_121", world)
        also = {"v1080_121": {"value": job2.lay["total_frames"] / (ms2 * 1e-3), "unit": "frames/s", "ms_per_step": ms2, "steps": k2,
                              "padded_work_efficiency": cfg2["padded_work_efficiency"],
                              "frames_computed_incl_4n1_padding": cfg2["frames_computed_incl_4n1_padding"],
                              "latent_frames_per_rank": [r["batches"][0][2][0] for r in job2.lay["ranks"]],
                              "workload": cfg2["workload"], "stitch": job2.stitch_mode}}
        del job2

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---- SURVEY §8f N3 beside the headline: one residual block of the VAE decoder's top level (128 channels at 1080 x 1920, 5
    # frames, fp16) — GroupNorm + SiLU, two causal 3-D convolutions, fused residual; reported, not part of `value`
    if also is not None and world == 1:
        for key, fn in (("vae_resnet_block_1080p", vae_block_sample), ("vae_decoder_1080p", vae_decoder_sample),
                        ("vae_encoder_1080p", vae_encoder_sample)):
            try:
                torch.cuda.empty_cache()
                also[key] = fn(svr2, dev)
            except Exception as e:      # noqa: BLE001 — a side measurement must not take the headline down
                also[key] = {"error": f"{type(e).__name__}: {e}"[:200]}

    pk, pk_src = peaks()
    attn_flops_step = job.attention_flops(svr2, heads)
    achieved_tf = attn_flops_step * steps_1b / (attn_ms * 1e-3) / 1e12 if attn_n else 0.0
    g0 = tok_grid(job.grids[0])
    plans = [svr2.get_plan(g0, (4, 3, 3), sh, TXT_LEN) for sh in (False, True)]
    short = all(pl.packed_rows <= 576 * pl.n_windows for pl in plans)          # the library's dispatch rule (window_attn.cu)
    attn_kernel = ("generation 8, window_attn_tc6_kernel: CTA pairs, tcgen05.mma.cta_group::2 + TMA + TMEM" if short else
                   "generation 4, window_attn_tc2_kernel: TMA + tcgen05 + TMEM")
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(name, {}).get("svr2_window_attn_fwd")
    except Exception:
        pass
    roofline = {
        "kernel": "svr2_window_attn_fwd (" + attn_kernel + ")", "bound": "tensor",
        "achieved": achieved_tf, "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s",
        "frac": achieved_tf / pk["bf16_tflops_sustained"], "frac_of_burst_peak": achieved_tf / pk["bf16_tflops"],
        "peak_source": pk_src + ", sustained bf16 figure (kernel timed inside a long step)",
        "algorithmic_flops_per_launch": attn_flops_step * steps_1b / max(attn_n, 1), "avg_launch_ms": attn_ms / max(attn_n, 1),
        "launches_timed": attn_n, "share_of_step": attn_ms / ms_eager_total if ms_eager_total else None, "traffic": traffic,
        "timing": f"CUDA events around every attention launch over {steps_1b} steps of this rank's forwards launched kernel by kernel "
                  "(region 1b); the headline region replays the same kernels from CUDA graphs",
    }
    # ---- tcgen05 GEMMs (csrc/gemm.cu): algorithmic FLOPs of the block linears of one step / their summed launch time
    mlp = 3 * D * 6912 if model == "3b" else 2 * D * 4 * D
    lin_flops = sum(layers * (tokens_of(g) + TXT_LEN) * 2 * (4 * D * D + mlp) for g in job.grids)
    gemm_names = ("svr2_linear_fwd", "svr2_linear_swiglu_fwd", "svr2_linear_gate_residual_fwd", "svr2_linear_qkv_pack_fwd")
    gemm_ms = sum(per_kernel.get(n, (0, 0.0))[1] for n in gemm_names)
    gemm_n = sum(per_kernel.get(n, (0, 0.0))[0] for n in gemm_names)
    gemm = None
    if gemm_n:
        gtf = lin_flops / (gemm_ms * 1e-3) / 1e12
        gemm = {"kernel": "svr2_linear_qkv_pack_fwd / _swiglu_fwd / _gate_residual_fwd / svr2_linear_fwd (linear_tc_kernel: CTA pairs, tcgen05.mma.cta_group::2, "
                          "TMA ring, double-buffered TMEM accumulator, fused epilogues)", "bound": "tensor", "achieved": gtf,
                "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s", "frac": gtf / pk["bf16_tflops_sustained"],
                "frac_of_burst_peak": gtf / pk["bf16_tflops"], "algorithmic_flops_per_step": lin_flops, "ms_per_step": gemm_ms,
                "launches_per_step": gemm_n, "share_of_step": gemm_ms / (ms_eager_total / steps_1b) if ms_eager_total else None,
                "timing": "CUDA events around every GEMM launch, one step launched kernel by kernel (per-kernel pass)"}
    if gemm is not None:      # the step's dominant kernels by time, summarised beside the attention roofline the metric names
        roofline["gemm"] = {k: gemm[k] for k in ("bound", "achieved", "peak", "unit", "frac", "share_of_step")}
    # ---- HBM-bound kernels: algorithmic bytes per launch, bf16
    qkvb = 3 * heads * 128 * 2
    n_b = len(job.grids)
    mean_L = sum(tokens_of(g) for g in job.grids) / n_b
    mean_P = sum(0.5 * (svr2.get_plan(tok_grid(g), (4, 3, 3), False, TXT_LEN).packed_rows +
                        svr2.get_plan(tok_grid(g), (4, 3, 3), True, TXT_LEN).packed_rows) for g in job.grids) / n_b
    kernels = {}
    for kname, (n, ms) in sorted(per_kernel.items()):
        kernels[kname] = {"launches_per_step": n, "ms_per_step": ms}
    if "svr2_qkv_pack_fwd" in kernels:
        n, ms = per_kernel["svr2_qkv_pack_fwd"]
        gbs = ((mean_L + TXT_LEN) * qkvb + mean_P * qkvb) * n / (ms * 1e-3) / 1e9
        kernels["svr2_qkv_pack_fwd"].update(achieved_gbs=gbs, frac_of_hbm_peak=gbs / pk["hbm_gbs"])
    # Row-wise kernels are launched on the video rows ([L,D], or the joined [L+Lt,D] of a shared-weight layer) and on the
    # few text rows; only the former are bandwidth-sized.  HBM fraction = algorithmic bytes of a video-row launch / the
    # mean duration of those launches (the ones lasting at least a third of the family's longest).
    Fh = {"3b": 6912, "7b": 4 * D}[model]
    for kname, per_row in (("svr2_modnorm_fwd", 4 * D), ("svr2_gate_residual_fwd", 6 * D),
                           ("svr2_gate_residual_modnorm_fwd", 8 * D), ("svr2_swiglu_fwd", 6 * Fh)):
        if kname in kernels and per_launch.get(kname):
            d = per_launch[kname]
            big = [x for x in d if x >= max(d) / 3]
            us = 1e3 * sum(big) / len(big)
            gbs = per_row * mean_L / (us * 1e-6) / 1e9
            kernels[kname].update(video_row_launches_per_step=len(big), us_per_video_row_launch=us,
                                  achieved_gbs=gbs, frac_of_hbm_peak=gbs / pk["hbm_gbs"], bytes_per_launch=per_row * mean_L)

    # ---- baselines (rank 0, N = 1 only): the reference's block on this box's host cores and on this GPU
    cpu_baseline = ref_gpu = None
    if world == 1 and not args.no_ref_gpu:
        try:
            ref_gpu = ref_torch_gpu(name, dev, value)
        except Exception as e:       # a baseline leg must never take the bench line down
            ref_gpu = {"unavailable": f"{type(e).__name__}: {e}"[:300]}
    if world == 1 and not args.no_cpu_baseline:
        by_kind, kind, threads, g0c, n_timed = cpu_block_times(name, steps=2, warmup=1)
        fps, per_forward = extrapolate_fps(name, by_kind)
        cpu_baseline = {"value": fps, "unit": "frames/s", "cores": threads, "kind": kind,
                        "sample": f"{'unmodified reference block (baseline/_ref)' if kind == 'reference' else 'oracle port of the reference block'}, fp32 CPU: "
                                  f"{n_timed} timed runs of ONE block of {layers} at token grid {g0c[0]}x{g0c[1]}x{g0c[2]} "
                                  f"({sum(by_kind.values()) / len(by_kind):.2f} s/block), extrapolated over layers and batches"}

    line = {
        "metric": "DiT frames/sec", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": warm, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "strong" if job.lay["temporal"] else "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": dict(workload_config(name, world), stitch=job.stitch_mode),
        "roofline": roofline, "gemm": gemm, "cpu_baseline": cpu_baseline, "ref_torch_gpu": ref_gpu,
        "e2e": {"value": e2e_value, "unit": "frames/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": job.h2d, "d2h_bytes_per_step": job.d2h},
        "gpu_launches": launches, "clocks": clk, "kernels": kernels, "also": also,
        "attention_tflops": achieved_tf,
        "launch_mode": ("cuda-graph replay, one graph per latent grid (%s library kernels per replay)" %
                        "/".join(str(g.launches_per_replay) for g in job.graphs.values())) if use_graph else "kernel by kernel",
        "ms_per_step_kernel_by_kernel": ms_eager_total / steps_1b,
    }
    emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


def ref_torch_gpu(name, dev, our_value):
    """The reference's own PyTorch path on THIS GPU (SURVEY §8d "the number to beat"): the stock block from baseline/_ref (or the
    oracle port) in bf16, CUDA-event timed per layer kind at the first batch's token grid and extrapolated over layers and batches
    like the CPU arm — with attention_mode='sdpa' (one F.scaled_dot_product_attention per window, the reference's default) and,
    where the flash-attn package runs on this GPU, attention_mode='flash_attn' (flash_attn_varlen_func, attention.py:64-91)."""
    import torch
    w = WORKLOADS[name]
    g0 = tok_grid(job_layout(name, 1)["ranks"][0]["batches"][0][2])

    def one_mode(mode):
        runner = RefBlockRunner(w["model"], dev, torch.bfloat16, attention_mode=mode)
        if mode != "sdpa" and runner.kind != "reference":
            raise RuntimeError("the oracle port restates the sdpa path only")
        vid, txt, emb = runner.inputs(g0)
        by_kind = {}
        for sh, sf, _ in layer_kinds(w["model"]):
            runner.run(sh, sf, g0, vid.clone(), txt.clone(), emb)
            torch.cuda.synchronize()
            ts = []
            for _ in range(3):
                v, t = vid.clone(), txt.clone()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                runner.run(sh, sf, g0, v, t, emb)
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1) * 1e-3)
            by_kind[(sh, sf)] = sorted(ts)[1]
        fps, per_forward = extrapolate_fps(name, by_kind)
        return runner.kind, fps, per_forward, by_kind

    kind, fps, per_forward, by_kind = one_mode("sdpa")
    out = {"value": fps, "unit": "frames/s", "kind": kind, "dtype": "bf16", "attention_mode": "sdpa",
           "ms_per_block": {f"{'shared' if sh else 'separate'}_{'swin' if sf else 'win'}": 1e3 * t for (sh, sf), t in by_kind.items()},
           "ms_per_forward": 1e3 * per_forward, "ours_over_reference_gpu": our_value / fps,
           "sample": f"median of 3 eager runs of one block per layer kind at token grid {g0[0]}x{g0[1]}x{g0[2]} on this GPU, "
                     "extrapolated over layers and batches (the reference's shell ops outside the blocks are not counted)"}
    try:
        _, fps_fa, per_forward_fa, _ = one_mode("flash_attn")
        out["flash_attn"] = {"value": fps_fa, "unit": "frames/s", "ms_per_forward": 1e3 * per_forward_fa,
                             "ours_over_reference_gpu": our_value / fps_fa, "attention_mode": "flash_attn (flash_attn_varlen_func)"}
    except Exception as e:       # package absent, or its kernels do not run on sm_100
        out["flash_attn"] = {"unavailable": f"{type(e).__name__}: {e}"[:200]}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="v1080", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-gpu", action="store_true", help="skip the reference-PyTorch-path-on-this-GPU baseline leg")
    ap.add_argument("--no-also", action="store_true", help="skip the 121-frame side measurement of the default workload")
    ap.add_argument("--no-graph", action="store_true", help="launch the forwards kernel by kernel instead of replaying their CUDA graphs")
    args = ap.parse_args()
    # stdout carries exactly ONE JSON line: everything else that writes to file descriptor 1 — Python prints of imported code, but
    # also C libraries (NCCL prints its version banner there) — is sent to stderr; emit() writes the line to the real stdout
    global _JSON_OUT
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


_JSON_OUT = None


def emit(line: dict) -> None:
    out = _JSON_OUT if _JSON_OUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


if __name__ == "__main__":
    sys.exit(main())
