#!/usr/bin/env python
"""bench.py — DiT frames/sec of the SeedVR2 one-step NaDiT hot path on B200 (BASELINE.json `metric`).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c1|c0|c2|c3] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

One "step" = one NaDiT forward (all 32/36 blocks + the thin patch/embedding glue) over one synthetic latent batch of the
named workload.  `value` = frames / step time with the inputs already resident in HBM; `e2e` = the same through the
public host-buffer call (`NaDiTStack.forward_host`: pinned host latents -> H2D -> forward -> D2H of the prediction).
N > 1: the path shards by independent temporal segments — every rank runs the same per-rank workload on its
own segment (weak scaling, no collective inside the forward) and the ranks' predictions are gathered to rank 0 over NCCL
inside the timed region (the reference's `mp.Queue` + shared-memory hand-off, inference_cli.py:1148-1153).

`--impl reference`: the reference's CPU implementation of the same path (its fp32 PyTorch code restated in `oracle/`,
pinned to the real reference by tests/golden) on all host cores; one step = ONE transformer block of the workload (1/layers
of a forward), frames/s extrapolated x layers.  Rank 0 only.
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

WORKLOADS = {
    # name: (model, latent grid (T, 2H, 2W) before 2x2 patching, pixel frames, description)   SURVEY §8 size table
    "c0": ("3b", (1, 64, 64), 1, "SeedVR2-3B 1 frame 512x512"),
    "c1": ("3b", (1, 256, 256), 1, "SeedVR2-3B 1 frame 2048x2048 image upscale"),
    "c2": ("3b", (9, 180, 320), 33, "SeedVR2-3B 33 frames 2560x1440"),
    "c3": ("7b", (31, 136, 240), 121, "SeedVR2-7B 121 frames 1920x1080"),
    "c4": ("7b", (10, 270, 480), 37, "SeedVR2-7B 33+4-frame streaming chunk at 3840x2160"),
    "c3r": ("7b", (6, 136, 240), 21, "SeedVR2-7B 21-frame per-rank segment of the 121-frame 1080p video at 8 GPUs"),
}
TXT_LEN, TXT_DIM = 58, 5120
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


def attention_flops_per_forward(svr2, model, tok_grid, layers, heads):
    """Algorithmic FLOPs, the reference's own formula heads*4*d*sum_w (Lw+Lt)^2 (attention.py:130-136), per layer kind."""
    win = svr2.get_plan(tok_grid, (4, 3, 3), False, TXT_LEN).attention_flops(heads)
    swin = svr2.get_plan(tok_grid, (4, 3, 3), True, TXT_LEN).attention_flops(heads)
    n_win = (layers + 1) // 2
    return n_win * win + (layers - n_win) * swin


# ------------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle (reference's fp32 PyTorch path restated) on host cores
# ------------------------------------------------------------------------------------------------------------------
def cpu_block_seconds(model, tok_grid, steps, warmup, threads=None):
    import torch
    from oracle import block as oblock
    from oracle import workload as owork
    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    cfg = owork.block_cfg(model, shared=False, shifted=False)
    params = owork.block_params(cfg, seed=0)
    vid, txt, emb = owork.block_inputs(cfg, tok_grid, TXT_LEN)
    times = []
    with torch.no_grad():
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            oblock.block_forward(cfg, params, vid, txt, tok_grid, emb)
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt)
    return sum(times) / len(times), threads


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    model, lat, frames, desc = WORKLOADS[args.workload]
    tok_grid = (lat[0], lat[1] // 2, lat[2] // 2)
    layers = 32 if model == "3b" else 36
    sec, threads = cpu_block_seconds(model, tok_grid, args.steps, args.warmup)
    per_rank_fps = frames / (sec * layers)
    sample = (f"one transformer block (layer 0: separate vid/txt weights, regular windows) of the {layers}-block forward per step, "
              f"fp32, torch CPU, {threads} threads; frames/s = frames / (block_s * {layers})")
    line = {
        "impl": "reference", "metric": "DiT frames/sec", "value": per_rank_fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.workload, 1),
        "cpu_baseline": {"value": per_rank_fps, "unit": "frames/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": per_rank_fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def workload_config(name, n):
    model, lat, frames, desc = WORKLOADS[name]
    T, H2, W2 = lat
    return {
        "workload": f"{desc}: one-step NaDiT forward, latent grid {T}x{H2}x{W2} -> token grid {T}x{H2 // 2}x{W2 // 2} "
                    f"({T * H2 * W2 // 4} video tokens + {TXT_LEN} text tokens), all {32 if model == '3b' else 36} blocks, random-init weights",
        "name": name, "frames_per_step_per_gpu": frames,
        "sharding": "replicas of the model, independent temporal segments per rank; final gather to rank 0 over NCCL" if n > 1 else "single GPU",
        "l2": "no explicit flush: each step streams the 6.8-16 GB of bf16 weights plus >80 MB activations per layer, far beyond the 126 MB L2",
    }


# ------------------------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------------------------
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

    model, lat, frames, desc = WORKLOADS[args.workload]
    T, H2, W2 = lat
    tok_grid = (T, H2 // 2, W2 // 2)
    L = tok_grid[0] * tok_grid[1] * tok_grid[2]

    # random-init weights of the named architecture, created on the device in bf16 (no checkpoints / network here)
    torch.manual_seed(0)
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    with torch.device(dev):
        stack = svr2.NaDiTStack(model)
    torch.set_default_dtype(prev)
    stack.eval()
    layers, heads, D = stack.cfg["num_layers"], stack.cfg["heads"], stack.cfg["vid_dim"]

    g = torch.Generator(device="cpu").manual_seed(1234 + rank)
    vid_host = torch.randn(T * H2 * W2, 33, generator=g).bfloat16().pin_memory()
    txt_host = torch.randn(TXT_LEN, TXT_DIM, generator=g).bfloat16().pin_memory()
    out_host = torch.empty(T * H2 * W2, 16, dtype=torch.bfloat16).pin_memory()
    vid_dev, txt_dev = vid_host.to(dev), txt_host.to(dev)
    txt_shape = torch.tensor([[TXT_LEN]], dtype=torch.long)
    tstep = torch.full((1,), 1000.0, device=dev)
    gather_buf = [torch.empty(T * H2 * W2, 16, dtype=torch.bfloat16, device=dev) for _ in range(world)] if (world > 1 and rank == 0) else None

    # The product path replays the CUDA graph of the forward (stack.GraphedForward): inputs are written into its static
    # buffers, one graph launch runs the ~1,500 kernels.  --no-graph launches them one by one (the round-1 behaviour).
    use_graph = not args.no_graph
    gf = None
    if use_graph:
        gf = stack.graph_for(lat, TXT_LEN, dev)
        gf.vid.copy_(vid_dev)
        gf.txt.copy_(txt_dev)
        gf.timestep.fill_(1000.0)

    def step_eager():
        return stack(vid_dev, txt_dev, None, txt_shape, tstep, grid=lat)

    def step_device():
        out = gf.replay() if use_graph else step_eager()
        if world > 1:
            dist.gather(out, gather_buf, dst=0)
        return out

    def step_host():
        stack.forward_host(vid_host, txt_host, lat, 1000.0, out_host, dev, graphed=use_graph)
        if world > 1:
            dist.gather(out_host.to(dev, non_blocking=True), gather_buf, dst=0)

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

    for _ in range(max(args.warmup, 3)):
        step_device()
    torch.cuda.synchronize()

    # ---- timed region 1: device-resident inputs (value)
    clocks = ClockSampler(local_rank) if rank == 0 else None
    n0 = ops.launch_count()
    ms_total = timed(step_device, args.steps)
    launches = ops.launch_count() - n0
    ms_per_step = ms_total / args.steps
    value = world * frames / (ms_per_step * 1e-3)

    # ---- timed region 1b: the same K steps launched kernel by kernel with CUDA events around every attention launch
    # (events cannot bracket a kernel inside a graph replay); the clock sampler keeps running over both regions
    timer = ops.KernelTimer(only={"svr2_window_attn_fwd"})
    for _ in range(2):
        step_eager()
    ops.set_timer(timer)
    ms_eager_total = timed(step_eager, args.steps)
    ops.set_timer(None)
    clk = clocks.stop() if clocks else None
    attn_n, attn_ms = timer.summary().get("svr2_window_attn_fwd", (0, 0.0))

    # ---- timed region 2: host buffers through the public call (e2e)
    for _ in range(2):
        step_host()
    ms_e2e = timed(step_host, args.steps) / args.steps
    e2e_value = world * frames / (ms_e2e * 1e-3)

    # ---- per-kernel pass (outside the headline regions): every C-ABI kernel bracketed by events
    timer_all = ops.KernelTimer()
    ops.set_timer(timer_all)
    torch.cuda.synchronize()
    for _ in range(2):
        step_eager()
    torch.cuda.synchronize()
    ops.set_timer(None)
    per_kernel = timer_all.summary()
    per_launch = timer_all.durations()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    pk, pk_src = peaks()
    attn_flops = attention_flops_per_forward(svr2, model, tok_grid, layers, heads)      # per forward
    attn_ms_per_launch = attn_ms / max(attn_n, 1)
    achieved_tf = (attn_flops / layers) / (attn_ms_per_launch * 1e-3) / 1e12 if attn_n else 0.0
    plans = [svr2.get_plan(tok_grid, (4, 3, 3), sh, TXT_LEN) for sh in (False, True)]
    short = all(pl.packed_rows <= 576 * pl.n_windows for pl in plans)          # the library's dispatch rule (window_attn.cu)
    attn_kernel = ("generation 8, window_attn_tc6_kernel: CTA pairs, tcgen05.mma.cta_group::2 + TMA + TMEM" if short else
                   "generation 4, window_attn_tc2_kernel: TMA + tcgen05 + TMEM")
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(args.workload, {}).get("svr2_window_attn_fwd")
    except Exception:
        pass
    roofline = {
        "kernel": "svr2_window_attn_fwd (" + attn_kernel + ")", "bound": "tensor",
        "achieved": achieved_tf, "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s",
        "frac": achieved_tf / pk["bf16_tflops_sustained"], "frac_of_burst_peak": achieved_tf / pk["bf16_tflops"],
        "peak_source": pk_src + ", sustained bf16 figure (kernel timed inside a long step)",
        "algorithmic_flops_per_launch": attn_flops / layers, "avg_launch_ms": attn_ms_per_launch, "launches_timed": attn_n,
        "share_of_step": attn_ms / ms_eager_total if ms_eager_total else None, "traffic": traffic,
        "timing": "CUDA events around every attention launch over K steps launched kernel by kernel (region 1b); "
                  "the headline region replays the same kernels from a CUDA graph",
    }
    # HBM-bound kernels: algorithmic bytes per launch, bf16
    qkvb = 3 * heads * 128 * 2
    P_avg = L + TXT_LEN * 0.5 * (svr2.get_plan(tok_grid, (4, 3, 3), False, TXT_LEN).n_windows + svr2.get_plan(tok_grid, (4, 3, 3), True, TXT_LEN).n_windows)
    alg_bytes = {
        "svr2_qkv_pack_fwd": (L + TXT_LEN) * qkvb + P_avg * qkvb,
    }
    kernels = {}
    for name, (n, ms) in sorted(per_kernel.items()):
        rec = {"launches_per_step": n / 2, "ms_per_step": ms / 2}
        b = alg_bytes.get(name)
        if b:
            gbs = b * n / (ms * 1e-3) / 1e9
            rec.update(achieved_gbs=gbs, frac_of_hbm_peak=gbs / pk["hbm_gbs"])
        kernels[name] = rec
    # Row-wise kernels are launched on the video rows ([L,D], or the joined [L+Lt,D] of a shared-weight layer) and on the
    # few text rows; only the former are bandwidth-sized.  HBM fraction = algorithmic bytes of a video-row launch / the
    # mean duration of those launches (the ones lasting at least a third of the family's longest).
    F = {"3b": 6912, "7b": 4 * D}[model]
    for name, per_row in (("svr2_modnorm_fwd", 4 * D), ("svr2_gate_residual_fwd", 6 * D),
                          ("svr2_gate_residual_modnorm_fwd", 8 * D), ("svr2_swiglu_fwd", 6 * F)):
        if name in kernels and per_launch.get(name):
            d = per_launch[name]
            big = [x for x in d if x >= max(d) / 3]
            us = 1e3 * sum(big) / len(big)
            gbs = per_row * L / (us * 1e-6) / 1e9
            kernels[name].update(video_row_launches_per_step=len(big) / 2, us_per_video_row_launch=us,
                                 achieved_gbs=gbs, frac_of_hbm_peak=gbs / pk["hbm_gbs"],
                                 bytes_per_launch=per_row * L)

    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        sec, threads = cpu_block_seconds(model, tok_grid, steps=2, warmup=1)
        cpu_baseline = {"value": frames / (sec * layers), "unit": "frames/s", "cores": threads, "kind": "port",
                        "sample": f"oracle (reference fp32 PyTorch path restated) — 2 timed runs of ONE block of {layers} at the same token grid, "
                                  f"{sec:.2f} s/block, extrapolated x{layers}"}

    line = {
        "metric": "DiT frames/sec", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": workload_config(args.workload, world),
        "roofline": roofline, "cpu_baseline": cpu_baseline,
        "e2e": {"value": e2e_value, "unit": "frames/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": vid_host.numel() * 2 + txt_host.numel() * 2, "d2h_bytes_per_step": out_host.numel() * 2},
        "gpu_launches": launches, "clocks": clk, "kernels": kernels,
        "attention_tflops": achieved_tf,
        "launch_mode": ("cuda-graph replay of the forward (%d library kernels per replay)" % gf.launches_per_replay) if use_graph else "kernel by kernel",
        "ms_per_step_kernel_by_kernel": ms_eager_total / args.steps,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="c1", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch the forward kernel by kernel instead of replaying its CUDA graph")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
