"""Generate the golden vectors under tests/golden/ by running the REAL reference (disk02/ComfyUI-SeedVR2, mounted
read-only at /root/reference) in the build container.  Run once:  python tests/golden/make_golden.py

The reference has no tests or fixtures of its own, so these files ARE the pin for the oracle:
  geometry.json     window bounds / window_shape / cu_seqlens / index arrays (or their sha256) from the reference's
                    window.py + na.window_idx for every grid in cases.GEOMETRY_GRIDS, both window methods
  stack_<case>.pt   output of the reference NaDiT.forward (whole model, narrowed width/depth; fp32 and bf16 CPU) on
                    cases.stack_inputs with cases.stack_params; also asserts that the product's NaDiTStack exposes the
                    reference NaDiT's state-dict names and shapes
  block_<case>.pt   outputs of the reference NaMMSRTransformerBlock (fp32 CPU, and bf16 CPU) on cases.block_inputs
                    with cases.block_params loaded through load_state_dict (which also pins the state-dict contract)
  sampler_<case>.pt condition, DiT input and one-step endpoint from the real VideoDiffusionInfer + EulerSampler with a
                    stand-in DiT (cases.SAMPLER_CASES)
Needs: /root/reference, oracle/shim (third-party restatements).  Never runs on the GPU box.
"""
import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[0:0] = [os.path.join(ROOT, "oracle", "shim"), "/root/reference", ROOT, HERE]

import torch  # noqa: E402

import cases  # noqa: E402


def sha(a) -> str:
    return hashlib.sha256(a.contiguous().numpy().tobytes()).hexdigest()


def gen_geometry():
    from src.models.dit_3b import na
    from src.models.dit_3b.window import get_window_op
    out = {}
    for (T, H, W) in cases.GEOMETRY_GRIDS:
        for method in ("720pwin_by_size_bysize", "720pswin_by_size_bysize"):
            op = get_window_op(method)
            slices = op((T, H, W), (4, 3, 3))
            bounds = [[s[0].start, s[0].stop, s[1].start, s[1].stop, s[2].start, s[2].stop] for s in slices]
            vid_shape = torch.tensor([[T, H, W]], dtype=torch.long)

            def make_window(x, op=op):
                t, h, w, _ = x.shape
                return [x[st, sh, sw] for (st, sh, sw) in op((t, h, w), (4, 3, 3))]

            part, rev, wshape, wcount = na.window_idx(vid_shape, make_window)
            L = T * H * W
            ar = torch.arange(L).unsqueeze(-1)
            tgt = part(ar).squeeze(-1)
            src = rev(ar).squeeze(-1)
            lens = wshape.prod(-1) + cases.TXT_LEN
            cu = torch.nn.functional.pad(lens.cumsum(0), (1, 0)).int()
            rec = {
                "bounds": bounds,
                "window_shape": wshape.tolist(),
                "window_count": int(wcount[0]),
                "cu_seqlens_txt58": cu.tolist(),
                "tgt_sha256": sha(tgt), "src_sha256": sha(src),
            }
            if L <= cases.GEOMETRY_FULL_LIMIT:
                rec["tgt_idx"] = tgt.tolist()
                rec["src_idx"] = src.tolist()
            out[f"{T}x{H}x{W}:{'swin' if 'swin' in method else 'win'}"] = rec
    with open(os.path.join(HERE, "geometry.json"), "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("geometry.json:", len(out), "entries")


def build_ref_block(c: cases.BlockCase):
    method = "720pswin_by_size_bysize" if c.shifted else "720pwin_by_size_bysize"
    if c.model == "3b":
        from src.models.dit_3b.modulation import get_ada_layer
        from src.models.dit_3b.nablocks import get_nablock
        from src.models.dit_3b.normalization import get_norm_layer
        blk = get_nablock("mmdit_sr")(
            vid_dim=c.dim, txt_dim=c.dim, emb_dim=6 * c.dim, heads=c.heads, head_dim=c.head_dim, expand_ratio=4,
            norm=get_norm_layer("fusedrms"), norm_eps=1e-5, ada=get_ada_layer("single"), qk_bias=False,
            qk_norm=get_norm_layer("fusedrms"), shared_weights=c.shared, mlp_type="swiglu", window=c.window,
            window_method=method, msa_type=None, mca_type=None, rope_type="mmrope3d", rope_dim=128,
            is_last_layer=c.is_last, attention_mode="sdpa")
    else:
        from src.models.dit_7b.modulation import get_ada_layer
        from src.models.dit_7b.nablocks import get_nablock
        from src.models.dit_7b.normalization import get_norm_layer
        blk = get_nablock("mmdit_sr")(
            vid_dim=c.dim, txt_dim=c.dim, emb_dim=6 * c.dim, heads=c.heads, head_dim=c.head_dim, expand_ratio=4,
            norm=get_norm_layer("fusedrms"), norm_eps=1e-5, ada=get_ada_layer("single"), qk_bias=False,
            qk_rope=True, qk_norm=get_norm_layer("fusedrms"), shared_qkv=False, shared_mlp=False,
            mlp_type="normal", window=c.window, window_method=method, temporal_window_size=None,
            temporal_shifted=None, attention_mode="sdpa")
    return blk.eval()


def gen_blocks():
    from src.common.cache import Cache
    from oracle import block as oblock
    for c in cases.BLOCK_CASES:
        blk = build_ref_block(c)
        sd = blk.state_dict()
        want = cases.block_param_shapes(c)
        got = {k: tuple(v.shape) for k, v in sd.items()}
        assert got == want, f"state-dict contract mismatch for {c.name}:\n{sorted(set(got.items()) ^ set(want.items()))}"
        ref_freqs = sd["attn.rope.rope.freqs"].clone()
        params = cases.block_params(c)
        assert torch.equal(ref_freqs, params["attn.rope.rope.freqs"]), "rope freqs formula differs from reference"
        blk.load_state_dict(params, strict=True)
        vid, txt, emb = cases.block_inputs(c)
        T, H, W = c.grid
        vid_shape = torch.tensor([[T, H, W]]); txt_shape = torch.tensor([[c.txt_len]])
        captured = {}
        hook_mod = blk.attn.proj_out
        h = hook_mod.register_forward_pre_hook(lambda m, a: captured.update(attn_vid=a[0].clone(), attn_txt=a[1].clone()))
        with torch.no_grad():
            cache = Cache(disable=False)
            v32, t32, _, _ = blk(vid=vid.clone(), txt=txt.clone(), vid_shape=vid_shape, txt_shape=txt_shape,
                                 emb=emb.clone(), cache=cache)
            a_v, a_t = captured["attn_vid"], captured["attn_txt"]
            blk16 = blk.bfloat16()
            blk16.attn.rope.rope.freqs.data = ref_freqs.clone()     # checkpoint buffer stays fp32 in the product too
            cache = Cache(disable=False)
            v16, t16, _, _ = blk16(vid=vid.bfloat16(), txt=txt.bfloat16(), vid_shape=vid_shape, txt_shape=txt_shape,
                                   emb=emb.bfloat16(), cache=cache)
            a_v16, a_t16 = captured["attn_vid"], captured["attn_txt"]
        h.remove()
        # oracle vs reference, same container
        cfg = oblock.BlockCfg(model=c.model, dim=c.dim, heads=c.heads, head_dim=c.head_dim, shared=c.shared,
                              is_last=c.is_last, shifted=c.shifted, window=c.window)
        with torch.no_grad():
            ov, ot = oblock.block_forward(cfg, params, vid, txt, c.grid, emb)
            p16 = {k: (v if k.endswith("rope.freqs") else v.bfloat16()) for k, v in params.items()}
            ov16, ot16 = oblock.block_forward(cfg, p16, vid.bfloat16(), txt.bfloat16(), c.grid, emb.bfloat16())
        print(f"{c.name}: oracle-vs-ref fp32 max|d| vid {float((ov - v32).abs().max()):.3e} txt {float((ot - t32).abs().max()):.3e}"
              f" | bf16 vid {float((ov16.float() - v16.float()).abs().max()):.3e} txt {float((ot16.float() - t16.float()).abs().max()):.3e}"
              f" | ref bf16-vs-fp32 vid {float((v16.float() - v32).abs().max()):.3e} txt {float((t16.float() - t32).abs().max()):.3e}"
              f" | |out| max {float(v32.abs().max()):.2f}")
        torch.save({
            "case": c.name,
            "in_digest": {"vid": cases.tensor_digest(vid), "txt": cases.tensor_digest(txt), "emb": cases.tensor_digest(emb)},
            "vid_out_fp32": v32, "txt_out_fp32": t32,
            "vid_out_bf16": v16, "txt_out_bf16": t16,
            "attn_txt_fp32": a_t,
            "attn_vid_bf16": a_v16, "attn_txt_bf16": a_t16,
        }, os.path.join(HERE, f"block_{c.name}.pt"))


def gen_stacks():
    """Whole NaDiT.forward of the real reference (dit_3b/nadit.py:171-248, dit_7b/nadit.py:152-191)."""
    import seedvr2_b200
    from oracle import nadit as onadit
    for c in cases.STACK_CASES:
        if c.model == "3b":
            from src.models.dit_3b.nadit import NaDiT
        else:
            from src.models.dit_7b.nadit import NaDiT
        ref = NaDiT(**cases.stack_kwargs(c)).eval()
        shapes = {k: tuple(v.shape) for k, v in ref.state_dict().items()}
        ours = seedvr2_b200.stack.NaDiTStack(c.model, num_layers=c.num_layers, vid_dim=c.dim, heads=c.heads,
                                             mm_layers=c.mm_layers, txt_in_dim=c.txt_in_dim)
        mine = {k: tuple(v.shape) for k, v in ours.state_dict().items()}
        assert mine == shapes, f"NaDiTStack state dict differs from the reference NaDiT for {c.name}:\n" \
                               f"{sorted(set(mine.items()) ^ set(shapes.items()))}"
        params = cases.stack_params(c, shapes)
        ref.load_state_dict(params, strict=True)
        vid, txt = cases.stack_inputs(c)
        vid_shape = torch.tensor([list(c.latent_grid)]); txt_shape = torch.tensor([[c.txt_len]])
        ts = torch.tensor([c.timestep])
        kw = dict(disable_cache=False) if c.model == "3b" else {}      # 7B keeps its default (disable_cache=True, quirk Q2)
        with torch.no_grad():
            o32 = ref(vid.clone(), txt.clone(), vid_shape.clone(), txt_shape.clone(), ts, **kw).vid_sample
            ref16 = ref.bfloat16()
            for m in ref16.modules():                    # checkpoint rope buffers stay fp32 in the product too
                if hasattr(m, "freqs") and isinstance(getattr(m, "freqs"), torch.Tensor):
                    m.freqs.data = cases.rope_freqs(c.model)
            o16 = ref16(vid.bfloat16(), txt.bfloat16(), vid_shape.clone(), txt_shape.clone(), ts, **kw).vid_sample
            okw = dict(dim=c.dim, heads=c.heads, num_layers=c.num_layers, mm_layers=c.mm_layers)
            oo32 = onadit.nadit_forward(c.model, params, vid, txt, c.latent_grid, c.timestep, **okw)
            p16 = {k: (v if k.endswith("rope.freqs") else v.bfloat16()) for k, v in params.items()}
            oo16 = onadit.nadit_forward(c.model, p16, vid.bfloat16(), txt.bfloat16(), c.latent_grid, c.timestep, **okw)
        print(f"{c.name}: out {tuple(o32.shape)} |out| max {float(o32.abs().max()):.3f} rms {float(o32.pow(2).mean().sqrt()):.3f}"
              f" | oracle-vs-ref fp32 {float((oo32 - o32).abs().max()):.3e} bf16 {float((oo16.float() - o16.float()).abs().max()):.3e}"
              f" | ref bf16-vs-fp32 max {float((o16.float() - o32).abs().max()):.3e} PSNR {cases.psnr(o16.float(), o32):.1f} dB")
        torch.save({"case": c.name, "in_digest": {"vid": cases.tensor_digest(vid), "txt": cases.tensor_digest(txt)},
                    "out_fp32": o32, "out_bf16": o16}, os.path.join(HERE, f"stack_{c.name}.pt"))


def gen_sampler():
    """The one-step SR call around the DiT forward, run through the REAL VideoDiffusionInfer / EulerSampler /
    LinearInterpolationSchedule / UniformTrailingSamplingTimesteps with a stand-in DiT (cases.standin_dit)."""
    from types import SimpleNamespace

    from src.common.diffusion.samplers.euler import EulerSampler
    from src.common.diffusion.schedules.lerp import LinearInterpolationSchedule
    from src.common.diffusion.timesteps.sampling.trailing import UniformTrailingSamplingTimesteps
    from src.common.diffusion.types import PredictionType
    from src.core.infer import VideoDiffusionInfer

    for c in cases.SAMPLER_CASES:
        dt = torch.bfloat16 if c.dtype == "bf16" else torch.float32
        latent, base_noise, aug_noise, mix = cases.sampler_inputs(c)
        runner = object.__new__(VideoDiffusionInfer)
        runner.config = SimpleNamespace(
            diffusion=SimpleNamespace(timesteps={"transform": True}, cfg=SimpleNamespace(scale=1.0, rescale=0.0, get=lambda k, d=None: d)),
            vae=SimpleNamespace(model={}))
        runner.schedule = LinearInterpolationSchedule(T=1000.0)                      # configs_3b/main.yaml:66-68
        runner.sampling_timesteps = UniformTrailingSamplingTimesteps(T=1000.0, steps=1, shift=1.0, device="cpu", dtype=dt)
        runner.sampler = EulerSampler(schedule=runner.schedule, timesteps=runner.sampling_timesteps,
                                      prediction_type=PredictionType.v_lerp)
        runner.sampler.get_progress_bar = lambda: SimpleNamespace(update=lambda: None)
        seen = {}
        fake = cases.standin_dit(mix)

        def dit(vid, txt, vid_shape, txt_shape, timestep):
            seen["vid"], seen["timestep"], seen["vid_shape"] = vid.clone(), timestep.clone(), vid_shape.clone()
            return SimpleNamespace(vid_sample=fake(vid, timestep))
        runner.dit = dit

        # generation_phases.py:685-701
        def _add_noise(x, aug):
            if c.latent_noise_scale == 0.0:
                return x
            t = torch.tensor([1000.0], dtype=dt) * c.latent_noise_scale
            shape = torch.tensor(x.shape[1:])[None]
            t = runner.timestep_transform(t, shape)
            return runner.schedule.forward(x, aug, t)
        blur = _add_noise(latent, aug_noise)
        condition = runner.get_condition(base_noise, task="sr", latent_blur=blur)
        txt = torch.zeros(58, 8, dtype=dt)
        out = runner.inference(noises=[base_noise], conditions=[condition], texts_pos=[txt], texts_neg=[txt])
        assert len(out) == 1 and out[0].shape == latent.shape
        torch.save({"latent_blur": blur, "condition": condition, "dit_input": seen["vid"], "timestep": seen["timestep"],
                    "vid_shape": seen["vid_shape"], "x0": out[0]}, os.path.join(HERE, f"sampler_{c.name}.pt"))
        print(f"sampler_{c.name}.pt: input {tuple(seen['vid'].shape)} {seen['vid'].dtype} timestep {seen['timestep'].tolist()} "
              f"x0 {out[0].dtype} blur dtype {blur.dtype}")


if __name__ == "__main__":
    torch.manual_seed(0)
    what = sys.argv[1:] or ["geometry", "blocks", "stacks", "sampler"]
    if "geometry" in what:
        gen_geometry()
    if "blocks" in what:
        gen_blocks()
    if "stacks" in what:
        gen_stacks()
    if "sampler" in what:
        gen_sampler()
