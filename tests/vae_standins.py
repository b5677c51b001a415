"""Stand-ins for the VAE modules that cannot be instantiated in this image (their base classes live in `diffusers`), composed from
the reference's real, importable parts in the order of the reference's forward methods.  Test and benchmark infrastructure."""
import torch
from einops import rearrange
from torch import nn


def reference_parts():
    from oracle import refstage
    refstage.activate()
    from src.models.video_vae_v3.modules.causal_inflation_lib import causal_norm_wrapper, init_causal_conv3d, remove_head
    from src.models.video_vae_v3.modules.types import MemoryState
    # the reference's own function, also after install_vae_blocks() has rebound the module global (the reference arm of a
    # comparison must not run this library's GroupNorm)
    return init_causal_conv3d, getattr(causal_norm_wrapper, "_svr2_original", causal_norm_wrapper), remove_head, MemoryState


def output_norm(module, sample):
    """`causal_norm_wrapper(self.conv_norm_out, sample)` as attn_video_vae.py calls it — through the module global, which
    install_vae_blocks() rebinds.  `module.use_rebound_norm` (set on the arm that had install_vae_blocks applied) selects the
    rebound function; the reference arm keeps the original even when both arms live in one process."""
    import src.models.video_vae_v3.modules.causal_inflation_lib as lib
    fn = lib.causal_norm_wrapper
    if not getattr(module, "use_rebound_norm", False):
        fn = getattr(fn, "_svr2_original", fn)
    return fn(module.conv_norm_out, sample)


class ResnetBlock3D(nn.Module):
    """ResnetBlock3D.forward (attn_video_vae.py:311-361) as the VAE configures it."""

    def __init__(self, ci, co, full=True):
        super().__init__()
        init_causal_conv3d, wrap, _, _ = reference_parts()
        self.wrap = wrap
        self.norm1 = nn.GroupNorm(32, ci, eps=1e-6)
        self.norm2 = nn.GroupNorm(32, co, eps=1e-6)
        k, p = (3, 1) if full else ((1, 3, 3), (0, 1, 1))
        self.conv1 = init_causal_conv3d(ci, co, kernel_size=k, stride=1, padding=p, inflation_mode="tail")
        self.conv2 = init_causal_conv3d(co, co, kernel_size=3, stride=1, padding=1, inflation_mode="tail")
        self.conv_shortcut = init_causal_conv3d(ci, co, kernel_size=1, stride=1, padding=0, inflation_mode="tail") if ci != co else None
        self.nonlinearity = nn.SiLU()
        self.time_emb_proj = self.upsample = self.downsample = None
        self.output_scale_factor = 1.0
        for n in (self.norm1, self.norm2):
            nn.init.normal_(n.weight, 1.0, 0.2)
            nn.init.normal_(n.bias, 0.0, 0.2)

    def forward(self, x, temb=None, memory_state=None):
        h = self.nonlinearity(self.wrap(self.norm1, x))
        h = self.conv1(h, memory_state=memory_state)
        h = self.nonlinearity(self.wrap(self.norm2, h))
        h = self.conv2(h, memory_state=memory_state)
        if self.conv_shortcut is not None:
            x = self.conv_shortcut(x, memory_state=memory_state)
        return (x + h) / self.output_scale_factor


class Upsample3D(nn.Module):
    """Upsample3D.forward (attn_video_vae.py:110-175): 1 x 1 x 1 expansion, pixel (and time) shuffle, the duplicated first frame
    removed on the first temporal slice, causal 3 x 3 x 3 convolution; with `slicing` the convolution gets two temporal pieces."""

    def __init__(self, channels, temporal_up=False, spatial_up=True, slicing=False):
        super().__init__()
        init_causal_conv3d, _, remove_head, MemoryState = reference_parts()
        self.remove_head, self.MemoryState = remove_head, MemoryState
        self.temporal_ratio, self.spatial_ratio = (2 if temporal_up else 1), (2 if spatial_up else 1)
        self.temporal_up, self.slicing = temporal_up, slicing
        self.channels, self.name, self.norm, self.use_conv, self.use_conv_transpose, self.interpolate = channels, "conv", None, True, False, False
        r = self.spatial_ratio ** 2 * self.temporal_ratio
        self.upscale_conv = nn.Conv3d(channels, channels * r, kernel_size=1, padding=0)
        self.conv = init_causal_conv3d(channels, channels, 3, padding=1, inflation_mode="tail")

    def forward(self, hidden_states, memory_state=None):
        if self.slicing:
            split = hidden_states.size(2) // 2
            hidden_states = list(hidden_states.split([split, hidden_states.size(2) - split], dim=2))
        else:
            hidden_states = [hidden_states]
        for i in range(len(hidden_states)):
            t = self.upscale_conv(hidden_states[i])
            hidden_states[i] = rearrange(t, "b (x y z c) f h w -> b c (f z) (h x) (w y)", x=self.spatial_ratio, y=self.spatial_ratio,
                                         z=self.temporal_ratio)
        if self.temporal_up and memory_state != self.MemoryState.ACTIVE:
            hidden_states[0] = self.remove_head(hidden_states[0])
        if not self.slicing:
            hidden_states = hidden_states[0]
        hidden_states = self.conv(hidden_states, memory_state=memory_state)
        return hidden_states if not self.slicing else torch.cat(hidden_states, dim=2)


class DecoderTop(nn.Module):
    """The upper half of Decoder3D.forward (attn_video_vae.py:1030-1041) for the product VAE (block_out_channels 128, 256, 512, 512;
    three residual blocks per up block): up block 2 (512 -> 256 at 540 x 960 for a 1080p clip, spatial up-sampler), up block 3
    (256 -> 128 at full resolution), conv_norm_out, SiLU, conv_out."""

    def __init__(self, c_in=512, c_mid=256, c_top=128, slicing=False):
        super().__init__()
        init_causal_conv3d, _, _, _ = reference_parts()
        self.up2 = nn.ModuleList([ResnetBlock3D(c_in, c_mid), ResnetBlock3D(c_mid, c_mid), ResnetBlock3D(c_mid, c_mid)])
        self.up2_sampler = Upsample3D(c_mid, temporal_up=False, spatial_up=True, slicing=slicing)
        self.up3 = nn.ModuleList([ResnetBlock3D(c_mid, c_top), ResnetBlock3D(c_top, c_top), ResnetBlock3D(c_top, c_top)])
        self.temporal = nn.Identity()
        self.conv_norm_out = nn.GroupNorm(32, c_top, eps=1e-6)
        self.conv_act = nn.SiLU()
        self.conv_out = init_causal_conv3d(c_top, 3, 3, padding=1, inflation_mode="tail")

    def forward(self, sample, memory_state=None):
        for r in self.up2:
            sample = self.temporal(r(sample, temb=None, memory_state=memory_state))
        sample = self.up2_sampler(sample, memory_state=memory_state)
        for r in self.up3:
            sample = self.temporal(r(sample, temb=None, memory_state=memory_state))
        sample = self.conv_act(output_norm(self, sample))
        return self.conv_out(sample, memory_state=memory_state)


class Decoder3D(nn.Module):
    """Decoder3D.forward (attn_video_vae.py:1000-1041) for the product VAE (s8_c16_t4_inflation_sd3.yaml: latent 16 channels,
    block_out_channels 128 / 256 / 512 / 512, 3 residual blocks per up block, temporal up-sampling in up blocks 0 and 1, spatial in 0-2,
    time_receptive_field "full") WITHOUT the mid block's attention layer (diffusers' `Attention`, not in this image)."""

    def __init__(self, latent=16, ch=(512, 512, 256, 128), attention=False):
        super().__init__()
        init_causal_conv3d, _, _, _ = reference_parts()
        self.conv_in = init_causal_conv3d(latent, ch[0], kernel_size=3, stride=1, padding=1, inflation_mode="tail")
        self.mid_block = UNetMidBlock3D(ch[0], attention)
        self.up_resnets = nn.ModuleList()
        self.up_samplers = nn.ModuleList()
        prev = ch[0]
        for i, c in enumerate(ch):
            self.up_resnets.append(nn.ModuleList([ResnetBlock3D(prev if j == 0 else c, c) for j in range(3)]))
            self.up_samplers.append(Upsample3D(c, temporal_up=i < 2, spatial_up=True) if i < len(ch) - 1 else nn.Identity())
            prev = c
        self.conv_norm_out = nn.GroupNorm(32, ch[-1], eps=1e-6)
        self.conv_act = nn.SiLU()
        self.conv_out = init_causal_conv3d(ch[-1], 3, 3, padding=1, inflation_mode="tail")

    def forward(self, sample, memory_state=None):
        sample = self.conv_in(sample, memory_state=memory_state)
        sample = self.mid_block(sample, None, memory_state=memory_state)
        for resnets, up in zip(self.up_resnets, self.up_samplers):
            for r in resnets:
                sample = r(sample, temb=None, memory_state=memory_state)
            if not isinstance(up, nn.Identity):
                sample = up(sample, memory_state=memory_state)
        sample = self.conv_act(output_norm(self, sample))
        return self.conv_out(sample, memory_state=memory_state)


class Downsample3D(nn.Module):
    """Downsample3D.forward (attn_video_vae.py:228-246) as the encoder builds it (`name="op"`, padding 0, convolution)."""

    def __init__(self, channels, temporal_down=False):
        super().__init__()
        init_causal_conv3d, _, _, _ = reference_parts()
        self.channels, self.out_channels, self.use_conv, self.padding, self.norm = channels, channels, True, 0, None
        self.spatial_down, self.temporal_down = True, temporal_down
        tk, tr = (3, 2) if temporal_down else (1, 1)
        self.conv = init_causal_conv3d(channels, channels, kernel_size=(tk, 3, 3), stride=(tr, 2, 2),
                                       padding=(1 if temporal_down else 0, 0, 0), inflation_mode="tail")

    def forward(self, hidden_states, memory_state=None):
        hidden_states = torch.nn.functional.pad(hidden_states, (0, 1, 0, 1), mode="constant", value=0)
        return self.conv(hidden_states, memory_state=memory_state)


class Encoder3D(nn.Module):
    """Encoder3D.forward (attn_video_vae.py:700-870) for the product VAE (3 -> 128 / 256 / 512 / 512, two residual blocks per down block,
    spatial down-sampling after blocks 0-2, temporal as well after blocks 1-2, double-z output) WITHOUT the mid block's attention layer."""

    def __init__(self, ch=(128, 256, 512, 512), latent=16, attention=False):
        super().__init__()
        init_causal_conv3d, _, _, _ = reference_parts()
        self.conv_in = init_causal_conv3d(3, ch[0], kernel_size=3, stride=1, padding=1, inflation_mode="tail")
        self.down_resnets = nn.ModuleList()
        self.down_samplers = nn.ModuleList()
        prev = ch[0]
        for i, c in enumerate(ch):
            self.down_resnets.append(nn.ModuleList([ResnetBlock3D(prev if j == 0 else c, c) for j in range(2)]))
            self.down_samplers.append(Downsample3D(c, temporal_down=i >= 1) if i < len(ch) - 1 else nn.Identity())
            prev = c
        self.mid_block = UNetMidBlock3D(ch[-1], attention)
        self.conv_norm_out = nn.GroupNorm(32, ch[-1], eps=1e-6)
        self.conv_act = nn.SiLU()
        self.conv_out = init_causal_conv3d(ch[-1], 2 * latent, 3, padding=1, inflation_mode="tail")

    def forward(self, sample, memory_state=None):
        sample = self.conv_in(sample, memory_state=memory_state)
        for resnets, down in zip(self.down_resnets, self.down_samplers):
            for r in resnets:
                sample = r(sample, temb=None, memory_state=memory_state)
            if not isinstance(down, nn.Identity):
                sample = down(sample, memory_state=memory_state)
        sample = self.mid_block(sample, None, memory_state=memory_state)
        sample = self.conv_act(output_norm(self, sample))
        return self.conv_out(sample, memory_state=memory_state)


class Attention(nn.Module):
    """The mid block's attention layer as UNetMidBlock3D builds it (attn_video_vae.py:600-640: diffusers `Attention(in_channels,
    heads=in_channels // attention_head_dim, dim_head=attention_head_dim, norm_num_groups=32, eps=1e-6, residual_connection=True,
    bias=True, upcast_softmax=True, rescale_output_factor=1)`) with diffusers' attribute names and the sequence of its
    AttnProcessor2_0.  diffusers is not in this image: this is a restatement, not the reference's code."""

    def __init__(self, channels, groups=32, eps=1e-6):
        super().__init__()
        self.heads, self.rescale_output_factor, self.residual_connection = 1, 1.0, True
        self.norm_cross = self.spatial_norm = self.added_kv_proj_dim = None
        self.group_norm = nn.GroupNorm(groups, channels, eps=eps, affine=True)
        self.to_q, self.to_k, self.to_v = (nn.Linear(channels, channels, bias=True) for _ in range(3))
        self.to_out = nn.ModuleList([nn.Linear(channels, channels, bias=True), nn.Dropout(0.0)])

    def forward(self, hidden_states, temb=None):
        residual = hidden_states
        b, c, h, w = hidden_states.shape
        x = hidden_states.reshape(b, c, h * w).transpose(1, 2)
        x = self.group_norm(x.transpose(1, 2)).transpose(1, 2)
        q, k, v = self.to_q(x), self.to_k(x), self.to_v(x)
        q, k, v = (t.view(b, -1, self.heads, c // self.heads).transpose(1, 2) for t in (q, k, v))
        x = torch.nn.functional.scaled_dot_product_attention(q, k, v, attn_mask=None, dropout_p=0.0, is_causal=False)
        x = x.transpose(1, 2).reshape(b, -1, c).to(q.dtype)
        x = self.to_out[1](self.to_out[0](x))
        x = x.transpose(-1, -2).reshape(b, c, h, w)
        if self.residual_connection:
            x = x + residual
        return x / self.rescale_output_factor


class UNetMidBlock3D(nn.Module):
    """UNetMidBlock3D.forward (attn_video_vae.py:656-668): residual block, attention frame by frame, residual block."""

    def __init__(self, channels, attention=True):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock3D(channels, channels), ResnetBlock3D(channels, channels)])
        self.attentions = nn.ModuleList([Attention(channels) if attention else None])

    def forward(self, hidden_states, temb=None, memory_state=None):
        video_length = hidden_states.size(2)
        hidden_states = self.resnets[0](hidden_states, temb, memory_state=memory_state)
        for attn, resnet in zip(self.attentions, self.resnets[1:]):
            if attn is not None:
                hidden_states = rearrange(hidden_states, "b c f h w -> (b f) c h w")
                hidden_states = attn(hidden_states, temb=temb)
                hidden_states = rearrange(hidden_states, "(b f) c h w -> b c f h w", f=video_length)
            hidden_states = resnet(hidden_states, temb, memory_state=memory_state)
        return hidden_states
