"""ORACLE — test infrastructure only (imported by tests/, __graft_entry__.smoke() and bench.py's CPU baseline).

Plain-torch fp32 CPU restatement of what the reference's hot loop calls as
``pipe.unet(x, t, encoder_hidden_states=emb).sample`` (Distribution/strategies/chunk_only.py:71,
fsdp_chunked_coherent.py:140, InferNet/neurons/miner.py:581).

The arithmetic lives in a third-party dependency that is NOT vendored in /root/reference and not installed in
this image: ``diffusers`` (InferNet/requirements.txt:10 ``diffusers>=0.19.0``, InferNet/setup.py:75
``>=0.21.0``, no lock file => effectively unpinned).  This file restates the published algorithm of
``diffusers.models.unet_3d_condition.UNet3DConditionModel`` (+ ``unet_3d_blocks``, ``transformer_temporal``,
``transformer_2d``, ``attention.BasicTransformerBlock``, ``resnet.{ResnetBlock2D,TemporalConvLayer,
Downsample2D,Upsample2D}``, ``embeddings.{Timesteps,TimestepEmbedding}``) for the Zeroscope v2 / ModelScope-T2V
configuration, with module names chosen so that ``state_dict()`` keys are identical to the diffusers checkpoint.

PARITY UNPINNED: the reference's own tests pin nothing numerically on this path and diffusers
cannot be imported here, so this restatement is anchored only on (a) the 1 411 2xx xxx parameter total that matches
Zeroscope's published 1.41 B, (b) the state_dict key set / shapes of the public checkpoint, (c) the tensor shapes
printed in Distribution/legacy/Latent Chunking/latent_chunking.ipynb cell 9.
"""
import math
from dataclasses import dataclass, field
from typing import Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F


@dataclass
class UNet3DConfig:
    in_channels: int = 4
    out_channels: int = 4
    block_out_channels: Tuple[int, ...] = (320, 640, 1280, 1280)
    layers_per_block: int = 2
    cross_attention_dim: int = 1024
    attention_head_dim: int = 64  # diffusers passes this as the *number of heads divisor*: heads = C // 64
    norm_num_groups: int = 32
    norm_eps: float = 1e-5
    transformer_in_heads: int = 8
    sample_size: int = 32

    def to_dict(self):
        return dict(self.__dict__)


ZEROSCOPE = UNet3DConfig()
# narrow variant for fast tests: same topology, every width a multiple of 64
TINY = UNet3DConfig(block_out_channels=(64, 128, 256, 256), cross_attention_dim=128, transformer_in_heads=2)


# ---------------------------------------------------------------------------------------------------------------
# embeddings
# ---------------------------------------------------------------------------------------------------------------
def timestep_sinusoid(timesteps: torch.Tensor, dim: int) -> torch.Tensor:
    """diffusers get_timestep_embedding(flip_sin_to_cos=True, downscale_freq_shift=0, scale=1)."""
    half = dim // 2
    exponent = -math.log(10000) * torch.arange(half, dtype=torch.float32, device=timesteps.device) / half
    emb = timesteps[:, None].float() * torch.exp(exponent)[None, :]
    return torch.cat([torch.cos(emb), torch.sin(emb)], dim=-1)


class TimestepEmbedding(nn.Module):
    def __init__(self, in_channels, time_embed_dim):
        super().__init__()
        self.linear_1 = nn.Linear(in_channels, time_embed_dim)
        self.linear_2 = nn.Linear(time_embed_dim, time_embed_dim)

    def forward(self, x):
        return self.linear_2(F.silu(self.linear_1(x)))


# ---------------------------------------------------------------------------------------------------------------
# attention blocks
# ---------------------------------------------------------------------------------------------------------------
class Attention(nn.Module):
    # bench.py's CPU arm sets this: the same attention through F.scaled_dot_product_attention (what diffusers'
    # default AttnProcessor2_0 calls) — the materialised [b, heads, s, s] score tensor below is 54 GB at the
    # ZeroscopeXL 16-frame window. Tests keep the explicit softmax.
    use_sdpa = False

    def __init__(self, query_dim, heads, dim_head, cross_attention_dim=None):
        super().__init__()
        inner = heads * dim_head
        self.heads = heads
        self.scale = dim_head ** -0.5
        kv_dim = cross_attention_dim if cross_attention_dim is not None else query_dim
        self.to_q = nn.Linear(query_dim, inner, bias=False)
        self.to_k = nn.Linear(kv_dim, inner, bias=False)
        self.to_v = nn.Linear(kv_dim, inner, bias=False)
        self.to_out = nn.ModuleList([nn.Linear(inner, query_dim), nn.Dropout(0.0)])

    def forward(self, x, context=None):
        ctx = x if context is None else context
        b, s, _ = x.shape
        q = self.to_q(x).view(b, s, self.heads, -1).transpose(1, 2)
        k = self.to_k(ctx).view(b, ctx.shape[1], self.heads, -1).transpose(1, 2)
        v = self.to_v(ctx).view(b, ctx.shape[1], self.heads, -1).transpose(1, 2)
        if Attention.use_sdpa:
            o = F.scaled_dot_product_attention(q, k, v, scale=self.scale).transpose(1, 2).reshape(b, s, -1)
        else:
            w = torch.softmax((q @ k.transpose(-1, -2)) * self.scale, dim=-1)
            o = (w @ v).transpose(1, 2).reshape(b, s, -1)
        return self.to_out[0](o)


class GEGLU(nn.Module):
    def __init__(self, dim_in, dim_out):
        super().__init__()
        self.proj = nn.Linear(dim_in, dim_out * 2)

    def forward(self, x):
        h, gate = self.proj(x).chunk(2, dim=-1)
        return h * F.gelu(gate)  # erf gelu


class FeedForward(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.net = nn.ModuleList([GEGLU(dim, dim * 4), nn.Dropout(0.0), nn.Linear(dim * 4, dim)])

    def forward(self, x):
        for m in self.net:
            x = m(x)
        return x


class BasicTransformerBlock(nn.Module):
    def __init__(self, dim, heads, dim_head, cross_attention_dim=None, double_self_attention=False):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim)
        self.attn1 = Attention(dim, heads, dim_head)
        self.norm2 = nn.LayerNorm(dim)
        self.attn2 = Attention(dim, heads, dim_head,
                               cross_attention_dim=None if double_self_attention else cross_attention_dim)
        self.norm3 = nn.LayerNorm(dim)
        self.ff = FeedForward(dim)
        self.double_self_attention = double_self_attention

    def forward(self, x, context=None):
        x = self.attn1(self.norm1(x)) + x
        x = self.attn2(self.norm2(x), None if self.double_self_attention else context) + x
        x = self.ff(self.norm3(x)) + x
        return x


class Transformer2DModel(nn.Module):
    """Spatial transformer, use_linear_projection=True, one block."""

    def __init__(self, heads, dim_head, in_channels, cross_attention_dim, groups):
        super().__init__()
        inner = heads * dim_head
        self.norm = nn.GroupNorm(groups, in_channels, eps=1e-6, affine=True)
        self.proj_in = nn.Linear(in_channels, inner)
        self.transformer_blocks = nn.ModuleList(
            [BasicTransformerBlock(inner, heads, dim_head, cross_attention_dim=cross_attention_dim)])
        self.proj_out = nn.Linear(inner, in_channels)

    def forward(self, x, context):
        n, c, h, w = x.shape
        res = x
        x = self.norm(x)
        x = x.permute(0, 2, 3, 1).reshape(n, h * w, c)
        x = self.proj_in(x)
        for blk in self.transformer_blocks:
            x = blk(x, context)
        x = self.proj_out(x)
        x = x.reshape(n, h, w, c).permute(0, 3, 1, 2)
        return x + res


class TransformerTemporalModel(nn.Module):
    """Attention over the frame axis; double_self_attention=True, no positional encoding."""

    def __init__(self, heads, dim_head, in_channels, groups):
        super().__init__()
        inner = heads * dim_head
        self.norm = nn.GroupNorm(groups, in_channels, eps=1e-6, affine=True)
        self.proj_in = nn.Linear(in_channels, inner)
        self.transformer_blocks = nn.ModuleList(
            [BasicTransformerBlock(inner, heads, dim_head, double_self_attention=True)])
        self.proj_out = nn.Linear(inner, in_channels)

    def forward(self, x, num_frames):
        bf, c, h, w = x.shape
        b = bf // num_frames
        res = x
        x = x[None, :].reshape(b, num_frames, c, h, w).permute(0, 2, 1, 3, 4)  # b c f h w
        x = self.norm(x)
        x = x.permute(0, 3, 4, 2, 1).reshape(b * h * w, num_frames, c)
        x = self.proj_in(x)
        for blk in self.transformer_blocks:
            x = blk(x)
        x = self.proj_out(x)
        x = x[None, None, :].reshape(b, h, w, num_frames, c).permute(0, 3, 4, 1, 2).contiguous()
        x = x.reshape(bf, c, h, w)
        return x + res


# ---------------------------------------------------------------------------------------------------------------
# convolutional blocks
# ---------------------------------------------------------------------------------------------------------------
class ResnetBlock2D(nn.Module):
    def __init__(self, in_channels, out_channels, temb_channels, groups, eps):
        super().__init__()
        self.norm1 = nn.GroupNorm(groups, in_channels, eps=eps, affine=True)
        self.conv1 = nn.Conv2d(in_channels, out_channels, 3, padding=1)
        self.time_emb_proj = nn.Linear(temb_channels, out_channels)
        self.norm2 = nn.GroupNorm(groups, out_channels, eps=eps, affine=True)
        self.conv2 = nn.Conv2d(out_channels, out_channels, 3, padding=1)
        self.conv_shortcut = nn.Conv2d(in_channels, out_channels, 1) if in_channels != out_channels else None

    def forward(self, x, temb):
        h = self.conv1(F.silu(self.norm1(x)))
        h = h + self.time_emb_proj(F.silu(temb))[:, :, None, None]
        h = self.conv2(F.silu(self.norm2(h)))
        if self.conv_shortcut is not None:
            x = self.conv_shortcut(x)
        return x + h


class TemporalConvLayer(nn.Module):
    def __init__(self, dim, groups):
        super().__init__()
        self.conv1 = nn.Sequential(nn.GroupNorm(groups, dim), nn.SiLU(), nn.Conv3d(dim, dim, (3, 1, 1), padding=(1, 0, 0)))
        self.conv2 = nn.Sequential(nn.GroupNorm(groups, dim), nn.SiLU(), nn.Dropout(0.1),
                                   nn.Conv3d(dim, dim, (3, 1, 1), padding=(1, 0, 0)))
        self.conv3 = nn.Sequential(nn.GroupNorm(groups, dim), nn.SiLU(), nn.Dropout(0.1),
                                   nn.Conv3d(dim, dim, (3, 1, 1), padding=(1, 0, 0)))
        self.conv4 = nn.Sequential(nn.GroupNorm(groups, dim), nn.SiLU(), nn.Dropout(0.1),
                                   nn.Conv3d(dim, dim, (3, 1, 1), padding=(1, 0, 0)))
        nn.init.zeros_(self.conv4[-1].weight)  # upstream: identity at init (tests re-randomise)
        nn.init.zeros_(self.conv4[-1].bias)

    def forward(self, x, num_frames):
        x = x[None, :].reshape((-1, num_frames) + x.shape[1:]).permute(0, 2, 1, 3, 4)
        identity = x
        x = self.conv4(self.conv3(self.conv2(self.conv1(x))))
        x = identity + x
        return x.permute(0, 2, 1, 3, 4).reshape((x.shape[0] * x.shape[2], -1) + x.shape[3:])


class Downsample2D(nn.Module):
    def __init__(self, channels):
        super().__init__()
        self.conv = nn.Conv2d(channels, channels, 3, stride=2, padding=1)

    def forward(self, x):
        return self.conv(x)


class Upsample2D(nn.Module):
    def __init__(self, channels):
        super().__init__()
        self.conv = nn.Conv2d(channels, channels, 3, padding=1)

    def forward(self, x):
        return self.conv(F.interpolate(x, scale_factor=2.0, mode="nearest"))


class Block3D(nn.Module):
    """CrossAttnDownBlock3D / DownBlock3D / UpBlock3D / CrossAttnUpBlock3D / UNetMidBlock3DCrossAttn share this
    container; only the module lists that exist upstream are registered so that state_dict keys match."""

    def __init__(self, cfg: UNet3DConfig, resnet_io, out_channels, temb, attn: bool, n_attn: int,
                 down: bool = False, up: bool = False):
        super().__init__()
        g, eps = cfg.norm_num_groups, cfg.norm_eps
        self.resnets = nn.ModuleList([ResnetBlock2D(i, o, temb, g, eps) for (i, o) in resnet_io])
        self.temp_convs = nn.ModuleList([TemporalConvLayer(o, g) for (_, o) in resnet_io])
        if attn:
            heads = out_channels // cfg.attention_head_dim
            self.attentions = nn.ModuleList(
                [Transformer2DModel(heads, cfg.attention_head_dim, out_channels, cfg.cross_attention_dim, g)
                 for _ in range(n_attn)])
            self.temp_attentions = nn.ModuleList(
                [TransformerTemporalModel(heads, cfg.attention_head_dim, out_channels, g) for _ in range(n_attn)])
        self.has_attn = attn
        if down:
            self.downsamplers = nn.ModuleList([Downsample2D(out_channels)])
        if up:
            self.upsamplers = nn.ModuleList([Upsample2D(out_channels)])


class UNet3DConditionModel(nn.Module):
    def __init__(self, cfg: UNet3DConfig = ZEROSCOPE):
        super().__init__()
        self.cfg = cfg
        ch = cfg.block_out_channels
        temb = ch[0] * 4
        self.conv_in = nn.Conv2d(cfg.in_channels, ch[0], 3, padding=1)
        self.time_embedding = TimestepEmbedding(ch[0], temb)
        self.transformer_in = TransformerTemporalModel(cfg.transformer_in_heads, cfg.attention_head_dim, ch[0],
                                                       cfg.norm_num_groups)
        self.down_blocks = nn.ModuleList()
        prev = ch[0]
        for i, c in enumerate(ch):
            last = i == len(ch) - 1
            io = [(prev if j == 0 else c, c) for j in range(cfg.layers_per_block)]
            self.down_blocks.append(Block3D(cfg, io, c, temb, attn=not last, n_attn=cfg.layers_per_block, down=not last))
            prev = c
        self.mid_block = Block3D(cfg, [(ch[-1], ch[-1]), (ch[-1], ch[-1])], ch[-1], temb, attn=True, n_attn=1)
        self.up_blocks = nn.ModuleList()
        rev = list(reversed(ch))
        prev_out = rev[0]
        for i, c in enumerate(rev):
            inc = rev[min(i + 1, len(ch) - 1)]
            nl = cfg.layers_per_block + 1
            io = []
            for j in range(nl):
                skip = inc if j == nl - 1 else c
                rin = prev_out if j == 0 else c
                io.append((rin + skip, c))
            self.up_blocks.append(Block3D(cfg, io, c, temb, attn=i > 0, n_attn=nl, up=i < len(ch) - 1))
            prev_out = c
        self.conv_norm_out = nn.GroupNorm(cfg.norm_num_groups, ch[0], eps=cfg.norm_eps)
        self.conv_out = nn.Conv2d(ch[0], cfg.out_channels, 3, padding=1)

    def forward(self, sample, timestep, encoder_hidden_states):
        b, _, f, h, w = sample.shape
        t = torch.as_tensor(timestep, dtype=torch.float32, device=sample.device).reshape(-1).expand(b)
        emb = self.time_embedding(timestep_sinusoid(t, self.cfg.block_out_channels[0]).to(sample.dtype))
        emb = emb.repeat_interleave(f, dim=0)
        ctx = encoder_hidden_states.repeat_interleave(f, dim=0)
        x = sample.permute(0, 2, 1, 3, 4).reshape(b * f, -1, h, w)
        x = self.conv_in(x)
        x = self.transformer_in(x, f)
        skips = [x]
        for blk in self.down_blocks:
            for j, (res, tc) in enumerate(zip(blk.resnets, blk.temp_convs)):
                x = tc(res(x, emb), f)
                if blk.has_attn:
                    x = blk.temp_attentions[j](blk.attentions[j](x, ctx), f)
                skips.append(x)
            if hasattr(blk, "downsamplers"):
                x = blk.downsamplers[0](x)
                skips.append(x)
        m = self.mid_block
        x = m.temp_convs[0](m.resnets[0](x, emb), f)
        x = m.temp_attentions[0](m.attentions[0](x, ctx), f)
        x = m.temp_convs[1](m.resnets[1](x, emb), f)
        for blk in self.up_blocks:
            for j, (res, tc) in enumerate(zip(blk.resnets, blk.temp_convs)):
                x = torch.cat([x, skips.pop()], dim=1)
                x = tc(res(x, emb), f)
                if blk.has_attn:
                    x = blk.temp_attentions[j](blk.attentions[j](x, ctx), f)
            if hasattr(blk, "upsamplers"):
                x = blk.upsamplers[0](x)
        x = self.conv_out(F.silu(self.conv_norm_out(x)))
        return x[None, :].reshape((-1, f) + x.shape[1:]).permute(0, 2, 1, 3, 4)


def random_init_(model: nn.Module, seed: int = 1234, round_bf16: bool = False) -> nn.Module:
    """SURVEY §8c random-init protocol: overwrite EVERY parameter (including the zero-initialised temp_convs.*.conv4)
    with seeded normal values scaled by fan-in; norm weights around 1. Optionally round to values that are exactly
    representable in bf16 AND in fp16 (bf16 rounding, magnitudes below fp16's smallest normal 2^-14 flushed to zero) so
    that the oracle and the kernels — in either operand precision — multiply identical weights."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in model.named_parameters():
            if p.ndim == 1:
                if name.endswith("weight"):  # norm gamma
                    v = 1.0 + 0.1 * torch.randn(p.shape, generator=g)
                else:
                    v = 0.05 * torch.randn(p.shape, generator=g)
            else:
                fan_in = p[0].numel()
                v = torch.randn(p.shape, generator=g) / math.sqrt(fan_in)
            if round_bf16:
                v = v.to(torch.bfloat16).to(torch.float32)
                v = torch.where(v.abs() < 2.0 ** -14, torch.zeros_like(v), v)
            p.copy_(v)
    return model


def param_count(cfg: UNet3DConfig = ZEROSCOPE) -> int:
    with torch.device("meta"):
        m = UNet3DConditionModel(cfg)
    return sum(p.numel() for p in m.parameters())
