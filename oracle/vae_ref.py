"""ORACLE — test infrastructure only (imported by tests/ and nothing on the product path).

Plain-torch fp32 CPU restatement of what the reference calls as ``pipe.vae.decode(z / 0.18215).sample`` followed by the
frame conversion ``(img * 0.5 + 0.5).clamp(0, 1) * 255 -> uint8`` (Distribution/strategies/chunk_only.py:152-158,
fsdp_chunked_coherent.py:219-227, InferNet/neurons/miner.py:594-600) — SURVEY.md §8f row N1 / U10.

The arithmetic is ``diffusers.models.autoencoder_kl.AutoencoderKL.decode`` (+ ``vae.Decoder``, ``unet_2d_blocks.
{UNetMidBlock2D,UpDecoderBlock2D}``, ``resnet.{ResnetBlock2D,Upsample2D}``, ``attention_processor.Attention``) — a
third-party dependency that is not vendored in /root/reference and not installed here (InferNet/requirements.txt:10
``diffusers>=0.19.0``, unpinned).  Restated for the Stable-Diffusion VAE configuration that the Zeroscope checkpoints
ship (``latent_channels=4, block_out_channels=(128,256,512,512), layers_per_block=2, norm_num_groups=32,
scaling_factor=0.18215``); module names are chosen so that ``state_dict()`` keys equal the checkpoint's
(``post_quant_conv``, ``decoder.conv_in``, ``decoder.mid_block.{resnets,attentions}``, ``decoder.up_blocks.i.
{resnets,upsamplers}``, ``decoder.conv_norm_out``, ``decoder.conv_out``).

PARITY UNPINNED: the reference holds no test or golden vector for this path and diffusers cannot be imported here.
Anchor: the decoder + post_quant_conv parameter total 49 490 199 (the public SD VAE decoder; `param_count()`).
"""
import math
from dataclasses import dataclass
from typing import Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F


@dataclass
class VAEConfig:
    latent_channels: int = 4
    out_channels: int = 3
    block_out_channels: Tuple[int, ...] = (128, 256, 512, 512)
    layers_per_block: int = 2
    norm_num_groups: int = 32
    norm_eps: float = 1e-6
    scaling_factor: float = 0.18215

    def to_dict(self):
        return dict(self.__dict__)


SD_VAE = VAEConfig()
# narrow variant for fast tests: same topology, widths multiples of 64
TINY = VAEConfig(block_out_channels=(64, 64, 128, 128))


class ResnetBlock2D(nn.Module):
    """diffusers ResnetBlock2D with temb_channels=None: GN -> SiLU -> conv3x3 -> GN -> SiLU -> conv3x3, + shortcut."""

    def __init__(self, cin, cout, groups, eps):
        super().__init__()
        self.norm1 = nn.GroupNorm(groups, cin, eps=eps)
        self.conv1 = nn.Conv2d(cin, cout, 3, padding=1)
        self.norm2 = nn.GroupNorm(groups, cout, eps=eps)
        self.conv2 = nn.Conv2d(cout, cout, 3, padding=1)
        self.conv_shortcut = nn.Conv2d(cin, cout, 1) if cin != cout else None

    def forward(self, x):
        h = self.conv1(F.silu(self.norm1(x)))
        h = self.conv2(F.silu(self.norm2(h)))
        return (x if self.conv_shortcut is None else self.conv_shortcut(x)) + h


class Attention(nn.Module):
    """The mid-block self attention: one head of width C over the h*w positions, GroupNorm in front, residual behind."""

    def __init__(self, c, groups, eps):
        super().__init__()
        self.group_norm = nn.GroupNorm(groups, c, eps=eps)
        self.to_q = nn.Linear(c, c)
        self.to_k = nn.Linear(c, c)
        self.to_v = nn.Linear(c, c)
        self.to_out = nn.ModuleList([nn.Linear(c, c), nn.Identity()])

    def forward(self, x):
        n, c, h, w = x.shape
        t = self.group_norm(x).view(n, c, h * w).transpose(1, 2)          # [n, S, C]
        q, k, v = self.to_q(t), self.to_k(t), self.to_v(t)
        p = torch.softmax(q @ k.transpose(1, 2) / math.sqrt(c), dim=-1)
        o = self.to_out[0](p @ v)
        return x + o.transpose(1, 2).reshape(n, c, h, w)


class MidBlock(nn.Module):
    def __init__(self, c, groups, eps):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(c, c, groups, eps), ResnetBlock2D(c, c, groups, eps)])
        self.attentions = nn.ModuleList([Attention(c, groups, eps)])

    def forward(self, x):
        return self.resnets[1](self.attentions[0](self.resnets[0](x)))


class Upsample2D(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, padding=1)

    def forward(self, x):
        return self.conv(F.interpolate(x, scale_factor=2.0, mode="nearest"))


class UpDecoderBlock2D(nn.Module):
    def __init__(self, cin, cout, n_layers, groups, eps, upsample):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(cin if i == 0 else cout, cout, groups, eps) for i in range(n_layers)])
        self.upsamplers = nn.ModuleList([Upsample2D(cout)]) if upsample else None

    def forward(self, x):
        for r in self.resnets:
            x = r(x)
        if self.upsamplers is not None:
            x = self.upsamplers[0](x)
        return x


class Decoder(nn.Module):
    def __init__(self, cfg: VAEConfig):
        super().__init__()
        ch = list(reversed(cfg.block_out_channels))
        g, eps = cfg.norm_num_groups, cfg.norm_eps
        self.conv_in = nn.Conv2d(cfg.latent_channels, ch[0], 3, padding=1)
        self.mid_block = MidBlock(ch[0], g, eps)
        blocks, prev = [], ch[0]
        for i, c in enumerate(ch):
            blocks.append(UpDecoderBlock2D(prev, c, cfg.layers_per_block + 1, g, eps, upsample=i != len(ch) - 1))
            prev = c
        self.up_blocks = nn.ModuleList(blocks)
        self.conv_norm_out = nn.GroupNorm(g, ch[-1], eps=eps)
        self.conv_out = nn.Conv2d(ch[-1], cfg.out_channels, 3, padding=1)

    def forward(self, z):
        x = self.mid_block(self.conv_in(z))
        for b in self.up_blocks:
            x = b(x)
        return self.conv_out(F.silu(self.conv_norm_out(x)))


class AutoencoderKLDecoder(nn.Module):
    """``decode(z)``: post_quant_conv (1x1) then the decoder. (The encoder half is never used by the reference.)"""

    def __init__(self, cfg: VAEConfig = SD_VAE):
        super().__init__()
        self.cfg = cfg
        self.post_quant_conv = nn.Conv2d(cfg.latent_channels, cfg.latent_channels, 1)
        self.decoder = Decoder(cfg)

    def decode(self, z):
        return self.decoder(self.post_quant_conv(z))


def frames_u8(model: AutoencoderKLDecoder, lat: torch.Tensor):
    """chunk_only.py:152-158: per frame ``vae.decode(z / 0.18215).sample`` -> ``(img*0.5+0.5).clamp(0,1)*255`` ->
    uint8 (truncation) [H,W,3]. lat: [1,4,T,h,w] fp32. Returns (uint8 [T,H,W,3], fp32 pre-quantisation images
    [T,H,W,3] in [0,255])."""
    u8, raw = [], []
    with torch.no_grad():
        for i in range(lat.shape[2]):
            img = model.decode(lat[:, :, i] / model.cfg.scaling_factor)
            img = (img[0].permute(1, 2, 0) * 0.5 + 0.5).clamp(0, 1) * 255
            raw.append(img)
            u8.append(img.to(torch.uint8))
    return torch.stack(u8), torch.stack(raw)


def param_count(cfg: VAEConfig = SD_VAE) -> int:
    with torch.device("meta"):
        m = AutoencoderKLDecoder(cfg)
    return sum(p.numel() for p in m.parameters())
