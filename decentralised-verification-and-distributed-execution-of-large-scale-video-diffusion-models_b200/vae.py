"""``pipe.vae`` of the reference on libb2v: the frame decode that follows the denoise loop.

Mirrors the call surface the reference uses on diffusers' ``AutoencoderKL``:
    img = pipe.vae.decode(z / 0.18215).sample                  (chunk_only.py:156, fsdp_chunked_coherent.py:223)
    pipe.vae.decode(x / pipe.vae.config.scaling_factor).sample (miner.py:596)
and adds the batched form the strategies use here, which does the reference's whole per-frame loop on the device:
    frames = pipe.vae.decode_frames(lat)   # [1,4,T,h,w] -> uint8 [T,H,W,3], `/0.18215`, `*0.5+0.5`, clamp, *255 fused

Every convolution / projection is a tcgen05 GEMM of csrc/gemm.cu (3x3 through TMA im2col), GroupNorm+SiLU is
csrc/norm.cu, the mid-block attention (ONE head of width 512) is S = Q K^T -> row softmax -> P V^T as two GEMMs and
``b2v_op_softmax_rows``; V is produced already transposed by swapping the operands of its projection, and its bias is
added after P V (rows of P sum to one). Weights are fed under their diffusers ``state_dict`` names. No CPU path.
"""
import math
import os
from types import SimpleNamespace

import torch

from . import _lib, ops

SD_VAE_CONFIG = dict(latent_channels=4, out_channels=3, block_out_channels=(128, 256, 512, 512), layers_per_block=2,
                     norm_num_groups=32, norm_eps=1e-6, scaling_factor=0.18215)
_PAD = 64  # conv_in input channels / conv_out output channels are zero padded to one 64-wide tile


class DecoderOutput:
    __slots__ = ("sample",)

    def __init__(self, sample):
        self.sample = sample

    def __getitem__(self, i):
        return (self.sample,)[i]


def parameter_shapes(cfg):
    """{state_dict name: shape} of post_quant_conv + decoder for a config."""
    ch = list(reversed(cfg["block_out_channels"]))
    L, out = cfg["latent_channels"], {}

    def conv(p, co, ci, k=3):
        out[p + ".weight"], out[p + ".bias"] = (co, ci, k, k), (co,)

    def norm(p, c):
        out[p + ".weight"], out[p + ".bias"] = (c,), (c,)

    def lin(p, co, ci):
        out[p + ".weight"], out[p + ".bias"] = (co, ci), (co,)

    def resnet(p, ci, co):
        norm(p + ".norm1", ci); conv(p + ".conv1", co, ci); norm(p + ".norm2", co); conv(p + ".conv2", co, co)
        if ci != co:
            conv(p + ".conv_shortcut", co, ci, 1)

    conv("post_quant_conv", L, L, 1)
    conv("decoder.conv_in", ch[0], L)
    resnet("decoder.mid_block.resnets.0", ch[0], ch[0])
    a = "decoder.mid_block.attentions.0"
    norm(a + ".group_norm", ch[0])
    for n in ("to_q", "to_k", "to_v", "to_out.0"):
        lin(a + "." + n, ch[0], ch[0])
    resnet("decoder.mid_block.resnets.1", ch[0], ch[0])
    prev = ch[0]
    for i, c in enumerate(ch):
        for j in range(cfg["layers_per_block"] + 1):
            resnet("decoder.up_blocks.%d.resnets.%d" % (i, j), prev if j == 0 else c, c)
        if i != len(ch) - 1:
            conv("decoder.up_blocks.%d.upsamplers.0.conv" % i, c, c)
        prev = c
    norm("decoder.conv_norm_out", ch[-1])
    conv("decoder.conv_out", cfg["out_channels"], ch[-1])
    return out


def random_state_dict(cfg, device, seed=4321):
    """Seeded random weights (no network for checkpoints): normal / sqrt(fan_in), norm gains around 1."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for name, shape in parameter_shapes(cfg).items():
        if len(shape) == 1:
            r = torch.randn(shape, generator=g)
            v = 1.0 + 0.1 * r if name.endswith("weight") else 0.05 * r
        else:
            v = torch.randn(shape, generator=g) / math.sqrt(math.prod(shape[1:]))
        sd[name] = v.to(torch.bfloat16).float().to(device)
    return sd


_LEGACY_ATTENTION_KEYS = {"query": "to_q", "key": "to_k", "value": "to_v", "proj_attn": "to_out.0"}


def canonical_keys(sd):
    """Checkpoints saved by older diffusers releases (the 2023 Zeroscope / SD VAE files among them) name the mid-block
    attention projections ``query / key / value / proj_attn``; diffusers renames them to ``to_q / to_k / to_v /
    to_out.0`` on load. Same here; encoder / quant_conv tensors (never used by decode) are dropped."""
    out = {}
    for k, v in sd.items():
        if k.startswith(("encoder.", "quant_conv.")):
            continue
        parts = k.split(".")
        if len(parts) >= 2 and ".attentions." in k and parts[-2] in _LEGACY_ATTENTION_KEYS:
            parts[-2:-1] = _LEGACY_ATTENTION_KEYS[parts[-2]].split(".")
            k = ".".join(parts)
        out[k] = v
    return out


def _gn_fuse():
    return os.environ.get("B2V_GN_FUSE", "1") != "0"


class AutoencoderKL:
    def __init__(self, config=None, device="cuda"):
        cfg = dict(SD_VAE_CONFIG)
        if config:
            cfg.update({k: v for k, v in dict(config).items() if k in cfg})
        cfg["block_out_channels"] = tuple(int(c) for c in cfg["block_out_channels"])
        self.config = SimpleNamespace(**cfg)
        dev = device if isinstance(device, torch.device) else torch.device(device)
        if dev.type != "cuda":
            raise _lib.B2VError("AutoencoderKL runs on an sm_100 CUDA device only (no CPU path)")
        if dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        self.device, self.dtype = dev, torch.bfloat16
        if any(c % 64 for c in cfg["block_out_channels"]) or cfg["latent_channels"] > _PAD or cfg["out_channels"] > 4:
            raise ValueError("decoder widths must be multiples of 64 (latent <= 64 channels, <= 4 output channels)")
        self.frames_per_batch = None  # None = sized from the output resolution (see _batch)
        self._w = None
        _lib.load()

    # ---- weights ------------------------------------------------------------------------------------------------
    def load_state_dict(self, sd):
        cfg, dev = self.config.__dict__, self.device
        shapes = parameter_shapes(cfg)
        sd = canonical_keys(sd)
        missing = [k for k in shapes if k not in sd]
        if missing:
            raise _lib.B2VError("vae state_dict misses %d tensors, e.g. %s" % (len(missing), missing[:3]))
        for k, shp in shapes.items():
            if len(shp) == 2 and tuple(sd[k].shape) == tuple(shp) + (1, 1):
                continue  # LDM-style checkpoints keep the attention projections as 1x1 convolutions
            if tuple(sd[k].shape) != tuple(shp):
                raise _lib.B2VError("vae tensor %s has shape %s, expected %s" % (k, tuple(sd[k].shape), shp))
        f32 = lambda k: sd[k].detach().to(dev, torch.float32).contiguous()  # noqa: E731
        W = {}

        def conv3(p, pad_in=0, pad_out=0):
            w = f32(p + ".weight").permute(0, 2, 3, 1)                   # [Cout,3,3,Cin]
            b = f32(p + ".bias")
            if pad_in:
                w = torch.nn.functional.pad(w, (0, pad_in - w.shape[3]))
            if pad_out:
                w = torch.nn.functional.pad(w, (0, 0, 0, 0, 0, 0, 0, pad_out - w.shape[0]))
                b = torch.nn.functional.pad(b, (0, pad_out - b.shape[0]))
            W[p] = (w.to(torch.bfloat16).contiguous(), b.contiguous())

        def mat(p, key=None):
            w = f32(p + ".weight")
            W[key or p] = (w.reshape(w.shape[0], -1).to(torch.bfloat16).contiguous(), f32(p + ".bias"))

        def norm(p):
            W[p] = (f32(p + ".weight"), f32(p + ".bias"))

        # post_quant_conv on padded rows; the `/ scaling_factor` of decode_frames is folded into a second copy
        L = cfg["latent_channels"]
        pq, pqb = f32("post_quant_conv.weight").reshape(L, L), f32("post_quant_conv.bias")
        for key, scale in (("pq", 1.0), ("pq_scaled", 1.0 / cfg["scaling_factor"])):
            w = torch.zeros((_PAD, _PAD), device=dev)
            w[:L, :L] = pq * scale
            b = torch.zeros(_PAD, device=dev)
            b[:L] = pqb
            W[key] = (w.to(torch.bfloat16).contiguous(), b)
        conv3("decoder.conv_in", pad_in=_PAD)
        for name in shapes:
            if name.endswith(".weight"):
                p = name[:-7]
                if p in W or p in ("post_quant_conv", "decoder.conv_in", "decoder.conv_out"):
                    continue
                if len(shapes[name]) == 1:
                    norm(p)
                elif len(shapes[name]) == 2 or shapes[name][2] == 1:
                    mat(p)
                else:
                    conv3(p)
        conv3("decoder.conv_out", pad_out=_PAD)
        self._w = W
        self.n_params = sum(math.prod(s) for s in shapes.values())
        return self

    @classmethod
    def from_random(cls, config=None, device="cuda", seed=4321):
        m = cls(config, device=device)
        return m.load_state_dict(random_state_dict(m.config.__dict__, m.device, seed))

    # ---- layers (activations: channels-last bf16 rows [n*H*W, C]) -----------------------------------------------
    def _gn(self, x, p, n, silu=True):
        g, b = self._w[p]
        pre = getattr(x, "_b2v_gn_stats", None)   # emitted by the conv / GEMM that produced x (see _conv)
        if pre is not None:
            return ops.group_norm_pre(x, g, b, n, self.config.norm_num_groups, self.config.norm_eps, silu, pre)
        return ops.group_norm(x, g, b, n, self.config.norm_num_groups, self.config.norm_eps, silu)

    def _stats_ok(self, M, Cout, rows_per_item):
        """May the producer of an [M, Cout] activation emit GroupNorm partials? The GEMM side needs whole 256-row tiles
        (b2v_gn_stats_ok); the consuming b2v_op_group_norm_pre folds them per item in 32-row units, so a frame whose
        H*W is not a multiple of 32 (20x20 latents) must take the two-pass norm instead, as unet.cu's gn() does."""
        return (_gn_fuse() and rows_per_item % 32 == 0
                and ops.gn_stats_ok(M, Cout, 0, self.config.norm_num_groups))

    def _conv(self, x, p, n, H, W_, residual=None, feeds_gn=True):
        """3x3 conv on rows [n*H*W, Cin]. Every conv of the decoder but conv_out feeds a GroupNorm: where the shape has a
        statistics epilogue (b2v_gn_stats_ok) the conv also emits the {sum, sumsq} partials of its output and the norm
        skips its first pass over the activation (B2V_GN_FUSE=0: off)."""
        w, b = self._w[p]
        Cout, M = w.shape[0], n * H * W_
        x4 = x.view(n, H, W_, x.shape[1])
        if feeds_gn and self._stats_ok(M, Cout, H * W_):
            out, st = ops.conv3x3_gnstats(x4, w, b, residual=residual, groups=self.config.norm_num_groups)
            out = out.view(M, Cout)
            out._b2v_gn_stats = st
            return out
        return ops.conv3x3(x4, w, b, residual=residual).view(M, Cout)

    def _resnet(self, x, p, n, H, W_):
        h = self._conv(self._gn(x, p + ".norm1", n), p + ".conv1", n, H, W_)
        h = self._gn(h, p + ".norm2", n)
        sc = x
        if p + ".conv_shortcut" in self._w:
            w, b = self._w[p + ".conv_shortcut"]
            sc = ops.gemm(x, w, bias=b)
        return self._conv(h, p + ".conv2", n, H, W_, residual=sc)

    def _attention(self, x, p, n, S):
        Cc = x.shape[1]
        t = self._gn(x, p + ".group_norm", n, silu=False)
        wq, bq = self._w[p + ".to_q"]
        wk, bk = self._w[p + ".to_k"]
        wv, bv = self._w[p + ".to_v"]
        q, k = ops.gemm(t, wq, bias=bq), ops.gemm(t, wk, bias=bk)
        wo, bo = self._w[p + ".to_out.0"]
        o = torch.empty_like(t)
        scores = torch.empty((S, S), dtype=torch.bfloat16, device=x.device)
        vt = torch.empty((Cc, S), dtype=torch.bfloat16, device=x.device)
        for f in range(n):
            rows = slice(f * S, (f + 1) * S)
            ops.gemm(q[rows], k[rows], out=scores)                       # S = Q K^T
            ops.softmax_rows_(scores, 1.0 / math.sqrt(Cc))               # P
            ops.gemm(wv, t[rows], out=vt)                                # V^T = Wv X^T (bias deferred)
            ops.gemm(scores, vt, bias=bv, out=o[rows])                   # O = P V + bv
        if self._stats_ok(o.shape[0], wo.shape[0], S):
            out, st = ops.gemm_gnstats(o, wo, bias=bo, residual=x, groups=self.config.norm_num_groups)
            out._b2v_gn_stats = st   # the next resnet's norm1 reads them
            return out
        return ops.gemm(o, wo, bias=bo, residual=x)

    def _batch(self, T, H, W_):
        if self.frames_per_batch:
            return max(1, min(T, int(self.frames_per_batch)))
        widest = max(self.config.block_out_channels[:2])                 # channels alive at the output resolution
        return max(1, min(T, (1 << 30) // max(1, H * W_ * widest)))    # < 2^30 elements per activation

    def _decode_rows(self, rows, n, h, w, scaled):
        """rows: [n*h*w, 64] padded latent rows -> decoder output rows [n*H*W, 64] (RGB in channels 0..2), H, W."""
        if self._w is None:
            raise _lib.B2VError("vae weights not loaded")
        ch = list(reversed(self.config.block_out_channels))
        pw, pb = self._w["pq_scaled" if scaled else "pq"]
        x = ops.gemm(rows, pw, bias=pb)
        x = self._conv(x, "decoder.conv_in", n, h, w)
        x = self._resnet(x, "decoder.mid_block.resnets.0", n, h, w)
        x = self._attention(x, "decoder.mid_block.attentions.0", n, h * w)
        x = self._resnet(x, "decoder.mid_block.resnets.1", n, h, w)
        H, W_ = h, w
        for i in range(len(ch)):
            for j in range(self.config.layers_per_block + 1):
                x = self._resnet(x, "decoder.up_blocks.%d.resnets.%d" % (i, j), n, H, W_)
            if i != len(ch) - 1:
                x = ops.upsample2x(x.view(n, H, W_, x.shape[1]))
                H, W_ = 2 * H, 2 * W_
                x = self._conv(x.view(n * H * W_, -1), "decoder.up_blocks.%d.upsamplers.0.conv" % i, n, H, W_)
        x = self._gn(x, "decoder.conv_norm_out", n)
        return self._conv(x, "decoder.conv_out", n, H, W_, feeds_gn=False), H, W_

    # ---- the reference's call: per-frame / per-batch decode of already scaled latents ---------------------------
    @torch.no_grad()
    def decode(self, z, return_dict=True):
        """z: [n,4,h,w] (already divided by the scaling factor, as the reference does) -> .sample [n,3,8h,8w]."""
        if z.dim() != 4 or z.shape[1] != self.config.latent_channels:
            raise ValueError("decode expects [n,%d,h,w], got %s" % (self.config.latent_channels, tuple(z.shape)))
        n, Cc, h, w = z.shape
        if (h * w) % 64:
            raise ValueError("latent h*w must be a multiple of 64 (attention tile), got %dx%d" % (h, w))
        zz = z.to(self.device).permute(1, 0, 2, 3).unsqueeze(0).contiguous()     # [1,C,n,h,w]: frames on the F axis
        outs = []
        nb = self._batch(n, 8 * h, 8 * w)
        for s in range(0, n, nb):
            e = min(n, s + nb)
            rows = ops.latent_rows(zz[:, :, s:e].contiguous(), _PAD)
            y, H, W_ = self._decode_rows(rows, e - s, h, w, scaled=False)
            outs.append(y.view(e - s, H, W_, _PAD)[..., :self.config.out_channels].permute(0, 3, 1, 2))
        img = torch.cat(outs).to(z.dtype if z.dtype.is_floating_point else torch.float32).contiguous()
        return DecoderOutput(img) if return_dict else (img,)

    @torch.no_grad()
    def decode_frames(self, lat):
        """The reference's whole decode loop (chunk_only.py:152-158) on the device: lat [1,4,T,h,w] (unscaled, any
        float dtype) -> uint8 [T,H,W,3] device tensor; `/scaling_factor` is folded into post_quant_conv and
        `*0.5+0.5, clamp, *255, byte()` is one kernel over the decoder output."""
        if lat.dim() != 5 or lat.shape[0] != 1 or lat.shape[1] != self.config.latent_channels:
            raise ValueError("decode_frames expects [1,%d,T,h,w], got %s" % (self.config.latent_channels, tuple(lat.shape)))
        _, _, T, h, w = lat.shape
        if (h * w) % 64:
            raise ValueError("latent h*w must be a multiple of 64 (attention tile), got %dx%d" % (h, w))
        lat = lat.to(self.device)
        out = torch.empty((T, 8 * h, 8 * w, 3), dtype=torch.uint8, device=self.device)
        nb = self._batch(T, 8 * h, 8 * w)
        for s in range(0, T, nb):
            e = min(T, s + nb)
            rows = ops.latent_rows(lat[:, :, s:e].contiguous(), _PAD)
            y, H, W_ = self._decode_rows(rows, e - s, h, w, scaled=True)
            ops.frames_u8(y, (e - s) * H * W_, out=out[s:e].view(-1, 3))
        return out

    def flops(self, T, h, w):
        """Algorithmic FLOPs of decoding T frames (convs, projections, attention; MAC = 2)."""
        cfg = self.config
        ch = list(reversed(cfg.block_out_channels))
        S, fl = h * w, 0.0
        conv = lambda px, ci, co, k=9: 2.0 * px * ci * co * k  # noqa: E731
        res = lambda px, ci, co: conv(px, ci, co) + conv(px, co, co) + (conv(px, ci, co, 1) if ci != co else 0)  # noqa: E731
        fl += conv(S, cfg.latent_channels, cfg.latent_channels, 1) + conv(S, cfg.latent_channels, ch[0])
        fl += 2 * res(S, ch[0], ch[0]) + 4 * conv(S, ch[0], ch[0], 1) + 4.0 * S * S * ch[0]
        px, prev = S, ch[0]
        for i, c in enumerate(ch):
            fl += res(px, prev, c) + cfg.layers_per_block * res(px, c, c)
            if i != len(ch) - 1:
                px *= 4
                fl += conv(px, c, c)
            prev = c
        fl += conv(px, ch[-1], cfg.out_channels)
        return T * fl
