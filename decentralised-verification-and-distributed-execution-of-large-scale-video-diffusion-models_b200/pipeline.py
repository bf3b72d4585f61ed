"""A ``pipe``-shaped object with the attributes the reference's hot loop touches on diffusers'
``TextToVideoSDPipeline`` (Distribution/strategies/chunk_only.py:44-74, InferNet/neurons/miner.py:83-100,531-600):

    pipe.unet(x, t, encoder_hidden_states=emb).sample      pipe.unet.config.in_channels
    pipe.scheduler.{set_timesteps, timesteps, scale_model_input, step, init_noise_sigma, alphas_cumprod}
    pipe.tokenizer(...), pipe.text_encoder(ids)[0]          pipe.device, pipe.to(device)
    pipe.vae.decode(z / scaling).sample                     (attached by the caller; "next" row N1 of SURVEY.md §8f)

Model fetch (HF hub) is out of scope: weights are random-initialised from a seed with the protocol of SURVEY.md §8c
(every parameter overwritten, including the zero-initialised temp_convs.*.conv4), or loaded from a diffusers
``state_dict``. The CLIP text tower is not on the accelerated path: ``SyntheticTextEncoder`` maps token ids to a
seeded [n,77,1024] embedding so the strategies' tokenizer -> text_encoder call sequence runs unchanged.
"""
import hashlib
import math
import os
from types import SimpleNamespace

import torch

from .scheduler import DDIMScheduler
from .unet import UNet3DConditionModel, ZEROSCOPE_CONFIG

# architecture of cerspense/zeroscope_v2_576w and zeroscope_v2_XL (identical UNet3D; they differ in weights)
KNOWN_MODELS = {"cerspense/zeroscope_v2_XL": ZEROSCOPE_CONFIG, "cerspense/zeroscope_v2_576w": ZEROSCOPE_CONFIG}


def parameter_shapes(cfg):
    """diffusers state_dict names -> shapes of UNet3DConditionModel for `cfg` (no memory is allocated)."""
    ch = list(cfg["block_out_channels"])
    temb, g_in, g_out = ch[0] * 4, cfg["in_channels"], cfg["out_channels"]
    hd, xd, L = cfg["attention_head_dim"], cfg["cross_attention_dim"], cfg["layers_per_block"]
    s = {}

    def lin(p, n, k, bias=True):
        s[p + ".weight"] = (n, k)
        if bias:
            s[p + ".bias"] = (n,)

    def norm(p, c):
        s[p + ".weight"] = (c,)
        s[p + ".bias"] = (c,)

    def conv(p, co, ci, k=3):
        s[p + ".weight"] = (co, ci, k, k)
        s[p + ".bias"] = (co,)

    def block(p, dim, cross):
        norm(p + ".norm1", dim)
        for a in ("attn1", "attn2"):
            kv = xd if (a == "attn2" and cross) else dim
            lin(p + "." + a + ".to_q", dim, dim, False)
            lin(p + "." + a + ".to_k", dim, kv, False)
            lin(p + "." + a + ".to_v", dim, kv, False)
            lin(p + "." + a + ".to_out.0", dim, dim)
        norm(p + ".norm2", dim)
        norm(p + ".norm3", dim)
        lin(p + ".ff.net.0.proj", 8 * dim, dim)
        lin(p + ".ff.net.2", dim, 4 * dim)

    def transformer(p, c, inner, cross):
        norm(p + ".norm", c)
        lin(p + ".proj_in", inner, c)
        block(p + ".transformer_blocks.0", inner, cross)
        lin(p + ".proj_out", c, inner)

    def resnet(p, ci, co):
        norm(p + ".norm1", ci)
        conv(p + ".conv1", co, ci)
        lin(p + ".time_emb_proj", co, temb)
        norm(p + ".norm2", co)
        conv(p + ".conv2", co, co)
        if ci != co:
            conv(p + ".conv_shortcut", co, ci, 1)

    def temp_conv(p, c):
        for gn, cv in ((".conv1.0", ".conv1.2"), (".conv2.0", ".conv2.3"), (".conv3.0", ".conv3.3"),
                       (".conv4.0", ".conv4.3")):
            norm(p + gn, c)
            s[p + cv + ".weight"] = (c, c, 3, 1, 1)
            s[p + cv + ".bias"] = (c,)

    conv("conv_in", ch[0], g_in)
    lin("time_embedding.linear_1", temb, ch[0])
    lin("time_embedding.linear_2", temb, temb)
    transformer("transformer_in", ch[0], cfg["transformer_in_heads"] * hd, False)
    prev = ch[0]
    for i, c in enumerate(ch):
        p = "down_blocks.%d" % i
        for j in range(L):
            resnet("%s.resnets.%d" % (p, j), prev if j == 0 else c, c)
            temp_conv("%s.temp_convs.%d" % (p, j), c)
            if i < len(ch) - 1:
                transformer("%s.attentions.%d" % (p, j), c, c, True)
                transformer("%s.temp_attentions.%d" % (p, j), c, c, False)
        if i < len(ch) - 1:
            conv(p + ".downsamplers.0.conv", c, c)
        prev = c
    for j in range(2):
        resnet("mid_block.resnets.%d" % j, ch[-1], ch[-1])
        temp_conv("mid_block.temp_convs.%d" % j, ch[-1])
    transformer("mid_block.attentions.0", ch[-1], ch[-1], True)
    transformer("mid_block.temp_attentions.0", ch[-1], ch[-1], False)
    rev = ch[::-1]
    prev_out = rev[0]
    for i, c in enumerate(rev):
        p = "up_blocks.%d" % i
        inc = rev[min(i + 1, len(ch) - 1)]
        for j in range(L + 1):
            skip = inc if j == L else c
            resnet("%s.resnets.%d" % (p, j), (prev_out if j == 0 else c) + skip, c)
            temp_conv("%s.temp_convs.%d" % (p, j), c)
            if i > 0:
                transformer("%s.attentions.%d" % (p, j), c, c, True)
                transformer("%s.temp_attentions.%d" % (p, j), c, c, False)
        if i < len(ch) - 1:
            conv(p + ".upsamplers.0.conv", c, c)
        prev_out = c
    norm("conv_norm_out", ch[0])
    conv("conv_out", g_out, ch[0])
    return s


def random_state_dict(cfg, device, seed=1234):
    """Seeded random weights generated directly on the device: fan-in scaled normals for matrices / kernels,
    1 + 0.1 N for norm gains, 0.05 N for biases (SURVEY.md §8c random-init protocol)."""
    g = torch.Generator(device=device).manual_seed(seed)
    sd = {}
    for name, shape in parameter_shapes(cfg).items():
        if len(shape) == 1:
            v = torch.randn(shape, device=device, generator=g)
            v = 1.0 + 0.1 * v if name.endswith("weight") else 0.05 * v
        else:
            v = torch.randn(shape, device=device, generator=g) / math.sqrt(math.prod(shape[1:]))
        sd[name] = v
    return sd


class SyntheticTokenizer:
    model_max_length = 77

    def __call__(self, text, padding="max_length", max_length=77, truncation=True, return_tensors="pt", **_):
        texts = [text] if isinstance(text, str) else list(text)
        ids = torch.zeros((len(texts), max_length), dtype=torch.int64)
        for i, s in enumerate(texts):
            h = hashlib.sha256(s.encode("utf-8")).digest()
            raw = (h * ((max_length * 2) // len(h) + 1))[:max_length * 2]
            ids[i] = torch.tensor([int.from_bytes(raw[2 * k:2 * k + 2], "big") % 49408 for k in range(max_length)])
        return SimpleNamespace(input_ids=ids, attention_mask=torch.ones_like(ids))


class SyntheticTextEncoder:
    """ids [n,77] -> ([n,77,dim],): unit-variance embeddings seeded by the token ids (not CLIP; random weights)."""

    def __init__(self, dim=1024, dtype=torch.bfloat16, device="cuda"):
        self.dim, self.dtype, self.device = dim, dtype, torch.device(device)

    def to(self, *a, **k):
        return self

    def __call__(self, input_ids=None, **kw):
        ids = input_ids.cpu()
        out = torch.empty((ids.shape[0], ids.shape[1], self.dim), dtype=torch.float32)
        for i in range(ids.shape[0]):
            g = torch.Generator().manual_seed(int(ids[i].sum()) * 2654435761 % (2 ** 31))
            out[i] = torch.randn(ids.shape[1], self.dim, generator=g)
        return (out.to(self.device, self.dtype),)


def unet_arith_dtype(dtype, unet_dtype=None):
    """Arithmetic dtype of the UNet for a pipeline whose latents are `dtype`: fp16 latents -> fp16 arithmetic (the
    reference's ``torch_dtype=torch.float16``), anything else -> bf16, unless stated explicitly."""
    if unet_dtype is not None:
        return unet_dtype
    return torch.float16 if dtype == torch.float16 else torch.bfloat16


class VideoPipeline:
    def __init__(self, unet, scheduler, text_dim=1024, dtype=torch.bfloat16, vae=None, tokenizer=None,
                 text_encoder=None):
        self.unet, self.scheduler, self.vae = unet, scheduler, vae
        self.device = unet.device
        self.dtype = dtype
        self.tokenizer = tokenizer if tokenizer is not None else SyntheticTokenizer()
        self.text_encoder = text_encoder if text_encoder is not None else SyntheticTextEncoder(text_dim, dtype, self.device)
        self.synthetic_text = text_encoder is None

    def to(self, *a, **k):
        return self

    @classmethod
    def from_random(cls, model_id="cerspense/zeroscope_v2_XL", device="cuda", seed=1234, dtype=torch.bfloat16,
                    config=None, unet_dtype=None):
        cfg = dict(config or KNOWN_MODELS.get(model_id) or ZEROSCOPE_CONFIG)
        unet = UNet3DConditionModel(cfg, device=device, dtype=unet_arith_dtype(dtype, unet_dtype))
        unet.load_state_dict(random_state_dict(unet.config.__dict__, unet.device, seed))
        return cls(unet, DDIMScheduler.from_pipeline_config(), text_dim=cfg["cross_attention_dim"], dtype=dtype)

    @classmethod
    def from_pretrained(cls, path, device="cuda", dtype=torch.bfloat16, torch_dtype=None, tokenizer=None,
                        text_encoder=None, allow_synthetic_text=False, unet_dtype=None, **_):
        """The reference's ``TextToVideoSDPipeline.from_pretrained(model_id, torch_dtype=...)`` (chunk_only.py:44-56) for
        a LOCAL diffusers-layout directory: unet (+ vae when present) weights and the scheduler config are read from
        disk. The CLIP text tower is outside the accelerated path: pass the real ``tokenizer=`` and
        ``text_encoder=`` (e.g. transformers' CLIPTokenizer / CLIPTextModel loaded from the same folder). Real UNet
        weights driven by the synthetic stand-in embeddings would silently ignore the prompt, so that combination has
        to be asked for with ``allow_synthetic_text=True`` (benchmarks, tests)."""
        from . import checkpoint
        from .vae import AutoencoderKL
        if not os.path.isdir(path):
            raise FileNotFoundError("%r is not a local directory (no network: pass a downloaded checkpoint folder)" % (path,))
        if (tokenizer is None or text_encoder is None) and not allow_synthetic_text:
            raise ValueError("from_pretrained(%r): no tokenizer= / text_encoder= given. The CLIP text tower is not part of "
                             "libb2v; with the synthetic stand-in the prompt would be ignored. Pass the real modules, or "
                             "allow_synthetic_text=True to run on seeded random embeddings on purpose." % (path,))
        dtype = torch_dtype or dtype
        sd, cfg = checkpoint.read_component(path, "unet")
        ucfg = dict(ZEROSCOPE_CONFIG)
        ucfg.update(checkpoint.unet_config(cfg))
        unet = UNet3DConditionModel(ucfg, device=device, dtype=unet_arith_dtype(dtype, unet_dtype)).load_state_dict(sd)
        sched = DDIMScheduler(**{**dict(DDIMScheduler.from_pipeline_config().config.__dict__),
                                 **checkpoint.scheduler_config(path)})
        vae = None
        if os.path.isdir(os.path.join(path, "vae")):
            vsd, vcfg = checkpoint.read_component(path, "vae")
            vae = AutoencoderKL(checkpoint.vae_config(vcfg), device=device).load_state_dict(vsd)
        return cls(unet, sched, text_dim=ucfg["cross_attention_dim"], dtype=dtype, vae=vae, tokenizer=tokenizer,
                   text_encoder=text_encoder)

    @classmethod
    def from_state_dict(cls, state_dict, config=None, device="cuda", dtype=torch.bfloat16, unet_dtype=None):
        cfg = dict(config or ZEROSCOPE_CONFIG)
        unet = UNet3DConditionModel(cfg, device=device, dtype=unet_arith_dtype(dtype, unet_dtype)).load_state_dict(state_dict)
        return cls(unet, DDIMScheduler.from_pipeline_config(), text_dim=cfg["cross_attention_dim"], dtype=dtype)
