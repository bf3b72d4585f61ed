"""not-gpu: the frame-decode oracle (oracle/vae_ref.py) and the host logic of b2v.vae (parameter table, FLOP count).
PARITY UNPINNED for the decoder arithmetic (diffusers is absent, the reference pins nothing here): the anchors are the
published parameter total of the SD VAE decoder and its state_dict key set."""
import math
from types import SimpleNamespace

import torch

from oracle import vae_ref as V


def test_parameter_total_and_keys():
    # SD VAE: decoder 49 490 179 + post_quant_conv 20 (the full autoencoder's published 83 653 863 minus the encoder
    # 34 163 592 and quant_conv 72)
    assert V.param_count() == 49490199 == 83653863 - 34163592 - 72
    sd = V.AutoencoderKLDecoder(V.SD_VAE).state_dict()
    assert len(sd) == 140
    for k in ("post_quant_conv.weight", "decoder.conv_in.weight", "decoder.mid_block.attentions.0.to_q.weight",
              "decoder.mid_block.attentions.0.to_out.0.bias", "decoder.mid_block.attentions.0.group_norm.weight",
              "decoder.up_blocks.2.resnets.0.conv_shortcut.weight", "decoder.up_blocks.0.upsamplers.0.conv.weight",
              "decoder.conv_norm_out.weight", "decoder.conv_out.bias"):
        assert k in sd, k
    assert "decoder.up_blocks.3.upsamplers.0.conv.weight" not in sd
    assert tuple(sd["decoder.up_blocks.3.resnets.0.conv1.weight"].shape) == (128, 256, 3, 3)


def test_host_parameter_table_matches_oracle():
    from b2v import vae
    for cfg in (V.SD_VAE, V.TINY):
        shapes = vae.parameter_shapes(cfg.to_dict())
        sd = V.AutoencoderKLDecoder(cfg).state_dict()
        assert set(shapes) == set(sd)
        assert all(tuple(sd[k].shape) == tuple(s) for k, s in shapes.items())


def test_decode_shapes_and_frame_bytes():
    from oracle.unet3d_ref import random_init_
    m = random_init_(V.AutoencoderKLDecoder(V.TINY), seed=3).eval()
    lat = torch.randn(1, 4, 2, 8, 8, generator=torch.Generator().manual_seed(0)) * 0.3
    u8, raw = V.frames_u8(m, lat)
    assert u8.shape == (2, 64, 64, 3) and u8.dtype == torch.uint8
    assert float(raw.min()) >= 0 and float(raw.max()) <= 255
    assert torch.equal(u8, raw.floor().to(torch.uint8))  # `.byte()` truncates
    with torch.no_grad():
        one = m.decode(lat[:, :, 1] / 0.18215)
    assert one.shape == (1, 3, 64, 64)
    assert torch.equal(((one[0].permute(1, 2, 0) * 0.5 + 0.5).clamp(0, 1) * 255).to(torch.uint8), u8[1])


def test_attention_is_one_head_of_full_width():
    """softmax(q k^T / sqrt(C)) v with the V bias moved behind P V (what the device path does) is the same function."""
    a = V.Attention(64, 32, 1e-6)
    x = torch.randn(1, 64, 4, 4, generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        want = a(x)
        t = a.group_norm(x).view(1, 64, 16).transpose(1, 2)
        p = torch.softmax(a.to_q(t) @ a.to_k(t).transpose(1, 2) / 8.0, dim=-1)
        vt = a.to_v.weight @ t[0].t()                                   # V^T without bias
        o = p[0] @ vt.t() + a.to_v.bias
        got = x + a.to_out[0](o).t().reshape(1, 64, 4, 4)
    assert float((got - want).abs().max()) < 1e-5


def test_flop_count_matches_module_walk():
    from b2v import vae
    for cfg, (h, w) in ((V.TINY, (8, 8)), (V.SD_VAE, (8, 16))):
        m = V.AutoencoderKLDecoder(cfg).eval()
        total = [0.0]

        def hook(mod, inp, out):
            if isinstance(mod, torch.nn.Conv2d):
                total[0] += 2.0 * out.numel() * mod.in_channels * mod.kernel_size[0] * mod.kernel_size[1]
            elif isinstance(mod, torch.nn.Linear):
                total[0] += 2.0 * out.numel() * mod.in_features

        hs = [mod.register_forward_hook(hook) for mod in m.modules()]
        with torch.no_grad():
            m.decode(torch.zeros(1, 4, h, w))
        for x in hs:
            x.remove()
        c = list(reversed(cfg.block_out_channels))[0]
        total[0] += 4.0 * (h * w) ** 2 * c  # Q K^T and P V
        obj = SimpleNamespace(config=SimpleNamespace(**cfg.to_dict()))
        got = vae.AutoencoderKL.flops(obj, 1, h, w)
        assert math.isclose(got, total[0], rel_tol=1e-9), (got, total[0])


def test_legacy_checkpoint_key_names_are_accepted():
    """Older diffusers checkpoints name the VAE attention projections query/key/value/proj_attn; they map onto the
    current names, encoder tensors are ignored, everything else passes through untouched."""
    from b2v import vae
    sd = V.AutoencoderKLDecoder(V.TINY).state_dict()
    legacy = {}
    for k, v in sd.items():
        for new, old in (("to_q", "query"), ("to_k", "key"), ("to_v", "value"), ("to_out.0", "proj_attn")):
            if ".attentions.0." + new + "." in k:
                k = k.replace("." + new + ".", "." + old + ".")
        legacy[k] = v
    assert any(".query." in k for k in legacy) and not any(".to_q." in k for k in legacy)
    legacy["encoder.conv_in.weight"] = torch.zeros(1)
    legacy["quant_conv.weight"] = torch.zeros(1)
    back = vae.canonical_keys(legacy)
    assert set(back) == set(sd) and all(back[k] is sd[k] for k in sd)
    assert vae.canonical_keys(sd).keys() == sd.keys()   # idempotent on current names


def test_weight_repack_host_logic():
    """AutoencoderKL.load_state_dict is torch-only plumbing (permute / pad / cast), so it can be driven on the CPU with a
    stand-in object: padded conv_in / conv_out / post_quant_conv tiles, `/scaling_factor` folded into the second
    post_quant copy, and a legacy-named checkpoint with 1x1-conv attention projections repacks to the same tensors."""
    from b2v import vae
    from oracle.unet3d_ref import random_init_
    ref = random_init_(V.AutoencoderKLDecoder(V.TINY), seed=1, round_bf16=True)
    cfg = SimpleNamespace(**{**vae.SD_VAE_CONFIG, **V.TINY.to_dict()})
    a = SimpleNamespace(config=cfg, device=torch.device("cpu"), _w=None)
    vae.AutoencoderKL.load_state_dict(a, ref.state_dict())
    assert a.n_params == V.param_count(V.TINY)
    w_in, _ = a._w["decoder.conv_in"]
    assert w_in.shape == (128, 3, 3, 64) and w_in.dtype == torch.bfloat16 and float(w_in[..., 4:].abs().max()) == 0
    assert torch.equal(w_in[..., :4].float(), ref.decoder.conv_in.weight.permute(0, 2, 3, 1))
    w_out, b_out = a._w["decoder.conv_out"]
    assert w_out.shape == (64, 3, 3, 64) and float(w_out[3:].abs().max()) == 0 and float(b_out[3:].abs().max()) == 0
    pq, pqb = a._w["pq"]
    pqs, _ = a._w["pq_scaled"]
    assert pq.shape == (64, 64) and float(pq[4:].abs().max()) == 0 and float(pq[:, 4:].abs().max()) == 0
    want = (ref.post_quant_conv.weight.view(4, 4) / 0.18215).to(torch.bfloat16)
    assert torch.equal(pqs[:4, :4], want) and torch.equal(pqb[:4], ref.post_quant_conv.bias)
    legacy = {}
    for k, v in ref.state_dict().items():
        for new, old in (("to_q", "query"), ("to_k", "key"), ("to_v", "value"), ("to_out.0", "proj_attn")):
            if ".attentions.0." + new + "." in k:
                k = k.replace("." + new + ".", "." + old + ".")
                if k.endswith("weight"):
                    v = v[:, :, None, None]
        legacy[k] = v
    b = SimpleNamespace(config=cfg, device=torch.device("cpu"), _w=None)
    vae.AutoencoderKL.load_state_dict(b, legacy)
    assert all(torch.equal(a._w[k][0], b._w[k][0]) and torch.equal(a._w[k][1], b._w[k][1]) for k in a._w)
    with __import__("pytest").raises(vae._lib.B2VError):
        vae.AutoencoderKL.load_state_dict(b, {k: v for k, v in legacy.items() if "conv_out" not in k})
