"""-m gpu: the frame decode after the denoise loop — `vae.decode(z / 0.18215).sample` and the uint8
frame conversion of chunk_only.py:152-158 — through the C-ABI, against the CPU fp32 oracle (oracle/vae_ref.py).

Byte work (frame quantisation, nearest upsample, latent staging) is bit-exact; the floating-point decoder is bf16
tensor-core arithmetic against fp32, tolerance written at each assert."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def rel_l2(a, b):
    a, b = a.float(), b.float()
    return float((a - b).norm() / b.norm().clamp_min(1e-12))


# ---- kernels ----------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("rows,cols", [(1, 8), (7, 64), (300, 2880), (128, 9216), (5, 16384)])
def test_softmax_rows(cuda, rows, cols):
    """fp32 softmax of the bf16 scores, one bf16 rounding at the store: |err| <= 2^-8 relative + 1e-7 absolute."""
    from b2v import ops
    g = torch.Generator(device="cuda").manual_seed(rows * 31 + cols)
    x = (6.0 * torch.randn(rows, cols, device=cuda, generator=g)).to(torch.bfloat16)
    scale = 1.0 / math.sqrt(512)
    want = torch.softmax(x.float() * scale, dim=-1)
    got = ops.softmax_rows_(x.clone(), scale).float()
    assert torch.isfinite(got).all()
    assert float(((got - want).abs() - want * 2.0 ** -8).max()) <= 1e-7
    assert float((got.sum(-1) - 1).abs().max()) <= 4e-3  # rows of P sum to one up to the bf16 rounding of its entries


def test_softmax_rows_strided_and_large_scores(cuda):
    from b2v import ops
    buf = torch.zeros(16, 256, dtype=torch.bfloat16, device=cuda)
    view = buf[:, :128]
    view.copy_(torch.linspace(-300, 300, 16 * 128, device=cuda).view(16, 128))
    ref = torch.softmax(view.float() * 0.5, dim=-1)
    from b2v._lib import check, load
    check(load().b2v_op_softmax_rows(ops.ptr(buf), 16, 128, 256, 0.5, ops.stream()))
    assert torch.isfinite(buf.float()).all()
    assert float((buf[:, :128].float() - ref).abs().max()) <= 2.0 ** -8
    assert float(buf[:, 128:].float().abs().max()) == 0.0  # columns beyond `cols` are untouched


@pytest.mark.parametrize("n_px", [1, 3, 4, 5, 1023, 64 * 64, 576 * 1024])
def test_frames_u8_bit_exact(cuda, n_px):
    """(v*0.5+0.5).clamp(0,1)*255 truncated == torch's `.byte()` on the same bf16 input, every byte."""
    from b2v import ops
    g = torch.Generator(device="cuda").manual_seed(n_px)
    rows = (1.2 * torch.randn(n_px, 64, device=cuda, generator=g)).to(torch.bfloat16)
    rows[0, :3] = torch.tensor([-1.0, 1.0, 0.0])
    want = ((rows[:, :3].float() * 0.5 + 0.5).clamp(0, 1) * 255).to(torch.uint8)
    got = ops.frames_u8(rows, n_px)
    assert got.dtype == torch.uint8 and got.shape == (n_px, 3)
    assert torch.equal(got, want)
    if n_px > 100:
        assert int(got.max()) == 255 and int(got.min()) == 0


def test_upsample_and_latent_rows_bit_exact(cuda):
    from b2v import ops
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.randn(2, 5, 7, 64, device=cuda, generator=g).to(torch.bfloat16)
    want = x.repeat_interleave(2, dim=1).repeat_interleave(2, dim=2)
    assert torch.equal(ops.upsample2x(x), want)
    for dt in (torch.float32, torch.float16, torch.bfloat16):
        lat = torch.randn(1, 4, 3, 8, 8, device=cuda, generator=g).to(dt)
        rows = ops.latent_rows(lat, 64)
        want = lat[0].permute(1, 2, 3, 0).reshape(-1, 4).to(torch.bfloat16)
        assert torch.equal(rows[:, :4], want) and float(rows[:, 4:].float().abs().max()) == 0.0


# ---- decoder vs oracle --------------------------------------------------------------------------------------------
def _pair(cuda, cfg, seed=4321):
    from oracle import vae_ref as V
    from oracle.unet3d_ref import random_init_
    from b2v.vae import AutoencoderKL
    ref = random_init_(V.AutoencoderKLDecoder(cfg), seed=seed, round_bf16=True).eval()
    net = AutoencoderKL(cfg.to_dict(), device=cuda).load_state_dict(ref.state_dict())
    return ref, net


def _compare(out, want, tag):
    out, want = out.float().cpu(), want.float()
    cos = float(torch.nn.functional.cosine_similarity(out.flatten(), want.flatten(), dim=0))
    rel = rel_l2(out, want)
    print(tag, "cos", cos, "rel-l2", rel, "max-abs", float((out - want).abs().max()), "max|ref|", float(want.abs().max()))
    return cos, rel


@pytest.mark.parametrize("n,h,w", [(1, 8, 8), (3, 16, 8), (2, 16, 24)])
def test_tiny_decoder_vs_oracle(cuda, n, h, w):
    """Narrow decoder (same topology): bf16 kernels vs the fp32 oracle; cos >= 0.9995, rel-l2 <= 2.5e-2."""
    from oracle import vae_ref as V
    ref, net = _pair(cuda, V.TINY)
    z = torch.randn(n, 4, h, w, generator=torch.Generator().manual_seed(n * h + w))
    with torch.no_grad():
        want = ref.decode(z)
    got = net.decode(z.to(cuda)).sample
    assert got.shape == (n, 3, 8 * h, 8 * w) and got.dtype == torch.float32
    cos, rel = _compare(got, want, "tiny vae %s" % ((n, h, w),))
    assert cos >= 0.9995 and rel <= 2.5e-2, (cos, rel)


def test_decoder_refuses_frames_off_the_attention_tile(cuda):
    """A 20x20 latent (400 rows per frame: not a multiple of 64, nor of the 32-row blocks the GroupNorm statistics
    epilogue is folded in) is refused up front by decode / decode_frames with a ValueError that names the rule, not by a
    B2V_CHECK deep inside a norm — also when 16 frames make n*h*w a whole number of 256-row GEMM tiles."""
    from oracle import vae_ref as V
    _, net = _pair(cuda, V.TINY)
    z = torch.randn(16, 4, 20, 20, device=cuda)
    with pytest.raises(ValueError, match="multiple of 64"):
        net.decode(z)
    with pytest.raises(ValueError, match="multiple of 64"):
        net.decode_frames(z.permute(1, 0, 2, 3).unsqueeze(0).contiguous())
    assert not net._stats_ok(16 * 400, 128, 400) and net._stats_ok(16 * 1600, 128, 1600)


def test_full_width_decoder_vs_oracle(cuda):
    """The SD VAE decoder at full width (49 490 199 parameters, the published decoder + post_quant_conv) on a 16 x 16
    latent: cos >= 0.9995, rel-l2 <= 2.5e-2 against fp32."""
    from oracle import vae_ref as V
    ref, net = _pair(cuda, V.SD_VAE)
    assert net.n_params == V.param_count() == 49490199
    z = torch.randn(2, 4, 16, 16, generator=torch.Generator().manual_seed(8)) * 3.0
    with torch.no_grad():
        want = ref.decode(z)
    got = net.decode(z.to(cuda)).sample
    cos, rel = _compare(got, want, "sd vae (2,16,16)")
    assert cos >= 0.9995 and rel <= 2.5e-2, (cos, rel)


def test_decode_frames_vs_oracle(cuda):
    """The reference's whole loop (chunk_only.py:152-158): latents [1,4,T,h,w] -> uint8 [T,H,W,3]. Against the oracle's
    bytes: the decoder deviates by bf16 arithmetic, so bytes may move by a few counts: mean |diff| <= 1.5 counts,
    99 % of the bytes within 6 counts, none further than 24 (of 255)."""
    from oracle import vae_ref as V
    ref, net = _pair(cuda, V.SD_VAE, seed=77)
    lat = torch.randn(1, 4, 3, 8, 16, generator=torch.Generator().manual_seed(2)) * 0.18215 * 2.0
    want_u8, want_raw = V.frames_u8(ref, lat)
    got = net.decode_frames(lat.to(cuda))
    assert got.dtype == torch.uint8 and got.shape == (3, 64, 128, 3) and got.is_cuda
    d = (got.cpu().int() - want_u8.int()).abs()
    print("decode_frames |diff| mean", float(d.float().mean()), "p99", float(d.float().quantile(0.99)), "max", int(d.max()),
          "oracle range", float(want_raw.min()), float(want_raw.max()))
    assert float(want_raw.max()) - float(want_raw.min()) > 100  # the comparison is not on a saturated image
    assert float(d.float().mean()) <= 1.5 and float(d.float().quantile(0.99)) <= 6 and int(d.max()) <= 24
    # half precision latents (what the reference feeds) give the same frames up to their own rounding
    got16 = net.decode_frames(lat.half().to(cuda))
    assert float((got16.int() - got.int()).abs().float().mean()) <= 1.5
    # and the reference-shaped per-frame call agrees with the batched one
    per = net.decode((lat[:, :, 1] / 0.18215).to(cuda)).sample
    u8 = ((per[0].permute(1, 2, 0) * 0.5 + 0.5).clamp(0, 1) * 255).to(torch.uint8)
    assert float((u8.int() - got[1].int()).abs().float().mean()) <= 1.5


def test_xl_frames_properties(cuda):
    """ZeroscopeXL output size (72 x 128 latent -> 576 x 1024 frames), full-width decoder: size-independent properties."""
    from b2v.vae import AutoencoderKL
    net = AutoencoderKL.from_random(device=cuda)
    lat = torch.randn(1, 4, 3, 72, 128, generator=torch.Generator().manual_seed(4)).to(cuda) * 0.3
    a = net.decode_frames(lat)
    assert a.shape == (3, 576, 1024, 3) and a.dtype == torch.uint8
    assert torch.equal(a, net.decode_frames(lat)), "decode must be bit-reproducible"
    # frames never mix: frame 1 decoded alone gives the same bytes up to the bf16 floor (GroupNorm partial sums fold in
    # another split order when the batch changes)
    b = net.decode_frames(lat[:, :, 1:2].contiguous())
    d = (a[1].int() - b[0].int()).abs().float()
    assert float(d.mean()) <= 1.0, float(d.mean())
    assert float((a[0].int() - a[1].int()).abs().float().mean()) > 5.0  # ... while different frames differ
    # frame batching is invisible
    net.frames_per_batch = 1
    c = net.decode_frames(lat)
    assert float((a.int() - c.int()).abs().float().mean()) <= 1.0
    assert int(a.max()) > int(a.min())


def test_strategy_decodes_and_writes_video(cuda, tmp_path, monkeypatch):
    """chunk_only.py:152-208 end to end at a small size: denoise -> blend -> decode_frames -> mp4 + boundary metrics,
    the frames equal to a direct decode of the blended latents."""
    import argparse
    import os
    from oracle import unet3d_ref as R
    from oracle import vae_ref as V
    from b2v.pipeline import VideoPipeline
    from b2v.scheduler import DDIMScheduler
    from b2v.strategies.chunk_only import ChunkedVideoDiffuser
    from b2v.unet import UNet3DConditionModel
    from b2v.vae import AutoencoderKL
    monkeypatch.chdir(tmp_path)
    ref = R.random_init_(R.UNet3DConditionModel(R.TINY), seed=1234, round_bf16=True).eval()
    net = UNet3DConditionModel(R.TINY.to_dict(), device=cuda).load_state_dict(ref.state_dict())
    vae = AutoencoderKL.from_random(V.TINY.to_dict(), device=cuda)
    pipe = VideoPipeline(net, DDIMScheduler.from_pipeline_config(), text_dim=128, dtype=torch.float32, vae=vae)
    cfg = argparse.Namespace(model_id="tiny", prompt="a rocket in space, 4k", num_frames=8, steps=2, guidance_scale=7.5,
                             chunk_size=5, overlap=1, fps=8, height=128, width=128, device="cuda",
                             out_csv=str(tmp_path / "m.csv"), emu_bw_mbps=0, emu_rtt_ms=0, emu_jitter_ms=0, seed=1234,
                             dtype="fp32", decode=True)
    d = ChunkedVideoDiffuser(cfg, pipe=pipe)
    res = d()
    assert os.path.getsize(tmp_path / "out.mp4") > 0
    assert res["temp_instab"] is not None and res["flow_err"] is not None
    frames = vae.decode_frames(d.latents)
    assert frames.shape == (8, 128, 128, 3)
    # reference quirk Q1 (chunk_only.py:140-147): frames covered only by ramp-zero weights have all-zero latents, so
    # they decode to the decoder's image of z = 0 — identical for every such frame
    assert float(d.latents[:, :, 0].abs().max()) == 0.0 and torch.equal(frames[0], frames[7])
