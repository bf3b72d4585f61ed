"""-m gpu: ``VideoPipeline.from_pretrained`` on a local diffusers-layout folder (chunk_only.py:44-52:
``TextToVideoSDPipeline.from_pretrained(model_id, torch_dtype=torch.float16)``). The folder is written by the test
(safetensors, no network); the loaded pipeline must equal ``from_state_dict`` on the same tensors BIT FOR BIT (same
repack, same kernels), and the real-weights + synthetic-text combination must be refused unless asked for."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _write_folder(tmp_path, fmt="safetensors", shard=False):
    from test_checkpoint_reader import _write
    return _write(tmp_path, shard=shard, fmt=fmt)


@pytest.mark.parametrize("fmt,shard,dtype", [("safetensors", False, torch.float16), ("safetensors", True, torch.bfloat16),
                                             ("bin", False, torch.float16)])
def test_from_pretrained_equals_from_state_dict(cuda, tmp_path, fmt, shard, dtype):
    from b2v.pipeline import VideoPipeline
    from oracle import unet3d_ref as R
    usd, vsd = _write_folder(tmp_path, fmt, shard)
    with pytest.raises(ValueError):     # real weights + stand-in text tower: must be asked for
        VideoPipeline.from_pretrained(str(tmp_path), device=cuda, torch_dtype=dtype)
    pipe = VideoPipeline.from_pretrained(str(tmp_path), device=cuda, torch_dtype=dtype, allow_synthetic_text=True)
    assert pipe.synthetic_text and pipe.vae is not None and pipe.dtype == dtype
    want = VideoPipeline.from_state_dict(usd, config=R.TINY.to_dict(), device=cuda, dtype=dtype)
    assert pipe.unet.param_count() == want.unet.param_count() == sum(v.numel() for v in usd.values())
    g = torch.Generator().manual_seed(11)
    x = torch.randn(2, 4, 3, 16, 16, generator=g).to(cuda, dtype)
    ehs = torch.randn(2, 77, R.TINY.cross_attention_dim, generator=g).to(cuda, dtype)
    a = pipe.unet(x, 501, encoder_hidden_states=ehs).sample
    b = want.unet(x, 501, encoder_hidden_states=ehs).sample
    assert torch.isfinite(a).all() and torch.equal(a, b)
    # the scheduler config read from disk is the pipeline's (U9)
    pipe.scheduler.set_timesteps(4, device="cuda")
    assert pipe.scheduler.host_timesteps() == [751, 501, 251, 1]
    # the VAE read from the same folder decodes (N1) and equals a decoder built from the same tensors
    from b2v.vae import AutoencoderKL
    from oracle import vae_ref as V
    ref_vae = AutoencoderKL(V.TINY.to_dict(), device=cuda).load_state_dict(vsd)
    z = (torch.randn(1, 4, 2, 8, 8, generator=g) * 0.3).to(cuda)
    assert torch.equal(pipe.vae.decode_frames(z), ref_vae.decode_frames(z))


def test_strategy_cli_accepts_checkpoint_folder(cuda, tmp_path):
    """``--model_id <folder>`` through the strategy entry point (chunk_only.py:224-281) with --synthetic_text."""
    from b2v.strategies import chunk_only
    _write_folder(tmp_path)
    out_csv = os.path.join(str(tmp_path), "bench.csv")
    cwd = os.getcwd()
    os.chdir(str(tmp_path))
    try:
        chunk_only.main(["--model_id", str(tmp_path), "--synthetic_text", "--num_frames", "4", "--steps", "2",
                         "--height", "64", "--width", "64", "--out_csv", out_csv, "--no_decode"])
    finally:
        os.chdir(cwd)
    rows = open(out_csv).read().strip().splitlines()
    assert len(rows) == 2 and rows[0].startswith("timestamp")
