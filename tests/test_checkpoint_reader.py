"""not-gpu: reading a diffusers-layout checkpoint directory from local disk (b2v.checkpoint, the host half of
``VideoPipeline.from_pretrained``; chunk_only.py:44-56). Files are written by the test from the oracle's modules."""
import json
import os

import pytest
import torch

from b2v import checkpoint
from oracle import unet3d_ref as R
from oracle import vae_ref as V


def _write(tmp_path, shard=False, fmt="safetensors"):
    from safetensors.torch import save_file
    unet = R.random_init_(R.UNet3DConditionModel(R.TINY), seed=3)
    vae = R.random_init_(V.AutoencoderKLDecoder(V.TINY), seed=4)
    usd = {k: v.half().contiguous() for k, v in unet.state_dict().items()}
    vsd = {k: v.half().contiguous() for k, v in vae.state_dict().items()}
    vsd["encoder.conv_in.weight"] = torch.zeros(2, 2)
    for name, sd, cfg in (("unet", usd, dict(R.TINY.to_dict(), _class_name="UNet3DConditionModel")),
                          ("vae", vsd, dict(V.TINY.to_dict(), _class_name="AutoencoderKL"))):
        d = tmp_path / name
        d.mkdir()
        cfg = {k: (list(v) if isinstance(v, tuple) else v) for k, v in cfg.items()}
        (d / "config.json").write_text(json.dumps(cfg))
        if fmt == "bin":
            torch.save(sd, d / "diffusion_pytorch_model.bin")
        elif shard and name == "unet":
            keys = sorted(sd)
            half = len(keys) // 2
            parts = {"diffusion_pytorch_model-00001-of-00002.safetensors": keys[:half],
                     "diffusion_pytorch_model-00002-of-00002.safetensors": keys[half:]}
            wm = {}
            for f, ks in parts.items():
                save_file({k: sd[k] for k in ks}, str(d / f))
                wm.update({k: f for k in ks})
            (d / "diffusion_pytorch_model.safetensors.index.json").write_text(json.dumps({"weight_map": wm}))
        else:
            save_file(sd, str(d / "diffusion_pytorch_model.fp16.safetensors"))
    (tmp_path / "scheduler").mkdir()
    (tmp_path / "scheduler" / "scheduler_config.json").write_text(json.dumps(
        {"_class_name": "DDIMScheduler", "num_train_timesteps": 1000, "beta_start": 0.00085, "beta_end": 0.012,
         "beta_schedule": "scaled_linear", "clip_sample": False, "set_alpha_to_one": False, "steps_offset": 1,
         "prediction_type": "epsilon", "trained_betas": None}))
    return usd, vsd


@pytest.mark.parametrize("shard,fmt", [(False, "safetensors"), (True, "safetensors"), (False, "bin")])
def test_read_components(tmp_path, shard, fmt):
    usd, vsd = _write(tmp_path, shard, fmt)
    sd, cfg = checkpoint.read_component(str(tmp_path), "unet")
    assert set(sd) == set(usd) and all(torch.equal(sd[k], usd[k]) for k in usd)
    ucfg = checkpoint.unet_config(cfg)
    assert ucfg["block_out_channels"] == (64, 128, 256, 256) and ucfg["cross_attention_dim"] == 128
    assert "_class_name" not in ucfg and "sample_size" not in ucfg
    assert ucfg["transformer_in_heads"] == 2   # repo-specific key of the narrow test config; absent in real checkpoints
    from b2v.unet import ZEROSCOPE_CONFIG
    real = {k: v for k, v in cfg.items() if k != "transformer_in_heads"}
    assert {**ZEROSCOPE_CONFIG, **checkpoint.unet_config(real)}["transformer_in_heads"] == 8
    vs, vcfg = checkpoint.read_component(str(tmp_path), "vae")
    assert set(vs) == set(vsd)
    assert checkpoint.vae_config(vcfg) == {"latent_channels": 4, "out_channels": 3, "layers_per_block": 2,
                                           "norm_num_groups": 32, "scaling_factor": 0.18215,
                                           "block_out_channels": (64, 64, 128, 128)}
    from b2v.vae import canonical_keys, parameter_shapes
    assert set(canonical_keys(vs)) == set(parameter_shapes(V.TINY.to_dict()))    # encoder tensors dropped
    sc = checkpoint.scheduler_config(str(tmp_path))
    from b2v.scheduler import PIPELINE_CONFIG
    assert sc == PIPELINE_CONFIG


def test_reader_errors(tmp_path):
    with pytest.raises(FileNotFoundError):
        checkpoint.read_component(str(tmp_path), "unet")
    (tmp_path / "unet").mkdir()
    with pytest.raises(FileNotFoundError):
        checkpoint.read_component(str(tmp_path), "unet")
    assert checkpoint.scheduler_config(str(tmp_path)) == {}
    (tmp_path / "scheduler").mkdir()
    (tmp_path / "scheduler" / "scheduler_config.json").write_text(json.dumps({"prediction_type": "v_prediction"}))
    with pytest.raises(NotImplementedError):
        checkpoint.scheduler_config(str(tmp_path))
    from b2v.pipeline import VideoPipeline
    with pytest.raises(FileNotFoundError):
        VideoPipeline.from_pretrained(os.path.join(str(tmp_path), "cerspense/zeroscope_v2_XL"))
