"""Reading a diffusers-layout checkpoint directory from local disk (the reference's
``TextToVideoSDPipeline.from_pretrained(model_id, torch_dtype=torch.float16)``, chunk_only.py:44-56 / miner.py:83-92):

    <dir>/unet/config.json + diffusion_pytorch_model[.fp16].safetensors | sharded *.safetensors + index | .bin
    <dir>/vae/...            (same file conventions)
    <dir>/scheduler/scheduler_config.json

Host file logic only — tensors come back on the CPU under their checkpoint names and are handed to
``UNet3DConditionModel.load_state_dict`` / ``AutoencoderKL.load_state_dict``, which repack them for the kernels.
There is no network here: ``model_id`` must be a local directory (a HF cache snapshot works as is)."""
import glob
import json
import os

_WEIGHT_NAMES = ("diffusion_pytorch_model.safetensors", "diffusion_pytorch_model.fp16.safetensors",
                 "diffusion_pytorch_model.bin", "diffusion_pytorch_model.fp16.bin")


def read_json(path):
    with open(path) as f:
        return json.load(f)


def _load_file(path):
    if path.endswith(".safetensors"):
        from safetensors.torch import load_file
        return load_file(path, device="cpu")
    import torch
    return torch.load(path, map_location="cpu", weights_only=True)


def read_component(root, name):
    """(state_dict on the CPU, config dict) of one pipeline component, e.g. ``read_component(dir, "unet")``."""
    d = os.path.join(root, name)
    if not os.path.isdir(d):
        raise FileNotFoundError("no %r component under %s" % (name, root))
    cfg_path = os.path.join(d, "config.json")
    config = read_json(cfg_path) if os.path.exists(cfg_path) else {}
    for w in _WEIGHT_NAMES:
        p = os.path.join(d, w)
        if os.path.exists(p):
            return _load_file(p), config
    index = glob.glob(os.path.join(d, "*.safetensors.index.json")) + glob.glob(os.path.join(d, "*.bin.index.json"))
    if index:
        files = sorted(set(read_json(index[0])["weight_map"].values()))
        sd = {}
        for f in files:
            sd.update(_load_file(os.path.join(d, f)))
        return sd, config
    raise FileNotFoundError("no weights (%s or a sharded index) in %s" % (", ".join(_WEIGHT_NAMES), d))


def scheduler_config(root):
    """scheduler/scheduler_config.json reduced to the keys b2v.scheduler.DDIMScheduler takes; {} when absent."""
    p = os.path.join(root, "scheduler", "scheduler_config.json")
    if not os.path.exists(p):
        return {}
    cfg = read_json(p)
    keys = ("num_train_timesteps", "beta_start", "beta_end", "beta_schedule", "clip_sample", "set_alpha_to_one",
            "steps_offset")
    if cfg.get("prediction_type", "epsilon") != "epsilon":
        raise NotImplementedError("prediction_type %r (the reference's pipelines are epsilon models)" % cfg["prediction_type"])
    return {k: cfg[k] for k in keys if k in cfg}


def unet_config(config):
    """diffusers UNet3DConditionModel config.json -> the keys b2v.unet.UNet3DConditionModel takes."""
    out = {}
    # transformer_in_heads is not a diffusers key (diffusers hard-codes 8 heads for transformer_in); the narrow test
    # configurations of this repo carry it, real checkpoints fall back to the default 8
    for k in ("in_channels", "out_channels", "layers_per_block", "cross_attention_dim", "attention_head_dim",
              "norm_num_groups", "norm_eps", "transformer_in_heads"):
        if k in config:
            out[k] = config[k]
    if "block_out_channels" in config:
        out["block_out_channels"] = tuple(config["block_out_channels"])
    if isinstance(out.get("attention_head_dim"), (list, tuple)):
        out["attention_head_dim"] = int(out["attention_head_dim"][0])
    return out


def vae_config(config):
    out = {}
    for k in ("latent_channels", "out_channels", "layers_per_block", "norm_num_groups", "scaling_factor"):
        if k in config:
            out[k] = config[k]
    if "block_out_channels" in config:
        out["block_out_channels"] = tuple(config["block_out_channels"])
    return out
