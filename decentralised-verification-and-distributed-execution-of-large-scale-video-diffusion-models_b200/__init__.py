"""B200-native (sm_100a) implementation of the chunked-coherent Zeroscope denoising loop of the reference's
``Distribution/strategies`` and of the Proof-of-Inference Merkle commit / verify of
``InferNet/template/validator/proof.py``, behind the reference's own Python call signatures.

Layout
  csrc/            hand-written CUDA kernels + the C-ABI (include/b2v.h) -> libb2v.so (built in-tree by build.py)
  _lib.py, ops.py  ctypes binding; torch supplies device memory and streams only
  scheduler.py     DDIMScheduler with the diffusers call surface the strategies/miner use
  unet.py          UNet3DConditionModel: pipe.unet(x, t, encoder_hidden_states=emb).sample
  vae.py           AutoencoderKL: pipe.vae.decode(z / 0.18215).sample and decode_frames (latents -> uint8 frames)
  pipeline.py      pipe-shaped object (unet, scheduler, vae, synthetic text encoder); from_pretrained(local dir)
  checkpoint.py    reader of diffusers-layout checkpoint folders (safetensors / sharded / .bin), host only
  planner.py       chunk planner / blend weights (host integer logic of chunk_only.py:80-105, 136-150)
  strategies/      chunk_only / fsdp_chunked_coherent / fsdp_chunked / fsdp entry points (same CLI + CSV)
  validator.py     spot-check round over many miners (validator.py:641-845): text in, four launches
  proof.py         derive_seed, build_merkle_root, verify_merkle_leaf, run_unet_step, ... (same signatures)

There is no CPU fallback: compute entry points raise when libb2v.so or an sm_100 device is missing.
The directory name is not a Python identifier; ``import b2v`` (repo-root shim) aliases this package.
"""
from . import _lib  # noqa: F401
from ._lib import B2VError, EXPORTED_SYMBOLS  # noqa: F401

__all__ = ["_lib", "ops", "scheduler", "unet", "vae", "pipeline", "planner", "proof", "validator", "miner", "exchange", "strategies", "B2VError"]
