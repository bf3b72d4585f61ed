"""ORACLE — test infrastructure only (tests/, __graft_entry__.smoke(), bench.py's CPU baseline / reference arm).

CPU restatement of the two DDIM schedulers the reference's hot path touches. The arithmetic lives in un-vendored,
unpinned ``diffusers`` (InferNet/requirements.txt:10 ``diffusers>=0.19.0``; absent from this image), so this file
restates the published algorithm of ``diffusers.schedulers.scheduling_ddim.DDIMScheduler`` (eta = 0,
epsilon-prediction, "leading" timestep spacing):

  * PIPELINE config — what ``pipe.scheduler`` is in Distribution/strategies/chunk_only.py:55,68,73 and
    InferNet/neurons/miner.py:531-586 (Zeroscope ``scheduler_config.json``: scaled_linear 0.00085..0.012,
    steps_offset 1, clip_sample False, set_alpha_to_one False)                         SURVEY.md §8a U9
  * DEFAULT config  — the bare ``DDIMScheduler()`` the verifier builds in InferNet/template/validator/proof.py:350-364
    and :492-495 (linear 1e-4..0.02, steps_offset 0, clip_sample True, set_alpha_to_one True)   SURVEY.md Appendix C

PARITY UNPINNED for the scheduler itself (no reference test pins it and diffusers cannot be imported here); the
closed-form DDIM(eta=0) update is additionally checked in tests/test_oracle_ddim.py against an independent float64
evaluation of x' = sqrt(a_p) * (x - sqrt(1-a_t) e)/sqrt(a_t) + sqrt(1-a_p) e.

Dtype behaviour follows torch's promotion rules exactly because the code below *is* torch: scheduler scalars are
0-dim fp32 tensors, so an fp16 sample stays fp16 with one rounding per elementwise op (what the verifier sees on CPU).
"""
from types import SimpleNamespace

import numpy as np
import torch

PIPELINE_CONFIG = dict(num_train_timesteps=1000, beta_start=0.00085, beta_end=0.012, beta_schedule="scaled_linear",
                       clip_sample=False, set_alpha_to_one=False, steps_offset=1, clip_sample_range=1.0)
DEFAULT_CONFIG = dict(num_train_timesteps=1000, beta_start=1e-4, beta_end=0.02, beta_schedule="linear",
                      clip_sample=True, set_alpha_to_one=True, steps_offset=0, clip_sample_range=1.0)


class DDIMSchedulerRef:
    def __init__(self, **overrides):
        cfg = dict(DEFAULT_CONFIG)
        cfg.update(overrides)
        self.config = SimpleNamespace(**cfg)
        n = cfg["num_train_timesteps"]
        if cfg["beta_schedule"] == "linear":
            betas = torch.linspace(cfg["beta_start"], cfg["beta_end"], n, dtype=torch.float32)
        elif cfg["beta_schedule"] == "scaled_linear":
            betas = torch.linspace(cfg["beta_start"] ** 0.5, cfg["beta_end"] ** 0.5, n, dtype=torch.float32) ** 2
        else:
            raise NotImplementedError(cfg["beta_schedule"])
        self.betas = betas
        self.alphas = 1.0 - betas
        self.alphas_cumprod = torch.cumprod(self.alphas, dim=0)
        self.final_alpha_cumprod = torch.tensor(1.0) if cfg["set_alpha_to_one"] else self.alphas_cumprod[0]
        self.init_noise_sigma = 1.0
        self.num_inference_steps = None
        self.timesteps = torch.from_numpy(np.arange(0, n)[::-1].copy().astype(np.int64))

    def scale_model_input(self, sample, timestep=None):
        return sample

    def set_timesteps(self, num_inference_steps, device=None):
        n = self.config.num_train_timesteps
        if num_inference_steps > n:
            raise ValueError("num_inference_steps > num_train_timesteps")
        self.num_inference_steps = num_inference_steps
        ratio = n // num_inference_steps
        ts = (np.arange(0, num_inference_steps) * ratio).round()[::-1].copy().astype(np.int64)
        ts += self.config.steps_offset
        self.timesteps = torch.from_numpy(ts).to(device) if device is not None else torch.from_numpy(ts)

    def step(self, model_output, timestep, sample, eta=0.0):
        if self.num_inference_steps is None:
            raise ValueError("set_timesteps first")
        if eta != 0.0:
            raise NotImplementedError("the reference's hot path only uses eta = 0")
        t = int(timestep)
        prev_t = t - self.config.num_train_timesteps // self.num_inference_steps
        a_t = self.alphas_cumprod[t]
        a_prev = self.alphas_cumprod[prev_t] if prev_t >= 0 else self.final_alpha_cumprod
        beta_t = 1 - a_t
        x0 = (sample - beta_t ** 0.5 * model_output) / a_t ** 0.5
        if self.config.clip_sample:
            x0 = x0.clamp(-self.config.clip_sample_range, self.config.clip_sample_range)
        direction = (1 - a_prev) ** 0.5 * model_output
        prev_sample = a_prev ** 0.5 * x0 + direction
        return SimpleNamespace(prev_sample=prev_sample, pred_original_sample=x0)

    # helper used by the kernels' host side: the two scalars a step needs
    def alpha_pair(self, timestep):
        t = int(timestep)
        prev_t = t - self.config.num_train_timesteps // self.num_inference_steps
        a_prev = self.alphas_cumprod[prev_t] if prev_t >= 0 else self.final_alpha_cumprod
        return float(self.alphas_cumprod[t]), float(a_prev)


def pipeline_scheduler():
    return DDIMSchedulerRef(**PIPELINE_CONFIG)


def default_scheduler():
    return DDIMSchedulerRef()
