"""DDIMScheduler with the call surface the reference's hot loop uses on diffusers' scheduler
(Distribution/strategies/chunk_only.py:55,66-73; InferNet/neurons/miner.py:531-586):

    sched.set_timesteps(n, device=...)      sched.timesteps (iterable of 0-dim int64 tensors)
    sched.scale_model_input(x, t)           sched.step(eps, t, lat).prev_sample
    sched.init_noise_sigma                  sched.alphas_cumprod[t]

Host side = the integer/scalar schedule (betas, cumulative products, timestep spacing; [UPSTREAM diffusers]
scheduling_ddim.DDIMScheduler, eta = 0, epsilon prediction, "leading" spacing). Device side = ONE fused kernel per
step (b2v_cfg_ddim_step): ``step`` launches it without CFG; the strategies call ``fused_cfg_step`` which folds
the CFG combine, the DDIM update, the next UNet input (+ context injection) and the fp16 leaf images into that launch.
CUDA tensors only — there is no CPU path.
"""
from types import SimpleNamespace

import numpy as np
import torch

from . import ops

# Zeroscope / ModelScope-T2V scheduler_config.json
PIPELINE_CONFIG = dict(num_train_timesteps=1000, beta_start=0.00085, beta_end=0.012, beta_schedule="scaled_linear",
                       clip_sample=False, set_alpha_to_one=False, steps_offset=1)
# bare DDIMScheduler() as constructed by the verifier (proof.py:350, :492)
DEFAULT_CONFIG = dict(num_train_timesteps=1000, beta_start=1e-4, beta_end=0.02, beta_schedule="linear",
                      clip_sample=True, set_alpha_to_one=True, steps_offset=0)


class DDIMSchedulerOutput:
    __slots__ = ("prev_sample",)

    def __init__(self, prev_sample):
        self.prev_sample = prev_sample


class DDIMScheduler:
    def __init__(self, **config):
        cfg = dict(DEFAULT_CONFIG)
        cfg.update(config)
        self.config = SimpleNamespace(**cfg)
        n = cfg["num_train_timesteps"]
        # fp32 on purpose: diffusers builds the schedule with torch.float32 and the verifier's booleans depend on it
        if cfg["beta_schedule"] == "linear":
            betas = torch.linspace(cfg["beta_start"], cfg["beta_end"], n, dtype=torch.float32)
        elif cfg["beta_schedule"] == "scaled_linear":
            betas = torch.linspace(cfg["beta_start"] ** 0.5, cfg["beta_end"] ** 0.5, n, dtype=torch.float32) ** 2
        else:
            raise NotImplementedError("beta_schedule %r" % cfg["beta_schedule"])
        self.betas = betas
        self.alphas = 1.0 - betas
        self.alphas_cumprod = torch.cumprod(self.alphas, dim=0)
        self.final_alpha_cumprod = torch.tensor(1.0) if cfg["set_alpha_to_one"] else self.alphas_cumprod[0]
        self.init_noise_sigma = 1.0
        self.num_inference_steps = None
        self.timesteps = torch.from_numpy(np.arange(0, n)[::-1].copy().astype(np.int64))

    @classmethod
    def from_pipeline_config(cls):
        return cls(**PIPELINE_CONFIG)

    def scale_model_input(self, sample, timestep=None):
        return sample

    def set_timesteps(self, num_inference_steps, device=None):
        n = self.config.num_train_timesteps
        if num_inference_steps > n:
            raise ValueError("`num_inference_steps`: %d cannot be larger than %d" % (num_inference_steps, n))
        self.num_inference_steps = int(num_inference_steps)
        ratio = n // self.num_inference_steps
        ts = (np.arange(0, self.num_inference_steps) * ratio).round()[::-1].copy().astype(np.int64)
        ts += self.config.steps_offset
        # kept on the host: the loop reads them as Python ints, never through a device sync
        self._timesteps_host = [int(t) for t in ts]
        self.timesteps = torch.from_numpy(ts)
        if device is not None and torch.device(device).type == "cuda" and torch.cuda.is_available():
            self.timesteps = self.timesteps.to(device)

    def host_timesteps(self):
        return list(self._timesteps_host)

    def alpha_pair(self, timestep):
        """(alpha_cumprod[t], alpha_cumprod[t_prev]) as Python floats (fp32 values)."""
        if self.num_inference_steps is None:
            raise ValueError("Number of inference steps is 'None', you need to run 'set_timesteps' first")
        t = int(timestep)
        prev_t = t - self.config.num_train_timesteps // self.num_inference_steps
        a_prev = self.alphas_cumprod[prev_t] if prev_t >= 0 else self.final_alpha_cumprod
        return float(self.alphas_cumprod[t]), float(a_prev)

    def step(self, model_output, timestep, sample, eta=0.0, **_ignored):
        if eta != 0.0:
            raise NotImplementedError("only eta = 0 (the reference's hot path) is implemented")
        a_t, a_prev = self.alpha_pair(timestep)
        lat = sample if sample.dim() == 5 else sample.reshape((1,) + tuple(sample.shape[-4:]))
        eps = model_output.reshape(lat.shape).to(lat.dtype)
        out, _ = ops.cfg_ddim_step(lat.contiguous(), eps.contiguous(), a_t, a_prev,
                                   clip_sample=self.config.clip_sample)
        return DDIMSchedulerOutput(out.reshape(sample.shape))

    def fused_cfg_step(self, lat, eps2, timestep, guidance_scale, ctx=None, ctx_w=0.0, n_dup=2, leaf_z=None,
                       leaf_eps=None, out=None, xnext=None):
        """CFG combine + DDIM update + next UNet input in one launch. Returns (lat', x_next)."""
        a_t, a_prev = self.alpha_pair(timestep)
        return ops.cfg_ddim_step(lat, eps2, a_t, a_prev, guidance_scale=guidance_scale, ctx=ctx, ctx_w=ctx_w,
                                 clip_sample=self.config.clip_sample, make_next_input=n_dup, leaf_z=leaf_z,
                                 leaf_eps=leaf_eps, out=out, xnext=xnext)
