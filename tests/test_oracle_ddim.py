"""not-gpu: the DDIM scheduler restatement (oracle/ddim_ref.py) against an independent float64 evaluation of the
closed-form eta=0 update, the values SURVEY.md quotes, and the product's host-side schedule (b2v.scheduler)."""
import numpy as np
import pytest
import torch

import b2v  # noqa: F401
from b2v.scheduler import DDIMScheduler
from oracle.ddim_ref import DDIMSchedulerRef, default_scheduler, pipeline_scheduler


def test_timesteps_quoted_in_survey():
    s = pipeline_scheduler()
    s.set_timesteps(4)
    assert s.timesteps.tolist() == [751, 501, 251, 1]
    d = default_scheduler()
    d.set_timesteps(8)
    assert d.timesteps.tolist() == [875, 750, 625, 500, 375, 250, 125, 0]   # Appendix C
    d.set_timesteps(50)
    assert d.timesteps[0] == 980 and d.timesteps[-1] == 0
    assert s.init_noise_sigma == 1.0 and s.scale_model_input(torch.ones(2)) is not None


@pytest.mark.parametrize("make,clip", [(pipeline_scheduler, False), (default_scheduler, True)])
def test_step_closed_form_float64(make, clip):
    s = make()
    s.set_timesteps(10)
    g = torch.Generator().manual_seed(0)
    x, e = torch.randn(2, 4, 3, 5, 5, generator=g, dtype=torch.float64), torch.randn(2, 4, 3, 5, 5, generator=g, dtype=torch.float64)
    n = s.config.num_train_timesteps
    betas = (np.linspace(s.config.beta_start ** 0.5, s.config.beta_end ** 0.5, n) ** 2
             if s.config.beta_schedule == "scaled_linear" else np.linspace(s.config.beta_start, s.config.beta_end, n))
    ac = np.cumprod(1 - betas)
    for t in s.timesteps.tolist():
        tp = t - n // 10
        a_t = ac[t]
        a_p = ac[tp] if tp >= 0 else (1.0 if s.config.set_alpha_to_one else ac[0])
        x0 = (x - (1 - a_t) ** 0.5 * e) / a_t ** 0.5
        if clip:
            x0 = x0.clamp(-1, 1)
        want = a_p ** 0.5 * x0 + (1 - a_p) ** 0.5 * e
        got = s.step(e, t, x).prev_sample
        assert torch.allclose(got, want, rtol=1e-5, atol=1e-5)


def test_fp16_stays_fp16():
    s = default_scheduler()
    s.set_timesteps(8)
    z = torch.randn(4, 8, 8).half()
    out = s.step(z, s.timesteps[2], z).prev_sample
    assert out.dtype == torch.float16


def test_product_schedule_equals_oracle():
    for mine, ref in ((DDIMScheduler.from_pipeline_config(), pipeline_scheduler()), (DDIMScheduler(), default_scheduler())):
        assert torch.equal(mine.alphas_cumprod, ref.alphas_cumprod)
        for n in (4, 8, 25, 50, 500):
            mine.set_timesteps(n)
            ref.set_timesteps(n)
            assert mine.host_timesteps() == ref.timesteps.tolist() == mine.timesteps.tolist()
            for t in mine.host_timesteps()[:3] + mine.host_timesteps()[-2:]:
                assert mine.alpha_pair(t) == ref.alpha_pair(t)
    with pytest.raises(ValueError):
        DDIMScheduler().set_timesteps(1001)
    with pytest.raises(ValueError):
        DDIMScheduler().alpha_pair(5)
