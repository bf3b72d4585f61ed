"""Test infrastructure (lives under tests/ because it runs the oracle). GPU box: error budget of the bf16 UNet path vs the fp32 CPU oracle on the TINY config: single forward, guided eps,
per-step latents for several guidance scales. Prints numbers used to state the tolerances in tests/."""
import sys
import torch
sys.path.insert(0, ".")
import b2v  # noqa
from b2v import ops
from b2v.unet import UNet3DConditionModel
from b2v.scheduler import DDIMScheduler
from oracle import unet3d_ref as R
from oracle.ddim_ref import pipeline_scheduler

dev = torch.device("cuda:0")
ref = R.random_init_(R.UNet3DConditionModel(R.TINY), seed=1234, round_bf16=True).eval()
net = UNet3DConditionModel(R.TINY.to_dict(), device=dev).load_state_dict(ref.state_dict())
g = torch.Generator().manual_seed(1)
cond, uncond = torch.randn(1, 77, 128, generator=g), torch.randn(1, 77, 128, generator=g)
emb = torch.cat([uncond, cond])
rl = lambda a, b: float((a.float().cpu() - b).norm() / b.norm())
for gs in (1.0, 3.0, 7.5):
    for steps in (4, 25):
        sr = pipeline_scheduler(); sr.set_timesteps(steps)
        sm = DDIMScheduler.from_pipeline_config(); sm.set_timesteps(steps)
        lat_r = torch.randn(1, 4, 5, 16, 16, generator=torch.Generator().manual_seed(0))
        lat_g = lat_r.to(dev)
        out = []
        for i, t in enumerate(sr.timesteps):
            with torch.no_grad():
                n_r = ref(torch.cat([lat_r] * 2), t, emb)
            n_g = net(torch.cat([lat_g] * 2), int(t), encoder_hidden_states=emb.to(dev)).sample
            # same-input forward error: feed the oracle's latent to the device net
            n_same = net(torch.cat([lat_r] * 2).to(dev), int(t), encoder_hidden_states=emb.to(dev)).sample
            u, c = n_r.chunk(2)
            e_r = u + gs * (c - u)
            us, cs_ = n_same.float().cpu().chunk(2)
            e_same = us + gs * (cs_ - us)
            lat_r = sr.step(e_r, t, lat_r).prev_sample
            lat_g, _ = sm.fused_cfg_step(lat_g, n_g, int(t), gs)
            out.append((i, rl(n_same, n_r), float((e_same - e_r).norm() / e_r.norm()), rl(lat_g, lat_r),
                        float((c - u).norm() / u.norm())))
        show = out if steps == 4 else out[::6] + out[-1:]
        print("gs=%.1f steps=%d" % (gs, steps))
        for o in show:
            print("  step %2d  fwd rel-l2 %.2e  guided-eps rel-l2 %.2e  latent rel-l2 %.2e  |c-u|/|u| %.2e" % o)
