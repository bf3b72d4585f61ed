"""-m gpu: per-step halo exchange between adjacent windows (not in the reference — its own
single-process oracle is oracle/strategy_ref.denoise_halo_sync). World = 1 here (all windows on one GPU, local
pointers); tests/test_gpu_multi.py runs the same plan on 2 GPUs over NVLink and demands bit-identical latents."""
import argparse

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ulp(dtype):
    return {torch.float32: 2.0 ** -23, torch.float16: 2.0 ** -10, torch.bfloat16: 2.0 ** -7}[dtype]


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("T,cs,ov", [(16, 6, 2), (12, 6, 2), (8, 5, 1), (24, 16, 4)])
def test_halo_sync_vs_oracle(cuda, dtype, T, cs, ov):
    """One synchronisation of random windows: fp32 crossfade rounded once (oracle) vs the blend kernel's per-chunk
    rounding: <= 2 ulp of the latent dtype; the two owners of an overlap frame hold bit-identical values; frames
    outside the overlaps are untouched."""
    from b2v import exchange, planner
    from oracle import strategy_ref as S
    plan = planner.plan(T, 1, cs, ov, variant="chunk_only")
    ranges = plan.of_rank(0)
    assert len(ranges) >= 2
    g = torch.Generator().manual_seed(T + cs)
    lats = [torch.randn(1, 4, e - s, 8, 12, generator=g).to(dtype) for s, e in ranges]
    want = S.halo_sync(lats, ranges)
    halo = exchange.HaloSync(plan, 1, 0, 4, 8, 12, dtype, cuda)
    for rep in range(3):  # both parity slots, repeatedly
        got = halo.sync([l.to(cuda) for l in lats])
    for (s, e), a, b, src in zip(ranges, got, want, lats):
        assert a.shape == b.shape and a.dtype == dtype
        err = (a.cpu().float() - b.float()).abs()
        assert bool((err <= 2 * _ulp(dtype) * (b.float().abs() + 1.0)).all()), float(err.max())
    order = sorted(range(len(ranges)), key=lambda i: ranges[i])
    for i, j in zip(order[:-1], order[1:]):
        o = ranges[i][1] - ranges[j][0]
        n_i = ranges[i][1] - ranges[i][0]
        assert o == plan.overlap or o > 0
        assert torch.equal(got[i][:, :, n_i - o:], got[j][:, :, :o]), "owners of an overlap frame disagree"
    s0, e0 = ranges[order[0]]
    keep = (e0 - s0) - (ranges[order[0]][1] - ranges[order[1]][0])
    assert torch.equal(got[order[0]][:, :, :keep].cpu(), lats[order[0]][:, :, :keep])


def test_halo_plan_rejects_triple_cover():
    from b2v import exchange, planner
    plan = planner.plan(24, 4, 0, 4, variant="coherent")   # 6-frame windows every 2 frames: 3 windows per frame
    with pytest.raises(ValueError):
        exchange.halo_neighbours(plan, 4)


def _tiny(cuda):
    from oracle import unet3d_ref as R
    from b2v.unet import UNet3DConditionModel
    from b2v.pipeline import VideoPipeline
    from b2v.scheduler import DDIMScheduler
    ref = R.random_init_(R.UNet3DConditionModel(R.TINY), seed=1234, round_bf16=True).eval()
    net = UNet3DConditionModel(R.TINY.to_dict(), device=cuda).load_state_dict(ref.state_dict())
    return ref, VideoPipeline(net, DDIMScheduler.from_pipeline_config(), text_dim=128, dtype=torch.float32)


def _cfg(**kw):
    d = dict(model_id="tiny", prompt="a rocket in space, 4k", num_frames=12, steps=3, guidance_scale=7.5, chunk_size=6,
             overlap=2, fps=8, height=128, width=128, device="cuda", out_csv="/tmp/b2v_halo.csv", emu_bw_mbps=0,
             emu_rtt_ms=0, emu_jitter_ms=0, seed=1234, dtype="fp32", halo_sync=True)
    d.update(kw)
    return argparse.Namespace(**d)


def test_lockstep_denoise_vs_oracle(cuda, tmp_path):
    """3 windows of 6 frames (overlap 2) through 3 DDIM steps with a halo exchange after each: every synchronised
    per-step latent against the fp32 oracle, rel-l2 <= 6e-2 (the CFG 7.5 budget of test_gpu_strategy.py)."""
    from b2v.strategies.chunk_only import ChunkedVideoDiffuser
    from b2v.strategies.common import base_noise
    from oracle import strategy_ref as S
    from oracle.ddim_ref import pipeline_scheduler
    ref, pipe = _tiny(cuda)
    cfg = _cfg(out_csv=str(tmp_path / "h.csv"))
    d = ChunkedVideoDiffuser(cfg, pipe=pipe)
    plan = d.make_plan()
    ranges = plan.of_rank(0)
    assert ranges == [(0, 6), (4, 10), (8, 12)]
    base = base_noise(4, 12, 16, 16, torch.float32, torch.device("cpu"))
    sched = pipeline_scheduler()
    sched.set_timesteps(3)
    want = {}
    S.denoise_halo_sync(ref, sched, base, ranges, d.uncond_emb.cpu().float(), d.cond_emb.cpu().float(), 7.5,
                        record=lambda i, t, lats: want.__setitem__(i, [l.clone() for l in lats]))
    got = {}
    mine = d.denoise_my_windows_lockstep(plan, base.to(cuda), on_sync=lambda i, t, lats: got.__setitem__(i, [l.cpu() for l in lats]))
    worst = 0.0
    for i in want:
        for a, b in zip(got[i], want[i]):
            worst = max(worst, float((a - b).norm() / b.norm()))
    print("halo lockstep worst per-step rel-l2 %.3e" % worst)
    assert worst <= 6e-2, worst
    # after the last exchange neighbouring windows agree bit for bit on their shared frames
    assert torch.equal(mine[0][2][:, :, 4:], mine[1][2][:, :, :2]) and torch.equal(mine[1][2][:, :, 4:], mine[2][2][:, :, :2])
    # and the exchange really changes the trajectory (independent windows disagree there)
    cfg2 = _cfg(halo_sync=False, out_csv=str(tmp_path / "h.csv"))
    ind = ChunkedVideoDiffuser(cfg2, pipe=pipe).denoise_my_windows(plan, base.to(cuda))
    assert not torch.equal(ind[0][2][:, :, 4:], ind[1][2][:, :, :2])
    # whole run through the CLI object: same windows, blended and finite
    res = d()
    assert res["num_frames"] == 12 and bool(torch.isfinite(d.latents).all())
