"""ORACLE — test infrastructure only. Generates the floating-point golden fixtures of tests/golden/ that are too heavy
to recompute inside every GPU test run (tens of CPU TFLOP each): the fp32 oracle (oracle/unet3d_ref.py +
oracle/strategy_ref.py + oracle/ddim_ref.py) executed HERE, on the shapes BASELINE.json names, with the seeded
random-init protocol of SURVEY.md §8c/§8d. The fixtures travel to the GPU box; this script and the seeds below are
all that is needed to regenerate them bit for bit (torch CPU generators are deterministic).

  config0_chunk_only.npz   BASELINE configs[0] exactly as written: chunk_only, 2 segments, FULL-WIDTH Zeroscope v2 576w
                           UNet3D (1 411 233 860 parameters, seed 1234), 8 frames of 256x256 (latent 32x32), 4 DDIM
                           steps (751, 501, 251, 1), guidance 7.5, base noise seed 0, text embeddings seed 1.
                           Path restated: Distribution/strategies/chunk_only.py:65-150. Stored: the plan, every
                           window's latent after every step (fp32), every guided eps (fp16), the blended video latent.
  xl_forward_f2.npz        one forward of the same network at the ZeroscopeXL latent size 72x128 (1024x576 video),
                           B = 2 (the CFG pair), F = 2, timestep 501 — the spatial tile / attention shapes (seq 9216)
                           of the headline benchmark. chunk_only.py:71 at --height 576 --width 1024.
(The attention / conv3x3 kernels at the XL shapes are compared with torch fp32 computed live inside the GPU tests:
their oracle is a fraction of a TFLOP and its output would be tens of MB as a fixture.)

usage: python oracle/gen_golden_unet.py [config0] [xl_forward]      (default: both; ~10 min on 8 cores)
"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tests", "golden")

from oracle import strategy_ref as S  # noqa: E402
from oracle import unet3d_ref as R  # noqa: E402
from oracle.ddim_ref import pipeline_scheduler  # noqa: E402

WEIGHT_SEED, TEXT_SEED, NOISE_SEED = 1234, 1, 0


def full_width_model():
    t0 = time.time()
    ref = R.random_init_(R.UNet3DConditionModel(R.ZEROSCOPE), seed=WEIGHT_SEED, round_bf16=True).eval()
    n = sum(p.numel() for p in ref.parameters())
    assert n == 1411233860, n
    print("model: %d parameters, init %.1fs" % (n, time.time() - t0), flush=True)
    return ref


def text_embeddings(dim=1024):
    g = torch.Generator().manual_seed(TEXT_SEED)
    cond, uncond = torch.randn(1, 77, dim, generator=g), torch.randn(1, 77, dim, generator=g)
    return cond, uncond


def gen_config0(ref):
    T, h, w, world, steps, gs = 8, 32, 32, 2, 4, 7.5
    cond, uncond = text_embeddings()
    base = torch.randn(1, 4, T, h, w, generator=torch.Generator().manual_seed(NOISE_SEED))
    sched = pipeline_scheduler()
    sched.set_timesteps(steps)
    assert [int(t) for t in sched.timesteps] == [751, 501, 251, 1]
    lat_steps, eps_steps = {}, {}

    def rec(rank, s, e, i, t, lat_in, eps, lat_out):
        lat_steps[(s, e, i)] = lat_out.clone()
        eps_steps[(s, e, i)] = eps.clone()
        print("  rank %d window (%d,%d) step %d t=%d |lat| %.4f" % (rank, s, e, i, t, float(lat_out.norm())), flush=True)

    t0 = time.time()
    out = S.run_strategy(ref, sched, base, uncond, cond, world, gs, 0, 4, coherent=False, record=rec)
    print("config0: %.1fs, plan cs=%d ov=%d %s" % (time.time() - t0, out["cs"], out["ov"], out["ranges"]), flush=True)
    assert out["ranges"] == [(0, 5), (4, 8)] and out["ov"] == 1
    arrays = dict(ranges=np.array(out["ranges"], dtype=np.int64), ov=np.int64(out["ov"]), cs=np.int64(out["cs"]),
                  timesteps=np.array([int(t) for t in sched.timesteps], dtype=np.int64),
                  lat=out["lat"].numpy(), full=out["full"].numpy())
    for (s, e, i), v in lat_steps.items():
        arrays["lat_%d_%d_%d" % (s, e, i)] = v.numpy()
        arrays["eps_%d_%d_%d" % (s, e, i)] = eps_steps[(s, e, i)].half().numpy()
    np.savez_compressed(os.path.join(OUT, "config0_chunk_only.npz"), **arrays)


def gen_xl_forward(ref):
    g = torch.Generator().manual_seed(11)
    B, F, H, W = 2, 2, 72, 128
    x = torch.randn(B, 4, F, H, W, generator=g)
    ehs = torch.randn(B, 77, 1024, generator=g)
    t0 = time.time()
    with torch.no_grad():
        want = ref(x, 501, ehs)
    print("xl forward: %.1fs |out| %.4f" % (time.time() - t0, float(want.norm())), flush=True)
    np.savez_compressed(os.path.join(OUT, "xl_forward_f2.npz"), out=want.numpy(), shape=np.array([B, F, H, W]),
                        timestep=np.int64(501), seed=np.int64(11))


if __name__ == "__main__":
    which = sys.argv[1:] or ["config0", "xl_forward"]
    torch.set_num_threads(os.cpu_count() or 8)
    if "config0" in which or "xl_forward" in which:
        model = full_width_model()
        if "config0" in which:
            gen_config0(model)
        if "xl_forward" in which:
            gen_xl_forward(model)
