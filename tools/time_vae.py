"""Frame decode timing on one B200: T frames of a ZeroscopeXL / 576w latent -> uint8 frames.

    python tools/time_vae.py [--frames 24] [--height 576] [--width 1024] [--iters 3] [--batch 0]

Prints one JSON line: ms per decode (CUDA events, after one warm-up), algorithmic TFLOP/s, launches, peak memory.
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import b2v  # noqa: E402,F401
from b2v import ops  # noqa: E402
from b2v.vae import AutoencoderKL  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=24)
    ap.add_argument("--height", type=int, default=576)
    ap.add_argument("--width", type=int, default=1024)
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--batch", type=int, default=0, help="frames per decoder batch (0 = auto)")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    net = AutoencoderKL.from_random(device=dev)
    if a.batch:
        net.frames_per_batch = a.batch
    h, w = a.height // 8, a.width // 8
    lat = (torch.randn(1, 4, a.frames, h, w, generator=torch.Generator().manual_seed(0)) * 0.3).to(dev)
    net.decode_frames(lat)
    torch.cuda.synchronize()
    torch.cuda.reset_peak_memory_stats()
    l0 = ops.launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(a.iters + 1)]
    ev[0].record()
    for i in range(a.iters):
        out = net.decode_frames(lat)
        ev[i + 1].record()
    torch.cuda.synchronize()
    ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(a.iters)]
    fl = net.flops(a.frames, h, w)
    best = min(ms)
    print(json.dumps({"workload": "vae decode_frames %d x %dx%d" % (a.frames, a.height, a.width), "ms": ms,
                      "ms_per_frame": best / a.frames, "tflops_algorithmic": fl / 1e12,
                      "tflops_per_s": fl / best / 1e9, "frames_per_batch": net._batch(a.frames, a.height, a.width),
                      "launches_per_decode": (ops.launch_count() - l0) // a.iters,
                      "peak_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30,
                      "out": list(out.shape), "mean_byte": float(out.float().mean())}), flush=True)


if __name__ == "__main__":
    main()
