"""Where does a --halo_sync step spend its time? One GPU, world = 1, ZeroscopeXL windows (0,16) + (12,24):
independent windows vs lockstep + per-step halo exchange, and the exchange alone. Prints one JSON line."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import b2v  # noqa: E402,F401
from b2v import exchange, planner  # noqa: E402
from b2v.pipeline import VideoPipeline  # noqa: E402
from b2v.strategies.common import CudaDenoiser, base_noise  # noqa: E402


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 6
    dev = torch.device("cuda:0")
    pipe = VideoPipeline.from_random("cerspense/zeroscope_v2_XL", device=dev, seed=1234, dtype=torch.bfloat16)
    pipe.scheduler.set_timesteps(steps, device="cuda")
    toks = pipe.tokenizer(["a rocket in space, 4k", ""])
    emb = pipe.text_encoder(toks.input_ids)[0]
    plan = planner.plan(24, 1, 16, 4, variant="chunk_only")
    ranges = plan.of_rank(0)
    base = base_noise(4, 24, 72, 128, torch.bfloat16, dev)
    den = CudaDenoiser(pipe, emb[1:], emb[:1], 7.5)
    halo = exchange.HaloSync(plan, 1, 0, 4, 72, 128, torch.bfloat16, dev)
    out = {"windows": ranges, "steps": steps}
    with torch.no_grad():
        for tag in ("warm", "plain", "lockstep"):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            if tag == "lockstep":
                halo.step = 0
                den.lockstep([base[:, :, s:e].clone() for s, e in ranges], halo)
            else:
                for s, e in ranges:
                    den(base[:, :, s:e].clone())
            torch.cuda.synchronize()
            out[tag + "_ms_per_step"] = (time.perf_counter() - t0) / steps * 1e3
        lats = [base[:, :, s:e].clone() for s, e in ranges]
        halo.sync(lats)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(50):
            lats = halo.sync(lats)
        t_host = (time.perf_counter() - t0) / 50 * 1e3
        torch.cuda.synchronize()
        out["sync_host_ms"] = t_host
        out["sync_total_ms"] = (time.perf_counter() - t0) / 50 * 1e3
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
