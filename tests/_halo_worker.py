"""torchrun worker of test_gpu_multi.py: --halo_sync on 2 GPUs (halo frames read out of the peer's HBM over NVLink every
step) must give, bit for bit, the windows a single process computes for the same plan."""
import argparse
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import b2v  # noqa: E402,F401
from b2v import exchange, planner  # noqa: E402
from b2v.pipeline import VideoPipeline  # noqa: E402
from b2v.strategies.chunk_only import ChunkedVideoDiffuser  # noqa: E402
from b2v.strategies.common import CudaDenoiser, base_noise  # noqa: E402

TINY = dict(in_channels=4, out_channels=4, block_out_channels=(64, 128, 256, 256), layers_per_block=2,
            cross_attention_dim=128, attention_head_dim=64, norm_num_groups=32, norm_eps=1e-5, transformer_in_heads=2,
            text_len=77)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    cfg = argparse.Namespace(model_id="tiny", prompt="a rocket in space, 4k", num_frames=16, steps=4, guidance_scale=7.5,
                             chunk_size=6, overlap=2, fps=8, height=128, width=128, device="cuda",
                             out_csv="/tmp/b2v_halo_w%d.csv" % rank, emu_bw_mbps=0, emu_rtt_ms=0, emu_jitter_ms=0, seed=1234,
                             dtype="bf16", halo_sync=True, exchange="p2p")
    pipe = VideoPipeline.from_random("tiny", device=dev, seed=1234, dtype=torch.bfloat16, config=TINY)
    d = ChunkedVideoDiffuser(cfg, pipe=pipe)          # initialises NCCL from the torchrun env
    plan = d.make_plan()
    base = base_noise(4, cfg.num_frames, 16, 16, torch.bfloat16, dev)
    mine = d.denoise_my_windows_lockstep(plan, base)
    # the same plan in ONE process: all windows local, same lockstep, local halo buffers
    flat = planner.plan(cfg.num_frames, world, cfg.chunk_size, cfg.overlap, variant="chunk_only")
    every = [r for rk in range(world) for r in flat.of_rank(rk)]

    class OneRank:
        ranges, overlap, chunk_size = every, flat.overlap, flat.chunk_size

        @staticmethod
        def of_rank(r):
            return every
    halo1 = exchange.HaloSync(OneRank, 1, 0, 4, 16, 16, torch.bfloat16, dev)
    den = CudaDenoiser(pipe, d.uncond_emb, d.cond_emb, 7.5)
    with torch.no_grad():
        single = den.lockstep([base[:, :, s:e].clone() for s, e in every], halo1)
    ref = {r: t for r, t in zip(every, single)}
    for s, e, t in mine:
        assert torch.equal(t, ref[(s, e)]), "rank %d window %r differs from the single-process run" % (rank, (s, e))
    res = d()                                          # whole strategy incl. the fused end-of-loop exchange + blend
    assert bool(torch.isfinite(d.latents).all())
    dist.barrier()
    if rank == 0:
        print("windows %s halos %s: 2-GPU halo exchange == single process" % (plan.ranges, halo1.halos), flush=True)
        print("HALO_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
