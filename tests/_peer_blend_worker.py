"""torchrun worker of test_gpu_multi.py (one rank per GPU, NCCL): the fused peer-memory exchange + blend must equal the
NCCL all-gather + blend and the single-process blend of the same windows, bit for bit."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import b2v  # noqa: E402,F401
from b2v import exchange, ops, planner  # noqa: E402
from b2v.strategies.common import cuda_blend  # noqa: E402


def window(s, e, C, h, w, dtype, dev):
    g = torch.Generator().manual_seed(1000 * s + e)
    return torch.randn(1, C, e - s, h, w, generator=g).to(dtype).to(dev)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    C, h, w = 4, 72, 128
    for (T, cs, ov, variant, dtype) in [(24, 0, 4, "coherent", torch.bfloat16), (24, 8, 2, "chunk_only", torch.float16),
                                        (96, 16, 4, "coherent", torch.bfloat16)]:
        try:
            plan = planner.plan(T, world, cs, ov, variant=variant)
        except planner.PlanError:
            continue
        mine = [(s, e, window(s, e, C, h, w, dtype, dev)) for s, e in plan.of_rank(rank)]
        pw = exchange.PeerWindows(plan, world, rank, C, h, w, dtype, dev)
        for rep in range(2):  # buffers are reused
            fused, dt = pw.gather_blend(mine, T, plan.overlap)
        wins, _ = exchange.nccl_all_gather(mine, plan, rank, world, dev)
        via_nccl = cuda_blend(wins, T, plan.overlap)
        local_all = [(s, e, window(s, e, C, h, w, dtype, dev)) for r in range(world) for s, e in plan.of_rank(r)]
        single = cuda_blend(local_all, T, plan.overlap)
        assert torch.equal(fused, via_nccl), "peer-memory blend differs from the NCCL path (T=%d)" % T
        assert torch.equal(fused, single), "peer-memory blend differs from the single-process blend (T=%d)" % T
        if rank == 0:
            print("T=%d windows=%s %s: fused exchange+blend == nccl == single (%.3f ms)" %
                  (T, plan.ranges, str(dtype).split(".")[-1], dt * 1e3), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("PEER_BLEND_OK", flush=True)


if __name__ == "__main__":
    main()
