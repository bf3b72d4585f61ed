"""BASELINE config 5 end to end: Proof-of-Inference on N GPUs of one node.

    torchrun --nproc_per_node=N tools/poi_run.py [--steps 50] [--frames 24] [--height 576] [--width 1024] [--reveal 5]

rank 0 plays the miner: records a `steps`-step ZeroscopeXL trajectory (InferNet/neurons/miner.py:571-589: batch 1, no CFG,
fp16 latents, random-init weights), commits it (SHA-256 leaves + Merkle tree on the device) and builds the wire fields.
The validator side (validator.py:641-845 + proof.py:397-416) draws `reveal` consecutive-pair start indices from the seed
derived from the challenge (10 % of the leaves at the defaults), and ALL ranks re-execute their share of the revealed
leaves (membership + UNet re-run + DDIM chaining, b2v.miner.reexecute_spot_check): independent units, no communication
besides shipping the revealed leaves and collecting the verdicts. Prints one JSON line on rank 0.
"""
import argparse
import json
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import b2v  # noqa: E402,F401
from b2v import miner, proof  # noqa: E402
from b2v.pipeline import VideoPipeline  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--frames", type=int, default=24)
    ap.add_argument("--height", type=int, default=576)
    ap.add_argument("--width", type=int, default=1024)
    ap.add_argument("--reveal", type=int, default=5, help="consecutive pairs to reveal (2 leaves each)")
    args = ap.parse_args()
    rank, world = int(os.getenv("RANK", 0)), int(os.getenv("WORLD_SIZE", 1))
    local = int(os.getenv("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    pipe = VideoPipeline.from_random("cerspense/zeroscope_v2_XL", device=dev, seed=1234, dtype=torch.float16)
    toks = pipe.tokenizer(["a rocket in space, 4k"])
    ehs = pipe.text_encoder(toks.input_ids)[0][:1].to(torch.float16)
    shape = (1, 4, args.frames, args.height // 8, args.width // 8)
    out = {"config": {"steps": args.steps, "latent": list(shape), "world": world}}

    payload = [None]
    if rank == 0:
        z0 = torch.randn(shape, generator=torch.Generator().manual_seed(0)).half().to(dev)
        miner.record_trajectory(pipe, z0, ehs, 2)  # warm-up: CUDA-graph capture for this shape
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        traj = miner.record_trajectory(pipe, z0, ehs, args.steps, commit="end")
        torch.cuda.synchronize()
        t_record = time.perf_counter() - t0
        t0 = time.perf_counter()
        root, leaf = traj.merkle(with_bytes=True)
        t_tree = time.perf_counter() - t0
        t0 = time.perf_counter()
        wire = traj.wire()
        t_wire = time.perf_counter() - t0
        challenge = b"\x07" * 32
        seed = proof.derive_seed(challenge, "5FminerHotkey")
        idx = proof.commit_then_reveal_merkle_spotcheck(root, args.steps, args.reveal, seed)
        revealed = {traj.timesteps[i]: leaf[traj.timesteps[i]] for i in sorted(set(idx))}
        payload = [{"root": root, "timesteps": traj.timesteps, "revealed": revealed}]
        leaf_bytes = sum(len(v[0]) + len(v[1]) + 2 for v in leaf.values())
        out["miner"] = {"record_s": t_record, "steps_per_s": args.steps / t_record,
                        "merkle_tree_and_leaf_bytes_s": t_tree, "wire_base64_s": t_wire,
                        "leaf_bytes": leaf_bytes, "root": root.hex(),
                        "wire_chars": sum(len(x) for x in wire["latents"]) + sum(len(x) for x in wire["noise_preds"])}
        out["spot_check"] = {"indices": idx, "revealed_leaves": len(revealed), "fraction": len(revealed) / args.steps}
    if world > 1:
        dist.broadcast_object_list(payload, src=0)
    pl = payload[0]
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    rep = miner.reexecute_spot_check(pipe, ehs, pl["revealed"], pl["root"], pl["timesteps"], shape, args.steps,
                                     rank=rank, world=world)
    torch.cuda.synchronize()
    t_mine = time.perf_counter() - t0
    reports = [None] * world
    if world > 1:
        dist.all_gather_object(reports, {"rank": rank, "s": t_mine, "rep": rep})
    else:
        reports = [{"rank": 0, "s": t_mine, "rep": rep}]
    if rank == 0:
        leaves = {}
        for r in reports:
            leaves.update(r["rep"]["leaves"])
        checked = [r for r in reports if r["rep"]["leaves"]]
        out["validator"] = {
            "ok": all(r["rep"]["ok"] for r in checked) and len(leaves) == len(pl["revealed"]),
            "verify_s_max_over_ranks": max(r["s"] for r in reports),
            "leaves_checked": len(leaves),
            "eps_exact": sum(1 for v in leaves.values() if v["eps_exact"]),
            "chained": sum(1 for v in leaves.values() if v["chain"] is True),
            "per_rank_leaves": [len(r["rep"]["leaves"]) for r in reports]}
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
