"""One validator round at ZeroscopeXL leaf size (validator.py:641-845): M miners, each with an
N-step trajectory of [1,4,24,72,128] fp16 leaves (3.5 MB per leaf image pair), k = 5 consecutive pairs revealed per miner
as base64 text (what `open_leaves` returns).

    python tools/validator_round.py [--miners 16] [--steps 12] [--pairs 5]

Times (wall clock around complete calls, inputs = the host-side reveal dictionaries):
  * b2v.validator.check_miners           all miners in one batched device pass (text in, four launches)
  * b2v.validator.check_revealed_leaves  miner by miner, leaf by leaf (the reference's sequence on the device kernels)
  * host byte work only                  base64.b64decode + hashlib.sha256 of every revealed leaf, Python stdlib, one thread —
                                         the part of the reference's loop that dominates it at this leaf size
Prints one JSON line.
"""
import argparse
import base64
import hashlib
import json
import os
import random
import sys
import time

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import b2v  # noqa: E402,F401
from b2v import ops, proof  # noqa: E402
from b2v.validator import Submission, check_miners, check_revealed_leaves, select_spot_check_pairs  # noqa: E402

SHAPE = (1, 4, 24, 72, 128)
CFG = {"latent_channels": 4, "latent_height": 72, "latent_width": 128}


def b64(b):
    return base64.b64encode(b).decode("ascii")


def miner(dev, seed, n_steps, k, cheat=False):
    g = torch.Generator(device=dev).manual_seed(seed)
    sched = proof._default_scheduler(n_steps)
    ts_sched = sched.host_timesteps()
    z = (0.5 * torch.randn(SHAPE, device=dev, generator=g)).half()
    zs, es = [], []
    for i in range(n_steps):
        e = (0.05 * torch.randn(SHAPE, device=dev, generator=g)).half()
        a_t, a_p = sched.alpha_pair(ts_sched[i])
        zs.append(z)
        es.append(e)
        x0 = ((z.float() - (1 - a_t) ** 0.5 * e.float()) / a_t ** 0.5).clamp(-1, 1)
        z = (a_p ** 0.5 * x0 + (1 - a_p) ** 0.5 * e.float()).half()
    labels = [int(t) for t in ts_sched]
    root, leaf = proof.build_merkle_tree(zs, es, labels)
    pick = select_spot_check_pairs(labels, k, random.Random(seed))
    reveal = {str(t): (b64(leaf[t][0]), b64(leaf[t][1]), [b64(p) for p in leaf[t][2]]) for t in pick}
    if cheat:
        t = str(pick[0])
        raw = bytearray(base64.b64decode(reveal[t][1]))
        raw[100] ^= 1
        reveal[t] = (reveal[t][0], b64(bytes(raw)), reveal[t][2])
    return Submission("uid%d" % seed, root, labels, [0.5] * n_steps, pick, reveal, dict(CFG))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--miners", type=int, default=16)
    ap.add_argument("--steps", type=int, default=12)
    ap.add_argument("--pairs", type=int, default=5)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    subs = [miner(dev, 10 + i, a.steps, a.pairs, cheat=(i % 4 == 1)) for i in range(a.miners)]
    n_leaves = sum(len(s.spot_check_indices) for s in subs)
    text_bytes = sum(len(v[0]) + len(v[1]) for s in subs for v in s.leaves_result.values())
    want = [i % 4 != 1 for i in range(a.miners)]
    check_miners(subs[:2])  # warm-up: pinned staging buffer, module load
    torch.cuda.synchronize()
    l0 = ops.launch_count()
    t0 = time.perf_counter()
    got = check_miners(subs)
    t_first = time.perf_counter() - t0    # includes growing the pinned staging buffer to this round's size
    launches = ops.launch_count() - l0
    assert got == want, got
    t_batch = 1e9
    for _ in range(3):                    # steady state: a validator runs round after round
        t0 = time.perf_counter()
        assert check_miners(subs) == want
        t_batch = min(t_batch, time.perf_counter() - t0)
    t0 = time.perf_counter()
    seq = [check_revealed_leaves(s) for s in subs]
    t_seq = time.perf_counter() - t0
    assert seq == want
    t0 = time.perf_counter()
    for s in subs:
        for t in s.spot_check_indices:
            zb, eb, _ = s.leaves_result[str(t)]
            hashlib.sha256(int(t).to_bytes(2, "big") + base64.b64decode(zb) + base64.b64decode(eb)).digest()
    t_host = time.perf_counter() - t0
    print(json.dumps({"workload": "%d miners x %d-step XL trajectories, %d revealed leaves, %.1f MB of base64 text"
                                  % (a.miners, a.steps, n_leaves, text_bytes / 1e6),
                      "check_miners_s": t_batch, "check_miners_first_call_s": t_first,
                      "check_miners_launches": launches,
                      "leaves_per_s": n_leaves / t_batch, "text_gb_per_s": text_bytes / t_batch / 1e9,
                      "sequential_device_s": t_seq, "host_stdlib_b64_sha256_s": t_host,
                      "verdicts_ok": got == want, "cheaters_caught": want.count(False)}), flush=True)


if __name__ == "__main__":
    main()
