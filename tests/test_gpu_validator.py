"""-m gpu: validator spot-check orchestration over many miners (validator.py:641-845): the batched
device path (`b2v.validator.check_miners`) gives, miner by miner, the verdict of the sequential path and of the host
oracle (oracle/validator_ref.py on hashlib + the restated default DDIM scheduler). Verdicts are booleans: exact."""
import base64
import random

import pytest
import torch

pytestmark = pytest.mark.gpu

C_, F_, H_, W_ = 4, 3, 8, 8
CFG = {"latent_channels": C_, "latent_height": H_, "latent_width": W_}
N_STEPS = 12


def _b64(b):
    return base64.b64encode(b).decode("ascii")


def _miner(cuda, seed, fault=None):
    """One miner's committed trajectory (random but bounded leaves) and its reveal, optionally tampered with."""
    from b2v import proof
    from b2v.validator import Submission, select_spot_check_pairs
    from oracle.ddim_ref import default_scheduler
    g = torch.Generator().manual_seed(seed)
    ts = [int(1000 - 80 * i - seed % 7) for i in range(N_STEPS)]   # the miner's own timestep labels
    # an honest trajectory under the scheduler the validator re-derives with (library-default DDIM, N_STEPS steps):
    # z_{i+1} = step(eps_i, timesteps[i], z_i) in fp16, small eps so that every later z stays inside the clip range
    sch = default_scheduler()
    sch.set_timesteps(N_STEPS)
    z = (0.5 * torch.randn(1, C_, F_, H_, W_, generator=g)).half()
    zs, es = [], []
    for i in range(N_STEPS):
        e = (0.05 * torch.randn(1, C_, F_, H_, W_, generator=g)).half()
        zs.append(z)
        es.append(e)
        z = sch.step(e, sch.timesteps[i], z).prev_sample.half()
    pick = select_spot_check_pairs(ts, 3, random.Random(seed))
    if fault == "incoherent":  # committed, in range, but z_{i+1} is nowhere near the DDIM step of (z_i, eps_i)
        i = ts.index(pick[0])
        zs[i] = torch.full_like(zs[i], 3.0)
        es[i] = torch.full_like(es[i], 3.0)
        zs[i + 1] = torch.full_like(zs[i + 1], -2.0)
    if fault == "unbounded":   # committed garbage: NaN in eps
        es[ts.index(pick[1])][0, 0, 0, 0, 0] = float("nan")
    root, leaf = proof.build_merkle_tree([z.to(cuda) for z in zs], [e.to(cuda) for e in es], ts)
    alphas = [0.5] * N_STEPS
    reveal = {str(t): (_b64(leaf[t][0]), _b64(leaf[t][1]), [_b64(p) for p in leaf[t][2]]) for t in pick}
    resp_ts = list(ts)
    if fault == "flip":
        t = str(pick[2])
        raw = bytearray(base64.b64decode(reveal[t][0]))
        raw[5] ^= 0x40
        reveal[t] = (_b64(bytes(raw)), reveal[t][1], reveal[t][2])
    elif fault == "missing":
        del reveal[str(pick[1])]
    elif fault == "malformed":
        reveal[str(pick[0])] = reveal[str(pick[0])][:2]
    elif fault == "path":
        t = str(pick[3])
        reveal[t] = (reveal[t][0], reveal[t][1], list(reversed(reveal[t][2]))[:-1] + [_b64(b"\x00" * 32)])
    elif fault == "unknown_t":
        resp_ts = [x for x in ts if x != pick[2]]
    elif fault == "short_alphas":
        alphas = alphas[:2]
    elif fault == "bad_b64":
        reveal[str(pick[0])] = ("@@@", reveal[str(pick[0])][1], reveal[str(pick[0])][2])
    elif fault == "wrong_root":
        root = bytes(32)
    elif fault == "truncated":
        t = str(pick[1])
        reveal[t] = (_b64(base64.b64decode(reveal[t][0])[:-2]), reveal[t][1], reveal[t][2])
    return Submission("uid%d" % seed, root, resp_ts, alphas, pick, reveal, dict(CFG))


FAULTS = [None, "flip", None, "missing", "malformed", "path", "unknown_t", "short_alphas", "incoherent", "unbounded",
          "bad_b64", "wrong_root", "truncated", None]


def test_batched_verdicts_equal_sequential_and_oracle(cuda):
    from b2v.validator import check_miners, check_revealed_leaves
    from oracle import validator_ref as VR
    subs = [_miner(cuda, 100 + i, f) for i, f in enumerate(FAULTS)]
    batch = check_miners(subs, request_id="r1")
    seq = [check_revealed_leaves(s, request_id="r1") for s in subs]
    want = [VR.check_revealed_leaves(s.merkle_root, s.timesteps, s.alphas, s.spot_check_indices, s.leaves_result,
                                     s.unet_config) for s in subs]
    print("verdicts", list(zip(FAULTS, want)))
    assert want == [f is None for f in FAULTS]  # every fault is caught, every honest miner passes
    assert seq == want
    assert batch == want


def test_check_miners_edge_cases(cuda):
    from b2v.validator import Submission, check_miners
    assert check_miners([]) == []
    honest = _miner(cuda, 7)
    # a miner asked for nothing passes vacuously (validator.py:768: empty loop, < 2 indices skips coherence)
    empty = Submission("e", honest.merkle_root, honest.timesteps, honest.alphas, [], {}, dict(CFG))
    # alphas missing from the proof fall back to the validator's configured schedule (validator.py:766)
    cfg = dict(CFG)
    cfg["alphas"] = [0.5] * N_STEPS
    fallback = Submission("f", honest.merkle_root, honest.timesteps, [], honest.spot_check_indices, honest.leaves_result, cfg)
    broken = Submission("b", honest.merkle_root, ["x"], honest.alphas, honest.spot_check_indices, honest.leaves_result,
                        dict(CFG))
    odd_root = Submission("o", honest.merkle_root[:31], honest.timesteps, honest.alphas, honest.spot_check_indices,
                          honest.leaves_result, dict(CFG))  # not a SHA-256 root: can only fail, and must not raise
    from b2v.validator import check_revealed_leaves
    subs = [empty, honest, fallback, broken, odd_root]
    assert check_miners(subs) == [True, True, True, False, False]
    assert [check_revealed_leaves(s) for s in subs] == [True, True, True, False, False]


def test_many_miners_one_round(cuda):
    """64 miners x 6 revealed leaves: still the sequential verdicts."""
    from b2v import ops
    from b2v.validator import check_miners, check_revealed_leaves
    subs = [_miner(cuda, 1000 + i, "flip" if i % 5 == 3 else None) for i in range(64)]
    l0 = ops.launch_count()
    got = check_miners(subs)
    launches = ops.launch_count() - l0
    assert got == [i % 5 != 3 for i in range(64)]
    assert got[:8] == [check_revealed_leaves(s) for s in subs[:8]]
    print("launches for 64 miners:", launches)
    assert launches <= 8  # hash + membership (one per path depth) + step re-check + coherence (sequential: >= 13 per miner)
