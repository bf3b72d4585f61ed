"""Validator spot-check orchestration (InferNet/neurons/validator.py:641-845) over MANY miners.

The reference walks its miners one by one and, per revealed leaf, base64-decodes, hashes on the host with hashlib,
folds the Merkle path, re-derives the DDIM step (``run_unet_step``) and finally checks temporal coherence — about
4 k + 1 small host/CPU jobs per miner. Here the same verdicts come from a handful of device launches for the whole round:

  * ``select_spot_check_pairs``  the index draw of validator.py:641-684 (global ``random``, duplicate timesteps dropped);
  * ``check_revealed_leaves``    one miner, the reference's order of checks (validator.py:768-830);
  * ``check_miners``             all miners at once: ONE ``b2v_sha256_leaves`` launch over every revealed leaf of every
                                 miner (bytes staged through one pinned buffer), ONE ``b2v_merkle_verify_roots``
                                 per path depth (leaf i against its own miner's root), ONE ``b2v_spotcheck_ddim`` launch per scheduler length for all
                                 P4 verdicts and ONE for all temporal-coherence pairs.

Both return exactly what a loop over ``check_revealed_leaves`` returns (``tests/test_gpu_validator.py``). Networking
(the ``open_leaves`` RPC), scoring and weights stay the reference's. Nothing here raises: a malformed submission is a
``False`` verdict with the reason logged, as in the reference (its try/except at validator.py:596-1020 scores 0).
"""
import base64
import logging
import random
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from . import ops, proof

LOG = logging.getLogger("b2v.validator")


def select_spot_check_pairs(all_timesteps: Sequence[int], k: int, rng=random) -> List[int]:
    """validator.py:641-684: k distinct start positions in [0, n-2] from ``random.sample``; pair (t[s], t[s+1]) is kept
    only when neither timestep was already taken by an earlier pair. Fewer than two timesteps: the list itself."""
    indices = list(all_timesteps)
    if len(indices) < 2:
        return indices
    max_start = len(indices) - 2
    starts = rng.sample(range(max_start + 1), min(k, max_start + 1))
    out, used = [], set()
    for s in starts:
        t1, t2 = indices[s], indices[s + 1]
        if t1 not in used and t2 not in used:
            out.extend([t1, t2])
            used.update((t1, t2))
    return out


class Submission:
    """What the validator holds for one miner after ``open_leaves`` (validator.py:686-766)."""
    __slots__ = ("uid", "merkle_root", "timesteps", "alphas", "spot_check_indices", "leaves_result", "unet_config")

    def __init__(self, uid, merkle_root: bytes, timesteps: Sequence[int], alphas: Sequence[float],
                 spot_check_indices: Sequence[int], leaves_result: Dict[str, Tuple[str, str, List[str]]],
                 unet_config: Dict):
        self.uid, self.merkle_root = uid, merkle_root
        self.timesteps, self.alphas = list(timesteps), list(alphas)
        self.spot_check_indices, self.leaves_result, self.unet_config = list(spot_check_indices), leaves_result, unet_config


def check_revealed_leaves(sub: Submission, request_id: Optional[str] = None, logger=None) -> bool:
    """One miner, the reference's sequence (validator.py:768-830): for each requested timestep the leaf must be present
    and well formed, hash to a member of the committed tree, map to a step with an alpha, and pass ``run_unet_step``;
    then the temporal-coherence check over all revealed leaves. First failure decides."""
    logger = logger or LOG
    try:
        t2s = {int(ts): i for i, ts in enumerate(sub.timesteps)}
        alphas = sub.alphas or sub.unet_config["alphas"]
        ok = True
        for t in sub.spot_check_indices:
            leaf = sub.leaves_result.get(str(t))
            if not leaf or len(leaf) != 3:
                logger.warning("[%s] missing or malformed leaf for t=%s from miner %s", request_id, t, sub.uid)
                ok = False
                break
            z, e = base64.b64decode(leaf[0]), base64.b64decode(leaf[1])
            path = [base64.b64decode(p) for p in leaf[2]]
            dev = proof._device()
            h = ops.sha256_leaves([proof._to_dev_bytes(z, dev)], [proof._to_dev_bytes(e, dev)], [t]).cpu().numpy()[0].tobytes()
            if not proof.verify_merkle_leaf(h, path, sub.merkle_root):
                logger.warning("[%s] Merkle verification failed for miner %s at t=%s", request_id, sub.uid, t)
                ok = False
                break
            step = t2s.get(t)
            if step is None or step >= len(alphas):
                logger.warning("[%s] invalid step index for t=%s: %s", request_id, t, step)
                ok = False
                break
            cfg = dict(sub.unet_config)
            cfg["alphas"] = alphas
            if not proof.run_unet_step(z, e, cfg, step):
                logger.warning("[%s] step verification failed for miner %s at t=%s", request_id, sub.uid, t)
                ok = False
                break
        if len(sub.spot_check_indices) >= 2:
            if not proof.verify_temporal_coherence_spot_check(sub.spot_check_indices, sub.leaves_result, t2s, alphas,
                                                              sub.unet_config, request_id, logger):
                ok = False
        return ok
    except Exception as ex:  # validator.py:1014-1020: any error scores the miner 0
        logger.error("[%s] spot check of miner %s failed: %r", request_id, sub.uid, ex)
        return False


def check_miners(subs: Sequence[Submission], request_id: Optional[str] = None, logger=None) -> List[bool]:
    """Every miner of a round in a handful of launches; element i == check_revealed_leaves(subs[i])."""
    logger = logger or LOG
    n = len(subs)
    ok = [True] * n
    dev = proof._device()
    # ---- host pass: decode, collect ------------------------------------------------------------------------------
    leaves = []   # (miner, t, z_bytes, eps_bytes, path, step, alphas, cfg)
    meta = []
    for m, sub in enumerate(subs):
        try:
            t2s = {int(ts): i for i, ts in enumerate(sub.timesteps)}
            alphas = sub.alphas or sub.unet_config["alphas"]
            meta.append((t2s, alphas))
            for t in sub.spot_check_indices:
                leaf = sub.leaves_result.get(str(t))
                if not leaf or len(leaf) != 3:
                    logger.warning("[%s] missing or malformed leaf for t=%s from miner %s", request_id, t, sub.uid)
                    ok[m] = False
                    break
                z, e = base64.b64decode(leaf[0]), base64.b64decode(leaf[1])
                path = [base64.b64decode(p) for p in leaf[2]]
                if not 0 <= int(t) < 65536:
                    raise OverflowError("int too big to convert")
                step = t2s.get(t)
                if step is None or step >= len(alphas):
                    # the reference reaches this only after the Merkle check passed; either way the verdict is False
                    logger.warning("[%s] invalid step index for t=%s: %s", request_id, t, step)
                    ok[m] = False
                    break
                leaves.append((m, int(t), z, e, path, step, alphas, sub.unet_config))
        except Exception as ex:
            logger.error("[%s] spot check of miner %s failed: %r", request_id, sub.uid, ex)
            ok[m] = False
            if len(meta) <= m:
                meta.append(({}, []))
    live = [lf for lf in leaves if ok[lf[0]]]
    # ---- leaf hashes: one pinned staging buffer, one copy, ONE launch ----------------------------------------------
    if live:
        total = sum(len(lf[2]) + len(lf[3]) for lf in live)
        host = proof._pinned(max(total, 1))[:total]
        hv, off, spans = memoryview(host.numpy()), 0, []
        for lf in live:
            for b in (lf[2], lf[3]):
                hv[off:off + len(b)] = b
                spans.append((off, len(b)))
                off += len(b)
        blob = torch.empty(max(total, 1), dtype=torch.uint8, device=dev)
        blob[:total].copy_(host, non_blocking=True)
        zs = [blob[o:o + ln] for o, ln in spans[0::2]]
        es = [blob[o:o + ln] for o, ln in spans[1::2]]
        digests = ops.sha256_leaves(zs, es, [lf[1] for lf in live])              # [n,32], stays on the device
        # ---- membership: every leaf of every miner against its own root, ONE launch per path depth ------------------
        by_depth: Dict[int, List[int]] = {}
        odd: Dict[int, List[int]] = {}
        for i, lf in enumerate(live):
            if len(subs[lf[0]].merkle_root) == 32 and all(len(p) == 32 for p in lf[4]):
                by_depth.setdefault(len(lf[4]), []).append(i)
            else:  # nodes that cannot come from this protocol: the generic per-miner path (same answer as hashlib)
                odd.setdefault(lf[0], []).append(i)
        for depth, idxs in by_depth.items():
            sel = torch.tensor(idxs, dtype=torch.long, device=dev)
            pb = b"".join(b"".join(live[i][4]) for i in idxs)
            paths = proof._to_dev_bytes(pb, dev).reshape(len(idxs), depth, 32) if depth else \
                torch.empty((len(idxs), 0, 32), dtype=torch.uint8, device=dev)
            roots = proof._to_dev_bytes(b"".join(subs[live[i][0]].merkle_root for i in idxs), dev).reshape(len(idxs), 32)
            res = ops.merkle_verify_roots(digests.index_select(0, sel).contiguous(), paths, roots).cpu().numpy()
            for i, r in zip(idxs, res):
                if not r:
                    logger.warning("[%s] Merkle verification failed for miner %s at t=%s", request_id,
                                   subs[live[i][0]].uid, live[i][1])
                    ok[live[i][0]] = False
        if odd:
            host_digests = digests.cpu().numpy()
            for m, idxs in odd.items():
                try:
                    res = proof.verify_merkle_leaves([host_digests[i].tobytes() for i in idxs], [live[i][4] for i in idxs],
                                                     subs[m].merkle_root)
                except Exception as ex:
                    logger.error("[%s] Merkle verification of miner %s failed: %r", request_id, subs[m].uid, ex)
                    res = [False]
                if not all(res):
                    logger.warning("[%s] Merkle verification failed for miner %s", request_id, subs[m].uid)
                    ok[m] = False
        # ---- run_unet_step verdicts: one launch per (scheduler length, latent shape) --------------------------------
        groups: Dict[Tuple, List[int]] = {}
        frames = {}
        for i, lf in enumerate(live):
            m, _, z, e, _, step, alphas, cfg = lf
            if not ok[m]:
                continue
            try:
                if len(e) != len(z):
                    raise ValueError("eps and z byte lengths differ")
                frames[i] = (proof._frame0_f16(z, cfg), proof._frame0_f16(e, cfg))
                key = (len(alphas), int(cfg["latent_channels"]), int(cfg["latent_height"]), int(cfg["latent_width"]))
                groups.setdefault(key, []).append(i)
            except Exception as ex:
                logger.error("[%s] Error running UNet step %s of miner %s: %s", request_id, step, subs[m].uid, ex)
                ok[m] = False
        for key, idxs in groups.items():
            cfg = live[idxs[0]][7]
            try:
                flags = proof.spotcheck_flags([frames[i][0] for i in idxs], [frames[i][1] for i in idxs],
                                              [None] * len(idxs), [live[i][5] % key[0] for i in idxs], key[0], cfg)
            except Exception as ex:
                logger.error("[%s] step verification launch failed: %r", request_id, ex)
                flags = [0] * len(idxs)
            for i, f in zip(idxs, flags):
                if not (f & proof._FLAG_FINITE and f & proof._FLAG_BOUNDED):
                    logger.warning("[%s] step verification failed for miner %s at t=%s", request_id, subs[live[i][0]].uid,
                                   live[i][1])
                    ok[live[i][0]] = False
    # ---- temporal coherence: host plans for every miner, ONE launch per scheduler length, verdicts in order ---------
    plans, cgroups = {}, {}
    for m, sub in enumerate(subs):
        if len(sub.spot_check_indices) < 2:
            continue
        t2s, alphas = meta[m]
        try:
            plan = proof._coherence_plan(sub.spot_check_indices, sub.leaves_result, t2s, alphas, sub.unet_config,
                                         request_id, logger)
        except Exception as ex:
            logger.error("[%s] temporal coherence of miner %s failed: %r", request_id, sub.uid, ex)
            ok[m] = False
            continue
        if plan[0] is not None:
            ok[m] = ok[m] and plan[0]
            continue
        plans[m] = plan
        cfg = sub.unet_config
        key = (len(alphas), int(cfg["latent_channels"]), int(cfg["latent_height"]), int(cfg["latent_width"]))
        cgroups.setdefault(key, []).append(m)
    for key, ms in cgroups.items():
        zs, es, zn, steps, owner = [], [], [], [], []
        for m in ms:
            _, a, b, c, st, _, _ = plans[m]
            zs += a; es += b; zn += c; steps += st
            owner += [m] * len(a)
        try:
            flags = proof.spotcheck_flags(zs, es, zn, steps, key[0], subs[ms[0]].unet_config) if zs else []
        except Exception as ex:
            logger.error("[%s] temporal coherence launch failed: %r", request_id, ex)
            for m in ms:
                ok[m] = False
            continue
        for m in ms:
            mine = [f for f, o in zip(flags, owner) if o == m]
            if not proof._coherence_verdict(mine, plans[m][5], plans[m][6], request_id, logger):
                logger.warning("[%s] temporal coherence check failed for miner %s", request_id, subs[m].uid)
                ok[m] = False
    return ok
