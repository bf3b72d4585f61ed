"""Validator spot-check orchestration (InferNet/neurons/validator.py:641-845) over MANY miners.

The reference walks its miners one by one and, per revealed leaf, base64-decodes, hashes on the host with hashlib,
folds the Merkle path, re-derives the DDIM step (``run_unet_step``) and finally checks temporal coherence — about
4 k + 1 small host/CPU jobs per miner. Here the same verdicts come from a handful of device launches for the whole round:

  * ``select_spot_check_pairs``  the index draw of validator.py:641-684 (global ``random``, duplicate timesteps dropped);
  * ``check_revealed_leaves``    one miner, the reference's order of checks (validator.py:768-830);
  * ``check_miners``             all miners at once: the revealed leaves cross PCIe ONCE, as the base64 text they
                                 arrived in (one pinned matrix per text length), ``b2v_base64_decode_rows`` decodes
                                 them on the device, ONE ``b2v_sha256_leaves`` launch hashes every leaf of every miner
                                 in place, ONE ``b2v_merkle_verify_roots`` per path depth checks leaf i against its
                                 own miner's root, and ONE ``b2v_spotcheck_ddim`` launch settles every run_unet_step
                                 verdict and every temporal-coherence pair from the same device buffers.

Both return exactly what a loop over ``check_revealed_leaves`` returns (``tests/test_gpu_validator.py``). Networking
(the ``open_leaves`` RPC), scoring and weights stay the reference's. Nothing here raises: a malformed submission is a
``False`` verdict with the reason logged, as in the reference (its try/except at validator.py:596-1020 scores 0).
"""
import base64
import logging
import random
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
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
            digest = ops.sha256_leaves([proof._to_dev_bytes(z, dev)], [proof._to_dev_bytes(e, dev)], [t])
            h = digest.cpu().numpy()[0].tobytes()
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


def _frames_of(nbytes: int, cfg: Dict) -> int:
    """Frame count of a leaf image under the reference's view rules (proof.py:321-343); raises where it would raise."""
    chw = int(cfg["latent_channels"]) * int(cfg["latent_height"]) * int(cfg["latent_width"])
    if nbytes % 2:
        raise ValueError("buffer size must be a multiple of element size")
    n = nbytes // 2
    if n > chw:
        if n % chw:
            raise ValueError("leaf of %d elements is not a whole number of %d-element frames" % (n, chw))
        return n // chw
    if n != chw:
        raise ValueError("leaf of %d elements is smaller than one %d-element frame" % (n, chw))
    return 1


def _decode_on_device(texts: List[bytes], dev) -> List[Optional[torch.Tensor]]:
    """base64 texts -> device uint8 views of their bytes. Texts of equal length travel as one pinned matrix and are
    decoded by ONE b2v_base64_decode_rows launch per length; rows the kernel flags as non-canonical (and lengths that
    are not a multiple of 4) go through Python's permissive base64.b64decode, whose exceptions give None."""
    out: List[Optional[torch.Tensor]] = [None] * len(texts)
    by_len: Dict[int, List[int]] = {}
    for i, t in enumerate(texts):
        by_len.setdefault(len(t), []).append(i)
    for chars, idxs in by_len.items():
        if chars == 0 or chars % 4:
            flagged = idxs
        else:
            stride = (chars + 15) // 16 * 16
            host = proof._pinned(len(idxs) * stride)[:len(idxs) * stride].view(len(idxs), stride)
            hv = host.numpy()
            for r, i in enumerate(idxs):
                hv[r, :chars] = np.frombuffer(texts[i], dtype=np.uint8)
            txt = torch.empty((len(idxs), stride), dtype=torch.uint8, device=dev)
            txt.copy_(host, non_blocking=True)
            rows, status = ops.base64_decode_rows(txt, chars)
            st = status.cpu().tolist()  # also orders the pinned buffer's reuse behind the copy
            flagged = []
            for r, i in enumerate(idxs):
                if st[r] & 0x100:
                    flagged.append(i)
                else:
                    out[i] = rows[r, :chars // 4 * 3 - (st[r] & 3)]
        for i in flagged:
            try:
                out[i] = proof._to_dev_bytes(base64.b64decode(texts[i]), dev)
            except Exception:
                out[i] = None
    return out


def check_miners(subs: Sequence[Submission], request_id: Optional[str] = None, logger=None) -> List[bool]:
    """Every miner of a round in a handful of launches; element i == check_revealed_leaves(subs[i]). The revealed
    leaves cross PCIe once, as the base64 text they arrived in, and are decoded, hashed and re-checked on the device."""
    logger = logger or LOG
    n = len(subs)
    ok = [True] * n
    dev = proof._device()

    def fail(m, msg, *a):
        logger.warning("[%s] miner %s: " + msg, request_id, subs[m].uid, *a)
        ok[m] = False

    # ---- host pass: structure only (which leaves, which steps); the payload stays text ------------------------------
    leaves = []   # [m, t, z_text, eps_text, path, step]
    meta: List[Tuple[Dict[int, int], List[float]]] = []
    for m, sub in enumerate(subs):
        meta.append(({}, []))
        try:
            t2s = {int(ts): i for i, ts in enumerate(sub.timesteps)}
            alphas = sub.alphas or sub.unet_config["alphas"]
            meta[m] = (t2s, alphas)
            mine = []
            for t in sub.spot_check_indices:
                leaf = sub.leaves_result.get(str(t))
                if not leaf or len(leaf) != 3:
                    fail(m, "missing or malformed leaf for t=%s", t)
                    break
                texts = [x.encode("ascii") if isinstance(x, str) else bytes(x) for x in (leaf[0], leaf[1])]
                path = [base64.b64decode(p) for p in leaf[2]]
                if not 0 <= int(t) < 65536:
                    raise OverflowError("int too big to convert")
                step = t2s.get(t)
                if step is None or step >= len(alphas):
                    fail(m, "invalid step index for t=%s: %s", t, step)
                    break
                mine.append([m, int(t), texts[0], texts[1], path, step])
            if ok[m]:
                leaves += mine
        except Exception as ex:
            fail(m, "spot check failed: %r", ex)
    leaves = [lf for lf in leaves if ok[lf[0]]]
    # ---- decode on the device, validate the shapes from the lengths -------------------------------------------------
    bufs = _decode_on_device([lf[2] for lf in leaves] + [lf[3] for lf in leaves], dev) if leaves else []
    nl = len(leaves)
    views: Dict[Tuple[int, int], Tuple[torch.Tensor, torch.Tensor, int]] = {}   # (miner, t) -> (z, eps, frames)
    for i, lf in enumerate(leaves):
        m, t = lf[0], lf[1]
        z, e = bufs[i], bufs[nl + i]
        try:
            if z is None or e is None:
                raise ValueError("base64 decode error")
            if e.numel() != z.numel():
                raise ValueError("eps and z byte lengths differ")
            views[(m, t)] = (z, e, _frames_of(z.numel(), subs[m].unet_config))
        except Exception as ex:
            fail(m, "leaf t=%s rejected: %s", t, ex)
    live = [lf for lf in leaves if ok[lf[0]]]
    if live:
        # ---- leaf hashes: ONE launch over every revealed leaf of every miner ----------------------------------------
        digests = ops.sha256_leaves([views[(lf[0], lf[1])][0] for lf in live], [views[(lf[0], lf[1])][1] for lf in live],
                                    [lf[1] for lf in live])
        # ---- membership: leaf i against its own miner's root, ONE launch per path depth -----------------------------
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
            res = ops.merkle_verify_roots(digests.index_select(0, sel).contiguous(), paths, roots).cpu().tolist()
            for i, r in zip(idxs, res):
                if not r:
                    fail(live[i][0], "Merkle verification failed at t=%s", live[i][1])
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
                    fail(m, "Merkle verification failed")
    # ---- DDIM re-checks on the device-resident leaves: run_unet_step for every leaf and the temporal-coherence pairs
    #      of every miner, ONE launch per (scheduler length, latent shape, frame count) -------------------------------
    jobs: Dict[Tuple, List[Tuple]] = {}   # key -> [(miner, kind, z, eps, z_next|None, step, label)]
    for lf in live:
        m, t, step = lf[0], lf[1], lf[5]
        if not ok[m]:
            continue
        z, e, F = views[(m, t)]
        cfg, alphas = subs[m].unet_config, meta[m][1]
        key = (len(alphas), int(cfg["latent_channels"]), int(cfg["latent_height"]), int(cfg["latent_width"]), F)
        jobs.setdefault(key, []).append((m, "step", z, e, None, step, t))
    for m, sub in enumerate(subs):
        if not ok[m] or len(sub.spot_check_indices) < 2:
            continue
        t2s, alphas = meta[m]
        cfg = sub.unet_config
        order = list(sub.spot_check_indices)        # every requested leaf is present and well formed here
        if len(order) > len(alphas):
            order = order[:len(alphas)]
        for t_i, t_j in zip(order[:-1], order[1:]):
            (z, e, F), (zj, _, Fj) = views[(m, t_i)], views[(m, t_j)]
            if Fj != F:
                # frame 0 of differently shaped leaves: outside the batched kernel's layout; the per-miner routine decides
                if not proof.verify_temporal_coherence_spot_check(sub.spot_check_indices, sub.leaves_result, t2s, alphas,
                                                                  cfg, request_id, logger):
                    fail(m, "temporal coherence check failed")
                break
            key = (len(alphas), int(cfg["latent_channels"]), int(cfg["latent_height"]), int(cfg["latent_width"]), F)
            jobs.setdefault(key, []).append((m, "pair", z, e, zj, t2s[t_i], (t_i, t_j)))
    for key, items in jobs.items():
        n_steps, C_, H_, W_, F = key
        try:
            sched = proof._default_scheduler(n_steps)
            ts = sched.host_timesteps()
            pairs = [sched.alpha_pair(ts[it[5]]) for it in items]
            flags = ops.spotcheck_ddim([it[2] for it in items], [it[3] for it in items],
                                       [it[4] if it[4] is not None else it[2] for it in items],
                                       [p[0] for p in pairs], [p[1] for p in pairs], C_, F, H_ * W_).cpu().tolist()
        except Exception as ex:
            logger.error("[%s] DDIM re-check launch failed: %r", request_id, ex)
            flags = [0] * len(items)
        for it, f in zip(items, flags):
            if it[1] == "step":
                if not (f & proof._FLAG_FINITE and f & proof._FLAG_BOUNDED):
                    fail(it[0], "step verification failed at t=%s", it[6])
            elif not f & proof._FLAG_CLOSE:
                fail(it[0], "cross-step check failed between t=%s and t=%s", *it[6])
    return ok
