"""ORACLE — test infrastructure only.

Host restatement of the validator's spot-check orchestration (InferNet/neurons/validator.py:641-845)
on top of oracle/merkle_ref.py (which is pinned to the reference's own proof.py by tests/golden/proof_golden.json):
the index draw (validator.py:641-684) and the per-miner sequence of checks (validator.py:768-830).
neurons/validator.py itself cannot be imported here (bittensor, diffusers, wandb at module level) and the reference has
no test for this loop, so the orchestration order is restated from the source; the arithmetic underneath is pinned.
"""
import base64
import random

from . import merkle_ref as M


def select_pairs(all_timesteps, k, rng=random):
    """validator.py:641-684."""
    indices = list(all_timesteps)
    if len(indices) < 2:
        return indices
    max_start = len(indices) - 2
    starts = rng.sample(range(max_start + 1), min(k, max_start + 1))
    picked, used = [], set()
    for s in starts:
        pair = [indices[s], indices[s + 1]]
        if pair[0] not in used and pair[1] not in used:
            picked.extend(pair)
            used.update(pair)
    return picked


def check_revealed_leaves(merkle_root, resp_timesteps, proof_alphas, spot_check_indices, leaves_result, unet_config):
    """validator.py:768-830 with the scheduler branch of run_unet_step / temporal coherence; any exception = False
    (validator.py:1014-1020 scores the miner 0)."""
    try:
        t2s = {int(ts): i for i, ts in enumerate(resp_timesteps)}
        alphas = proof_alphas or unet_config["alphas"]
        ok = True
        for t in spot_check_indices:
            leaf = leaves_result.get(str(t))
            if not leaf or len(leaf) != 3:
                ok = False
                break
            z, e = base64.b64decode(leaf[0]), base64.b64decode(leaf[1])
            path = [base64.b64decode(p) for p in leaf[2]]
            if not M.verify_merkle_leaf(M.leaf_hash(t, z, e), path, merkle_root):
                ok = False
                break
            step = t2s.get(t)
            if step is None or step >= len(alphas):
                ok = False
                break
            cfg = dict(unet_config)
            cfg["alphas"] = alphas
            if not M.run_unet_step_scheduler(z, e, cfg, step):
                ok = False
                break
        if len(spot_check_indices) >= 2 and not M.temporal_coherence(spot_check_indices, leaves_result, t2s, alphas,
                                                                     unet_config):
            ok = False
        return ok
    except Exception:
        return False
