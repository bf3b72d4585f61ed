"""ORACLE — test infrastructure only (tests/, __graft_entry__.smoke(), bench.py's CPU baseline / reference arm).

CPU restatement of the reference's Proof-of-Inference commitment, on the reference's own primitive (hashlib / OpenSSL
SHA-256):
  leaf_hash / commit_trajectory   InferNet/neurons/miner.py:199-212  leaf = t.to_bytes(2,'big') + z_bytes + eps_bytes
  tree_levels / sibling_paths     InferNet/neurons/miner.py:216-249 == template/validator/proof.py:260-286
  build_merkle_root               template/validator/proof.py:250-288 (hashes RAW leaves first; [] -> (b"", []))
  verify_merkle_leaf              template/validator/proof.py:291-304
  derive_seed                     template/validator/proof.py:11-18
  spotcheck_indices               template/validator/proof.py:397-416
  run_unet_step_scheduler         template/validator/proof.py:307-372 (branch taken when diffusers imports)
  run_unet_step_fallback          template/validator/proof.py:374-388 (branch taken when it does not)
  temporal_coherence              template/validator/proof.py:419-626

PINNED: tests/golden/proof_golden.json holds outputs of the reference's proof.py itself, imported from
/root/reference by oracle/gen_golden.py (bittensor stubbed to a logger; for the scheduler branch a stub ``diffusers``
module exposing oracle/ddim_ref.DDIMSchedulerRef — that branch is therefore pinned only up to ddim_ref's own
restatement); tests/test_oracle_merkle.py checks every function here against those vectors.
"""
import base64
import hashlib
import hmac
import random
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch

from .ddim_ref import DDIMSchedulerRef


def sha256(b):
    return hashlib.sha256(b).digest()


def leaf_hash(t, z_bytes, eps_bytes):
    return sha256(int(t).to_bytes(2, "big") + z_bytes + eps_bytes)


def _parent(l, r):
    return sha256(l + r) if l < r else sha256(r + l)


def tree_levels(leaf_hashes):
    levels = [list(leaf_hashes)]
    while len(levels[-1]) > 1:
        cur = levels[-1]
        levels.append([_parent(cur[i], cur[i + 1] if i + 1 < len(cur) else cur[i]) for i in range(0, len(cur), 2)])
    return levels


def sibling_paths(levels):
    out = []
    for idx in range(len(levels[0])):
        path, i = [], idx
        for lv in levels[:-1]:
            s = i ^ 1
            path.append(lv[s] if s < len(lv) else lv[i])
            i //= 2
        out.append(path)
    return out


def build_merkle_root(leaves):
    if not leaves:
        return b"", []
    levels = tree_levels([sha256(x) for x in leaves])
    return levels[-1][0], sibling_paths(levels)


def verify_merkle_leaf(leaf_hash_, path, root):
    cur = leaf_hash_
    for sib in path:
        cur = _parent(cur, sib)
    return cur == root


def commit_trajectory(latents, noise_preds, timesteps, threads=1):
    """Miner._build_merkle_tree: tensors -> (root, {t: (z_bytes, eps_bytes, path)}). threads > 1 hashes leaves on a
    thread pool (hashlib releases the GIL) — the multi-core CPU baseline of bench.py."""
    zb = [z.cpu().numpy().tobytes() for z in latents]
    eb = [e.cpu().numpy().tobytes() for e in noise_preds]
    if threads > 1:
        with ThreadPoolExecutor(threads) as ex:
            hashes = list(ex.map(lambda a: leaf_hash(*a), zip(timesteps, zb, eb)))
    else:
        hashes = [leaf_hash(t, z, e) for t, z, e in zip(timesteps, zb, eb)]
    levels = tree_levels(hashes)
    paths = sibling_paths(levels)
    return levels[-1][0], {t: (zb[i], eb[i], paths[i]) for i, t in enumerate(timesteps)}


def derive_seed(challenge, wallet_address):
    return int.from_bytes(hmac.new(wallet_address.encode(), challenge, hashlib.sha256).digest()[:8], "big")


def spotcheck_indices(num_leaves, num_to_reveal, seed):
    rng = random.Random(seed)
    if num_leaves < 2:
        return list(range(min(num_to_reveal, num_leaves)))
    starts = rng.sample(range(num_leaves - 1), min(num_to_reveal, num_leaves - 1))
    out = []
    for s in starts:
        out += [s, s + 1]
    return out


def _frame0(buf, cfg, shape_from=None):
    """fp16 frame 0 of a leaf buffer. The reference derives the 5-D view from the z buffer and applies the same view
    to eps (proof.py:330-343): pass shape_from=z_bytes for eps. A size mismatch raises, as `.view` does there."""
    x = torch.from_numpy(np.frombuffer(bytearray(buf), dtype=np.float16))
    C, H, W = cfg["latent_channels"], cfg["latent_height"], cfg["latent_width"]
    n = x.numel() if shape_from is None else len(shape_from) // 2
    if n > C * H * W:
        return x.view(1, C, n // (C * H * W), H, W)[0, :, 0]
    return x.view(C, H, W)


def run_unet_step_scheduler(z_bytes, eps_bytes, config, step_i):
    if step_i >= len(config["alphas"]):
        return False
    try:
        z, eps = _frame0(z_bytes, config), _frame0(eps_bytes, config, z_bytes)
        s = DDIMSchedulerRef()
        s.set_timesteps(len(config["alphas"]))
        prev = s.step(eps, s.timesteps[step_i], z).prev_sample
        return bool(torch.isfinite(prev).all() and torch.abs(prev).max() < 10.0)
    except Exception:
        return False


def run_unet_step_fallback(z_bytes, eps_bytes, config, step_i):
    if step_i >= len(config["alphas"]):
        return False
    try:
        z, eps = _frame0(z_bytes, config), _frame0(eps_bytes, config, z_bytes)
        a = torch.tensor(config["alphas"][step_i], dtype=torch.float16)
        pred = (z - torch.sqrt(1.0 - a) * eps) / torch.sqrt(a)
        return bool(torch.isfinite(pred).all() and torch.abs(pred).max() < 10.0)
    except Exception:
        return False


def temporal_coherence(spot_check_indices, leaves_result, timestep_to_step, proof_alphas, unet_config):
    """Scheduler branch of verify_temporal_coherence_spot_check; leaves_result maps str(t) -> (z_b64, eps_b64, path)."""
    if len(spot_check_indices) < 2:
        return True
    data = {}
    for t in spot_check_indices:
        leaf = leaves_result.get(str(t))
        if leaf and len(leaf) == 3:
            data[t] = (leaf[0], leaf[1])
    if len(data) < 2:
        return True
    order = [t for t in spot_check_indices if t in data][:len(proof_alphas)] \
        if len([t for t in spot_check_indices if t in data]) > len(proof_alphas) \
        else [t for t in spot_check_indices if t in data]
    s = DDIMSchedulerRef()
    s.set_timesteps(len(proof_alphas))
    for a, b in zip(order[:-1], order[1:]):
        try:
            z_i = base64.b64decode(data[a][0]); e_i = base64.b64decode(data[a][1]); z_j = base64.b64decode(data[b][0])
        except Exception:
            return False
        si, sj = timestep_to_step.get(a), timestep_to_step.get(b)
        if si is None or sj is None:
            return False
        try:
            prev = s.step(_frame0(e_i, unet_config, z_i), s.timesteps[si], _frame0(z_i, unet_config)).prev_sample
            if not torch.allclose(prev, _frame0(z_j, unet_config), rtol=1, atol=1):
                return False
        except Exception:
            return False
    return True
