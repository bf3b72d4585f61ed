"""Proof-of-Inference commit / verify on libb2v — same function names, argument meaning and error behaviour as
``InferNet/template/validator/proof.py`` (and ``Miner._build_merkle_tree`` of ``InferNet/neurons/miner.py:180-264``).

Device work (CUDA kernels of csrc/sha256.cu and csrc/step.cu through the C-ABI):
  * SHA-256 leaf hashing of the per-step latents in place on the device            b2v_sha256_leaves
  * tree levels + sibling paths, batched leaf verification                         b2v_merkle_build / _verify
  * the default-config DDIM re-check of revealed leaves (frame 0, fp16 per-op)     b2v_spotcheck_ddim
Host work (kept in Python like the reference): HMAC seed derivation, ``random.Random`` spot-check sampling, base64,
sr25519 signature verification (a callback: bittensor is the reference's dependency, not ours).

Like the reference, the verify-side functions never raise (``derive_seed`` excepted): errors are logged and turn into
``False``. There is no CPU fallback: without libb2v / an sm_100 device they log the error and return ``False``.
"""
import base64
import hashlib
import hmac
import logging
import random
import threading
from typing import Callable, Dict, List, Optional, Tuple

import numpy as np
import torch

from . import ops
from .scheduler import DDIMScheduler

LOG = logging.getLogger("b2v.proof")

# sr25519 verification hook: callable(hotkey_ss58, message, signature) -> bool. The reference uses bt.Keypair
# (proof.py:34-105); install one with set_signature_verifier() or pass `signature_verifier=` explicitly.
_signature_verifier: Optional[Callable[[str, bytes, bytes], bool]] = None


def set_signature_verifier(fn):
    global _signature_verifier
    _signature_verifier = fn


def _device():
    if not torch.cuda.is_available():
        raise ops._lib.B2VError("no CUDA device: libb2v has no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _to_dev_bytes(b, dev):
    if len(b) == 0:
        return torch.empty(0, dtype=torch.uint8, device=dev)
    return torch.frombuffer(bytearray(b), dtype=torch.uint8).to(dev)


def _rows(t):
    """CUDA uint8 [n,32] -> list of n bytes objects."""
    a = t.cpu().numpy()
    return [a[i].tobytes() for i in range(a.shape[0])]


# ---------------------------------------------------------------------------------------------------------------
# proof.py:11-18
# ---------------------------------------------------------------------------------------------------------------
def derive_seed(challenge: bytes, wallet_address: str) -> int:
    try:
        return int.from_bytes(hmac.new(wallet_address.encode(), challenge, hashlib.sha256).digest()[:8], byteorder="big")
    except Exception as e:
        LOG.error("Error deriving seed: %s", e)
        raise


# ---------------------------------------------------------------------------------------------------------------
# proof.py:250-288 build_merkle_root / proof.py:291-304 verify_merkle_leaf
# ---------------------------------------------------------------------------------------------------------------
def _paths_as_lists(paths, n):
    if paths is None or paths.shape[1] == 0:
        return [[] for _ in range(n)]
    a = paths.cpu().numpy()
    return [[a[i, k].tobytes() for k in range(a.shape[1])] for i in range(n)]


def build_merkle_root(leaves: List[bytes]) -> Tuple[bytes, List[List[bytes]]]:
    """Raw leaves -> (root, per-leaf sibling paths). Leaves are uploaded once as one buffer and hashed on the device."""
    if not leaves:
        return b"", []
    dev = _device()
    blob = _to_dev_bytes(b"".join(leaves), dev)
    views, off = [], 0
    for x in leaves:
        views.append(blob[off:off + len(x)])
        off += len(x)
    digests = ops.sha256_leaves(views)
    root, paths = ops.merkle_build(digests)
    return root.cpu().numpy().tobytes(), _paths_as_lists(paths, len(leaves))


def verify_merkle_leaves(leaf_hashes: List[bytes], proof_paths: List[List[bytes]], root: bytes) -> List[bool]:
    """Batched verify_merkle_leaf (all paths of equal depth go in one launch)."""
    dev = _device()
    out = [False] * len(leaf_hashes)
    by_depth: Dict[int, List[int]] = {}
    for i, (h, p) in enumerate(zip(leaf_hashes, proof_paths)):
        if len(h) != 32 or any(len(s) != 32 for s in p) or len(root) != 32:
            # hashlib would happily hash other lengths; a 32-byte root can then only match by collision
            out[i] = _verify_host_odd_lengths(h, p, root)
            continue
        by_depth.setdefault(len(p), []).append(i)
    root_d = _to_dev_bytes(root, dev) if len(root) == 32 else None
    for depth, idxs in by_depth.items():
        lh = _to_dev_bytes(b"".join(leaf_hashes[i] for i in idxs), dev).reshape(len(idxs), 32)
        pb = b"".join(b"".join(proof_paths[i]) for i in idxs)
        pt = _to_dev_bytes(pb, dev).reshape(len(idxs), depth, 32)
        ok = ops.merkle_verify(lh, pt, root_d).cpu().numpy()
        for j, i in enumerate(idxs):
            out[i] = bool(ok[j])
    return out


def _verify_host_odd_lengths(leaf_hash, path, root):
    # Nodes that are not 32 bytes cannot come from this protocol; they are folded on the device one pair at a time
    # through the generic leaf hasher so the answer still equals the reference's.
    dev = _device()
    cur = leaf_hash
    for sib in path:
        pair = cur + sib if cur < sib else sib + cur
        cur = ops.sha256_leaves([_to_dev_bytes(pair, dev)]).cpu().numpy()[0].tobytes()
    return cur == root


def verify_merkle_leaf(leaf_hash: bytes, proof_path: List[bytes], root: bytes) -> bool:
    try:
        return verify_merkle_leaves([leaf_hash], [proof_path], root)[0]
    except Exception as e:
        LOG.error("Error verifying Merkle leaf: %s", e)
        return False


# ---------------------------------------------------------------------------------------------------------------
# miner.py:180-264 Miner._build_merkle_tree — commit side
# ---------------------------------------------------------------------------------------------------------------
def commit_trajectory_device(latents, noise_preds, timesteps, mode=0):
    """Device-resident commit: per-step fp16 CUDA tensors -> (root uint8[32], digests [n,32], paths [n,depth,32]),
    all on the device, nothing copied to the host. Leaf i = SHA256(be16(t_i) ‖ bytes(z_i) ‖ bytes(eps_i))."""
    if not (len(latents) == len(noise_preds) == len(timesteps)):
        raise ValueError("latents, noise_preds and timesteps must have the same length")
    if not latents:
        raise ValueError("empty trajectory")
    zs = [z.contiguous() for z in latents]
    es = [e.contiguous() for e in noise_preds]
    digests = ops.sha256_leaves(zs, es, prefixes=[int(t) for t in timesteps], mode=mode)
    root, paths = ops.merkle_build(digests)
    return root, digests, paths


_PINNED = threading.local()


def _pinned(nbytes):
    """Page-locked staging buffer, grown on demand and reused (pinning 177 MB costs more than copying it). One per
    thread: the miner serves requests from several axon worker threads without a lock of its own (miner.py:571-600)."""
    buf = getattr(_PINNED, "buf", None)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, pin_memory=True)
        _PINNED.buf = buf
    return buf


def leaf_bytes_to_host(tensors, stream=None):
    """Byte images of CUDA tensors as Python ``bytes`` (what ``z.cpu().numpy().tobytes()`` returns in miner.py:209-214)
    through ONE pinned staging buffer: adjacent tensors (views of one trajectory buffer) travel as a single DMA, all
    copies are asynchronous on `stream`, and the host slices them apart afterwards. Returns a closure that waits for
    the copies and yields the list — so the caller can hash on the device while the bytes are in flight."""
    if not tensors:
        return lambda: []
    dev = tensors[0].device
    flat = [t.contiguous().view(-1).view(torch.uint8) for t in tensors]
    sizes = [f.numel() for f in flat]
    host = _pinned(sum(sizes))
    stream = stream or torch.cuda.Stream(device=dev)
    stream.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(stream):
        i, off = 0, 0
        while i < len(flat):  # coalesce runs that are contiguous in device memory
            j, run = i + 1, sizes[i]
            base = flat[i].untyped_storage()
            while (j < len(flat) and flat[j].data_ptr() == flat[i].data_ptr() + run
                   and flat[j].untyped_storage().data_ptr() == base.data_ptr()):  # same allocation only
                run += sizes[j]
                j += 1
            if j == i + 1:
                src = flat[i]
            else:
                src = torch.empty(0, dtype=torch.uint8, device=dev).set_(flat[i].untyped_storage(),
                                                                        flat[i].storage_offset(), (run,), (1,))
            host[off:off + run].copy_(src, non_blocking=True)
            off += run
            i = j
    done = torch.cuda.Event()
    done.record(stream)

    keep = [flat, tensors]  # the sources stay referenced until the DMA has finished

    def finish():
        done.synchronize()
        keep.clear()
        mv = memoryview(host.numpy())
        out, off = [], 0
        for n in sizes:
            out.append(bytes(mv[off:off + n]))
            off += n
        return out
    return finish


def build_merkle_tree(latents, noise_preds, timesteps, with_bytes=True):
    """Drop-in for ``Miner._build_merkle_tree``: returns (root bytes, {t: (z_bytes, eps_bytes, [sibling, ...])}).
    The hashing happens on the device in place; the byte images are copied to the host only because the reference's
    return value contains them (with_bytes=False leaves them empty) — as one pinned DMA on a side stream UNDER the
    hashing, and the host-side slicing into ``bytes`` also runs under it."""
    fetch = None
    if with_bytes:
        fetch = leaf_bytes_to_host([z.contiguous() for z in latents] + [e.contiguous() for e in noise_preds])
    root, _, paths = commit_trajectory_device(latents, noise_preds, timesteps)   # launches only: nothing waited for yet
    n = len(timesteps)
    # the DMA (a few ms) ends long before the serial SHA-256 chains (54 ms for 50 XL leaves): cut the staging buffer into
    # the per-leaf bytes objects NOW, while the device is still hashing, and only then wait for the tree
    blobs = fetch() if fetch is not None else [b""] * (2 * n)
    plist = _paths_as_lists(paths, n)
    leaf_data = {}
    for i, t in enumerate(timesteps):
        leaf_data[int(t)] = (blobs[i], blobs[n + i], plist[i])
    return root.cpu().numpy().tobytes(), leaf_data


# ---------------------------------------------------------------------------------------------------------------
# proof.py:397-416
# ---------------------------------------------------------------------------------------------------------------
def commit_then_reveal_merkle_spotcheck(merkle_root: bytes, num_leaves: int, num_to_reveal: int,
                                        random_seed: int) -> List[int]:
    rng = random.Random(random_seed)
    if num_leaves < 2:
        return list(range(min(num_to_reveal, num_leaves)))
    last_start = num_leaves - 2
    picked = rng.sample(range(last_start + 1), min(num_to_reveal, last_start + 1))
    reveal: List[int] = []
    for s in picked:
        reveal.extend((s, s + 1))
    return reveal


# ---------------------------------------------------------------------------------------------------------------
# proof.py:307-392 run_unet_step / proof.py:419-626 verify_temporal_coherence_spot_check
# (the branch the reference takes when diffusers imports: a default-config DDIMScheduler, frame 0 only)
# ---------------------------------------------------------------------------------------------------------------
_FLAG_FINITE, _FLAG_BOUNDED, _FLAG_CLOSE = 1, 2, 4


def _frame0_f16(buf: bytes, cfg: dict) -> np.ndarray:
    """fp16 [C,H,W] of frame 0, with the reference's shape rules (proof.py:321-343): raises where it would raise."""
    x = np.frombuffer(buf, dtype=np.float16)
    C, H, W = int(cfg["latent_channels"]), int(cfg["latent_height"]), int(cfg["latent_width"])
    chw = C * H * W
    if x.size > chw:
        nf = x.size // chw
        if x.size != chw * nf:
            raise ValueError("shape '[1, %d, %d, %d, %d]' is invalid for input of size %d" % (C, nf, H, W, x.size))
        return np.ascontiguousarray(x.reshape(C, nf, H, W)[:, 0])
    if x.size != chw:
        raise ValueError("shape '[%d, %d, %d]' is invalid for input of size %d" % (C, H, W, x.size))
    return x.reshape(C, H, W)


def _default_scheduler(n_steps: int) -> DDIMScheduler:
    s = DDIMScheduler()
    s.set_timesteps(n_steps)
    return s


def spotcheck_flags(z_list, eps_list, znext_list, step_indices, n_steps, cfg):
    """Batched device re-check. z/eps/znext: lists of fp16 [C,H,W] numpy arrays (znext entries may be None ->
    coherence bit set). Returns a list of flag bytes (bit0 finite, bit1 |prev|<10, bit2 allclose(prev, z_next))."""
    dev = _device()
    sched = _default_scheduler(n_steps)
    ts = sched.host_timesteps()
    pairs = [sched.alpha_pair(ts[i]) for i in step_indices]
    C, H, W = int(cfg["latent_channels"]), int(cfg["latent_height"]), int(cfg["latent_width"])
    m = len(z_list)
    stack = np.stack(list(z_list) + list(eps_list) + [zn if zn is not None else z_list[i]
                                                      for i, zn in enumerate(znext_list)]).astype(np.float16, copy=False)
    d = torch.from_numpy(stack).to(dev)
    zs, es, zn = [d[i] for i in range(m)], [d[m + i] for i in range(m)], [d[2 * m + i] for i in range(m)]
    flags = ops.spotcheck_ddim(zs, es, zn, [p[0] for p in pairs], [p[1] for p in pairs], C, 1, H * W).cpu().numpy()
    out = []
    for i in range(m):
        f = int(flags[i])
        if znext_list[i] is None:
            f |= _FLAG_CLOSE
        out.append(f)
    return out


def run_unet_step(z_bytes: bytes, eps_bytes: bytes, config: dict, step_i: int) -> bool:
    """Re-derive the DDIM step of one revealed leaf and check it is finite and bounded (despite its name the
    reference executes no UNet here)."""
    try:
        alphas = config["alphas"]
        if step_i >= len(alphas):
            LOG.error("Step index %s out of bounds for alphas array (length: %d)", step_i, len(alphas))
            return False
        z = _frame0_f16(z_bytes, config)
        # the reference views eps with z's frame count: a size mismatch raises there too
        if len(eps_bytes) != len(z_bytes):
            raise ValueError("eps and z byte lengths differ")
        eps = _frame0_f16(eps_bytes, config)
        if step_i < -len(alphas):
            raise IndexError("step index %d outside the scheduler's %d timesteps" % (step_i, len(alphas)))
        f = spotcheck_flags([z], [eps], [None], [step_i % len(alphas)], len(alphas), config)[0]
        return bool(f & _FLAG_FINITE) and bool(f & _FLAG_BOUNDED)
    except Exception as e:
        LOG.error("Error running UNet step %s: %s", step_i, e)
        return False


def _coherence_plan(spot_check_indices, leaves_result, timestep_to_step, proof_alphas, unet_config, request_id, logger):
    """Host half of the temporal-coherence check (proof.py:419-626): everything the reference decides before any
    arithmetic. Returns (verdict, zs, es, zn, steps, order, early_false): `verdict` is the final answer when no device
    work is needed, else None and the fp16 frame-0 arrays of every consecutive pair to settle in one launch."""
    if len(spot_check_indices) < 2:
        return True, [], [], [], [], [], False
    leaf = {}
    for t in spot_check_indices:
        item = leaves_result.get(str(t))
        if item and len(item) == 3:
            leaf[t] = (item[0], item[1])
        else:
            logger.warning("[%s] Missing or malformed leaf for timestep %s" % (request_id, t))
    if len(leaf) < 2:
        return True, [], [], [], [], [], False
    order = [t for t in spot_check_indices if t in leaf]
    if len(order) > len(proof_alphas):
        order = order[:len(proof_alphas)]
    try:
        _default_scheduler(len(proof_alphas))
    except Exception as e:
        logger.error("[%s] Error setting up scheduler: %s" % (request_id, e))
        return False, [], [], [], [], order, False
    # Collect every consecutive pair, stop at the first one the reference would reject before any arithmetic
    zs, es, zn, steps = [], [], [], []
    early_false = False
    for t_i, t_j in zip(order[:-1], order[1:]):
        try:
            z_i = base64.b64decode(leaf[t_i][0])
            e_i = base64.b64decode(leaf[t_i][1])
            z_j = base64.b64decode(leaf[t_j][0])
        except Exception as e:
            logger.error("[%s] Base64 decode error: %s" % (request_id, e))
            early_false = True
            break
        step_i, step_j = timestep_to_step.get(t_i), timestep_to_step.get(t_j)
        if step_i is None or step_j is None:
            early_false = True
            break
        try:
            if len(e_i) != len(z_i):
                raise ValueError("eps and z byte lengths differ")
            a, b, c = _frame0_f16(z_i, unet_config), _frame0_f16(e_i, unet_config), _frame0_f16(z_j, unet_config)
            if not -len(proof_alphas) <= step_i < len(proof_alphas):
                raise IndexError("step index %d outside the scheduler's %d timesteps" % (step_i, len(proof_alphas)))
            step_i %= len(proof_alphas)  # Python negative indexing, as scheduler.timesteps[step_i] would
        except Exception as e:
            logger.error("[%s] Error in temporal coherence computation: %s" % (request_id, e))
            early_false = True
            break
        zs.append(a); es.append(b); zn.append(c); steps.append(step_i)
    return None, zs, es, zn, steps, order, early_false


def _coherence_verdict(flags, order, early_false, request_id, logger):
    """Apply the arithmetic verdicts in order (first failing pair decides), then the host-side early rejection."""
    for k, f in enumerate(flags):
        if not f & _FLAG_CLOSE:
            logger.warning("[%s] Cross-step check failed between t=%s and t=%s" % (request_id, order[k], order[k + 1]))
            return False
    return not early_false


def verify_temporal_coherence_spot_check(spot_check_indices: List[int],
                                         leaves_result: Dict[str, Tuple[str, str, List[str]]],
                                         timestep_to_step: Dict[int, int], proof_alphas: List[float],
                                         unet_config: Dict, request_id: str = None, logger=None) -> bool:
    if logger is None:
        logger = LOG
    verdict, zs, es, zn, steps, order, early_false = _coherence_plan(
        spot_check_indices, leaves_result, timestep_to_step, proof_alphas, unet_config, request_id, logger)
    if verdict is not None:
        return verdict
    try:
        flags = spotcheck_flags(zs, es, zn, steps, len(proof_alphas), unet_config) if zs else []
        return _coherence_verdict(flags, order, early_false, request_id, logger)
    except Exception as e:
        logger.error("[%s] Error in temporal coherence computation: %s" % (request_id, e))
        return False


# ---------------------------------------------------------------------------------------------------------------
# proof.py:21-127 verify_proof_signature / proof.py:130-246 verify_proof_of_inference
# ---------------------------------------------------------------------------------------------------------------
def signed_message(challenge: bytes, seed: int, video_bytes: bytes, merkle_root: bytes) -> bytes:
    """challenge ‖ seed as little-endian u64 ‖ SHA256(video) ‖ merkle_root (proof.py:56-58, miner.py:288-308)."""
    return challenge + int(seed).to_bytes(8, byteorder="little", signed=False) + hashlib.sha256(video_bytes).digest() \
        + merkle_root


def _bittensor_verifier(hotkey_ss58, message, signature):
    import bittensor as bt  # the reference's dependency; absent here -> ImportError -> False (logged by the caller)
    pub = bt.Keypair(ss58_address=hotkey_ss58).public_key
    return bool(bt.Keypair(public_key=pub.hex(), crypto_type=1).verify(message, signature))


def verify_proof_signature(miner_hotkey_ss58: str, challenge: bytes, seed: int, video_bytes: bytes,
                           merkle_root: bytes, signature: bytes, signature_verifier=None) -> bool:
    try:
        message = signed_message(challenge, seed, video_bytes, merkle_root)
        fn = signature_verifier or _signature_verifier or _bittensor_verifier
        try:
            ok = bool(fn(miner_hotkey_ss58, message, signature))
        except Exception as e:
            LOG.error("Error in signature verification: %s", e)
            ok = False
        if not ok:
            LOG.error("SIGNATURE VERIFICATION FAILED")
        return ok
    except Exception as e:
        LOG.error("Signature verification failed with exception: %s", e)
        return False


def _b64(x):
    return base64.b64decode(x) if isinstance(x, str) else x


def verify_proof_of_inference(miner_hotkey_ss58: str, video_bytes: bytes, proof: Dict, config_for_unet: Dict,
                              signature_verifier=None) -> bool:
    try:
        challenge = _b64(proof["challenge"])
        seed = proof["seed"]
        merkle_root = _b64(proof["merkle_root"])
        signature = _b64(proof["signature"])
        if not isinstance(video_bytes, (bytes, bytearray)):
            video_bytes = base64.b64decode(video_bytes)
        if not verify_proof_signature(miner_hotkey_ss58, challenge, seed, video_bytes, merkle_root, signature,
                                      signature_verifier=signature_verifier):
            return False

        decoded: Dict[int, Tuple[bytes, bytes, List[bytes]]] = {}
        for key, (z64, e64, path64) in proof.get("leaf_data", {}).items():
            decoded[int(key)] = (base64.b64decode(z64), base64.b64decode(e64), [base64.b64decode(p) for p in path64])
        order = list(decoded.items())
        t_to_step = {t: i for i, (t, _) in enumerate(order)}
        alphas = proof.get("alphas", [])
        if not alphas:
            alphas = config_for_unet["alphas"]
        step_cfg = {"alphas": alphas, **config_for_unet}

        # every revealed leaf: hash on the device, verify its path, re-check its DDIM step — batched, then the
        # verdicts are applied in the reference's order (first failure wins; all failures mean False anyway)
        if decoded:
            dev = _device()
            ts = [t for t, _ in order]
            zs = [_to_dev_bytes(v[0], dev) for _, v in order]
            es = [_to_dev_bytes(v[1], dev) for _, v in order]
            digests = _rows(ops.sha256_leaves(zs, es, prefixes=ts))
            oks = verify_merkle_leaves(digests, [v[2] for _, v in order], merkle_root)
            for (t, (zb, eb, _)), ok in zip(order, oks):
                if not ok:
                    LOG.warning("Merkle path failed at timestep %s", t)
                    return False
                if not run_unet_step(zb, eb, step_cfg, t_to_step[t]):
                    LOG.warning("UNet step check failed at timestep %s (step %s)", t, t_to_step[t])
                    return False

        if len(decoded) >= 2:
            leaves = {str(t): (base64.b64encode(zb).decode("utf-8"), base64.b64encode(eb).decode("utf-8"), [])
                      for t, (zb, eb, _) in decoded.items()}
            if not verify_temporal_coherence_spot_check(spot_check_indices=[t for t, _ in order],
                                                        leaves_result=leaves, timestep_to_step=t_to_step,
                                                        proof_alphas=alphas, unet_config=config_for_unet,
                                                        request_id="proof_verification", logger=LOG):
                LOG.warning("Temporal coherence check failed in proof verification")
                return False
        return True
    except Exception as e:
        LOG.error("Proof-of-inference verification error: %s", e)
        return False
