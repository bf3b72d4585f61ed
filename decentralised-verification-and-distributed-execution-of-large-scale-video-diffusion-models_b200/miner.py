"""The miner's recording loop and commitment (InferNet/neurons/miner.py:571-589 + 180-264), on libb2v.

Reference loop, batch 1, NO classifier-free guidance, fp16 latents::

    for t in scheduler.timesteps:
        latents.append(z); timesteps.append(t.item())
        eps = pipe.unet(z, t, encoder_hidden_states=ehs).sample
        noise_preds.append(eps)
        z = scheduler.step(eps, t, z).prev_sample
    root, leaf_data = self._build_merkle_tree(latents, noise_preds, timesteps)   # after the loop, on the CPU

Here every step is one UNet forward (CUDA-graph replay) plus ONE fused kernel (``b2v_cfg_ddim_step`` with guidance off)
that performs the DDIM update and writes the step's two leaf images (z_t and eps_t as little-endian fp16, the exact bytes
the reference hashes) into a preallocated trajectory buffer. Nothing is copied to the host unless the caller asks for
the reference's ``leaf_data`` byte images.

When to hash (``commit=``): a leaf is ONE serial 3.5 MB Merkle-Damgard chain (~55 ms at ZeroscopeXL) and chains only
parallelise ACROSS leaves, so
* ``"end"`` (default): all n leaves are hashed together after the loop, one CTA per leaf on n SMs — ~55 ms for the whole
  50-step trajectory, ~1 % of the denoise time;
* ``"loop"``: leaf i is hashed on a side stream under the UNet of step i+1 by one shared-memory-free warp
  (``b2v_sha256_leaves`` mode 3, so it fits beside the UNet's resident one-CTA-per-SM kernels). The digest is ready one
  step after its leaf, but the warp shares an SM with one CTA of every persistent GEMM and that CTA paces the kernel:
  measured +15 % per step (bench.py ``merkle.in_loop``). Use it only when the commitment must stream out during the run.
"""
import time
from typing import Dict, List, Optional, Tuple

import torch

from . import ops, proof


class Trajectory:
    """What the recording loop leaves behind: device-resident fp16 leaf images and the per-leaf digests."""

    def __init__(self, latents, noise_preds, timesteps, alphas, final, digests):
        self.latents: List[torch.Tensor] = latents          # z_t before the step, fp16 [1,C,F,h,w]
        self.noise_preds: List[torch.Tensor] = noise_preds  # eps_t, fp16
        self.timesteps: List[int] = timesteps
        self.alphas: List[float] = alphas                   # alphas_cumprod[t] as Python floats (miner.py:536-543)
        self.final: torch.Tensor = final                    # z after the last step (goes to the VAE)
        self.digests: Optional[torch.Tensor] = digests      # uint8 [n,32] on the device, or None (commit=False)

    def merkle(self, with_bytes=True) -> Tuple[bytes, Dict[int, Tuple[bytes, bytes, List[bytes]]]]:
        """(root, {t: (z_bytes, eps_bytes, [sibling, ...])}) exactly like ``Miner._build_merkle_tree``."""
        if self.digests is None:
            return proof.build_merkle_tree(self.latents, self.noise_preds, self.timesteps, with_bytes=with_bytes)
        n = len(self.timesteps)
        fetch = proof.leaf_bytes_to_host(self.latents + self.noise_preds) if with_bytes else None
        root, paths = ops.merkle_build(self.digests)
        plist = proof._paths_as_lists(paths, n)
        blobs = fetch() if fetch is not None else [b""] * (2 * n)
        leaf_data = {int(t): (blobs[i], blobs[n + i], plist[i]) for i, t in enumerate(self.timesteps)}
        return root.cpu().numpy().tobytes(), leaf_data

    def wire(self) -> Dict:
        """The base64 fields the miner puts on the synapse (miner.py:643-666): ``merkle_root``, ``timesteps``, ``latents``,
        ``noise_preds``, ``alphas`` and ``leaf_data`` = {t: (z_b64, eps_b64, [sibling_b64, ...])}. The reference encodes
        every leaf image twice (leaf_data and latents / noise_preds); here each is encoded once, on the device, and the
        two views share the string."""
        import base64
        root, leaf = self.merkle(with_bytes=False)
        zs, es = leaves_b64(self.latents), leaves_b64(self.noise_preds)
        leaf_b64 = {int(t): (zs[i], es[i], [base64.b64encode(p).decode("ascii") for p in leaf[int(t)][2]])
                    for i, t in enumerate(self.timesteps)}
        return {"merkle_root": base64.b64encode(root).decode("ascii"), "timesteps": list(self.timesteps),
                "latents": zs, "noise_preds": es, "alphas": list(self.alphas), "leaf_data": leaf_b64}


    def reveal(self, indices) -> Dict[int, Optional[Tuple[str, str, List[str]]]]:
        """Only the requested leaves, in the wire form of ``open_leaves`` / ``forward_reveal_leaves`` (miner.py:813-888):
        {i: (z_b64, eps_b64, [sibling_b64, ...]) or None}. As in the reference ``i`` is looked up among the KEYS of
        leaf_data, i.e. the committed timesteps. Only the requested leaf images are base64-encoded (on the device) and
        cross PCIe: k of n leaves cost k/n of ``wire()`` — at ZeroscopeXL 9.4 MB of text per leaf instead of 472 MB for
        a 50-step trajectory."""
        import base64
        indices = [int(i) for i in indices]
        pos = {int(t): k for k, t in enumerate(self.timesteps)}
        hit = [i for i in dict.fromkeys(indices) if i in pos]
        out: Dict[int, Optional[Tuple[str, str, List[str]]]] = {i: None for i in indices}
        if not hit:
            return out
        digests = self.digests
        if digests is None:
            digests = ops.sha256_leaves(self.latents, self.noise_preds, prefixes=self.timesteps)
            self.digests = digests
        _, paths = ops.merkle_build(digests)
        plist = proof._paths_as_lists(paths, len(self.timesteps))
        zs = leaves_b64([self.latents[pos[i]] for i in hit])
        es = leaves_b64([self.noise_preds[pos[i]] for i in hit])
        for k, i in enumerate(hit):
            out[i] = (zs[k], es[k], [base64.b64encode(pp).decode("ascii") for pp in plist[pos[i]]])
        return out


class LeafStore:
    """The miner's ``_leaf_data_by_request`` (miner.py:131, 636): committed trajectories parked per
    ``(caller_hotkey, request_id)`` until the validator asks for leaves, DEVICE-resident (a 50-step XL trajectory is
    177 MB of HBM; the reference keeps the same bytes on the host) and deleted after one reveal like the reference does
    (miner.py:420, 841, 877).

    Unlike the reference's dict the store is BOUNDED: a request that no validator ever spot-checks would otherwise stay
    in HBM for the life of the process, until the UNet workspace or the VAE can no longer be allocated. ``put`` drops
    the oldest entries beyond ``max_entries`` (default 64: about 11 GB of XL trajectories) and, with ``ttl_s``, those
    older than that many seconds; a reveal for a dropped request answers like one for an unknown request (every index
    -> None), which is what the reference answers for a request it does not hold. ``evicted`` counts the drops."""

    def __init__(self, max_entries: Optional[int] = 64, ttl_s: Optional[float] = None, clock=time.monotonic):
        if max_entries is not None and int(max_entries) < 1:
            raise ValueError("LeafStore: max_entries must be >= 1 or None")
        self.max_entries = None if max_entries is None else int(max_entries)
        self.ttl_s = None if ttl_s is None else float(ttl_s)
        self._clock = clock
        self._born: Dict[Tuple[str, str], float] = {}
        self.evicted = 0
        self._leaf_data_by_request: Dict[Tuple[str, str], Trajectory] = {}   # insertion-ordered: oldest first

    def _drop(self, key):
        self._leaf_data_by_request.pop(key, None)
        self._born.pop(key, None)
        self.evicted += 1

    def _expire(self):
        if self.ttl_s is not None:
            now = self._clock()
            for key in [k for k, t in self._born.items() if now - t > self.ttl_s]:
                self._drop(key)
        if self.max_entries is not None:
            while len(self._leaf_data_by_request) > self.max_entries:
                self._drop(next(iter(self._leaf_data_by_request)))

    def put(self, caller_hotkey, request_id, trajectory: Trajectory):
        key = (caller_hotkey, request_id)
        self._leaf_data_by_request.pop(key, None)          # a re-commit of the same request becomes the newest entry
        self._leaf_data_by_request[key] = trajectory       # miner.py:636
        self._born[key] = self._clock()
        self._expire()

    def __len__(self):
        return len(self._leaf_data_by_request)

    def open_leaves(self, indices, request_id=None, caller_hotkey=None):
        """miner.py:813-842 (the RPC): only the requested leaves; unknown request -> every index maps to None."""
        indices = [int(i) for i in indices]
        if caller_hotkey is None:
            caller_hotkey = "unknown"
        if request_id is None:
            return {i: None for i in indices}
        key = (caller_hotkey, request_id)
        self._expire()
        traj = self._leaf_data_by_request.get(key)
        if not traj:
            return {i: None for i in indices}
        result = traj.reveal(indices)
        del self._leaf_data_by_request[key]
        self._born.pop(key, None)
        return result

    def forward_reveal_leaves(self, synapse):
        """miner.py:844-888: ``synapse`` is a RevealLeavesSynapse (protocol.py:74-95: indices, request_id,
        caller_hotkey, leaves) or anything with those attributes; never raises."""
        try:
            indices = [int(i) for i in synapse.indices]
            key = (synapse.caller_hotkey, synapse.request_id)
            self._expire()
            traj = self._leaf_data_by_request.get(key)
            if not traj:
                synapse.leaves = {i: None for i in indices}
                return synapse
            result = traj.reveal(indices)
            del self._leaf_data_by_request[key]
            self._born.pop(key, None)
            synapse.leaves = result
            return synapse
        except Exception:
            synapse.leaves = {i: None for i in synapse.indices}
            return synapse

    def spot_check_all(self, caller_hotkey, request_id):
        """The spot-check fast path of ``forward`` (miner.py:352-456: seed 0, empty challenge): the indices are lost in
        serialisation there, so the reference returns EVERY leaf of the request as ``proof['leaf_data']``. Same here
        (this is the expensive form; validators that can should send RevealLeavesSynapse)."""
        empty = {"merkle_root": "", "signature": "", "seed": 0, "challenge": "", "leaf_data": {}}
        if caller_hotkey is None:
            return empty
        key = (caller_hotkey, request_id)
        self._expire()
        traj = self._leaf_data_by_request.get(key)
        if not traj:
            return empty
        try:
            result = traj.reveal(traj.timesteps)
            del self._leaf_data_by_request[key]
            self._born.pop(key, None)
            return dict(empty, leaf_data=result)
        except Exception:
            return empty


def decode_final_frames(pipe, z, reference_view=True):
    """InferNet/neurons/miner.py:593-600: the final latents [B,C,T,h,w] -> ``pipe.vae.decode(flat / scaling).sample``
    over all B*T frames at once -> numpy [T,H,W,3] in the decoder's [-1,1] range (the array ``export_to_video`` gets).

    Reference quirk, reproduced by default because the video bytes enter the signed message (proof.py:56-58):
    ``flat = z.view(B*T, C, H, W)`` REINTERPRETS the channel-major memory of [B,C,T,h,w] as T images — image i is made of
    4 consecutive (channel, frame) planes, not of frame i's 4 channels. ``reference_view=False`` decodes the actual
    frames (``z.permute(0,2,1,3,4)``). ``pipe.vae`` is b2v.vae.AutoencoderKL (or anything with that call surface)."""
    B, Cc, T, H, W = z.shape
    if reference_view:
        flat = z.contiguous().view(B * T, Cc, H, W)
    else:
        flat = z.permute(0, 2, 1, 3, 4).reshape(B * T, Cc, H, W)
    decoded = pipe.vae.decode(flat / pipe.vae.config.scaling_factor).sample
    decoded = decoded.view(B, T, *decoded.shape[-3:])
    return decoded[0].permute(0, 2, 3, 1).float().cpu().numpy()


def leaves_b64(tensors) -> List[str]:
    """``[base64.b64encode(t.cpu().numpy().tobytes()).decode('ascii') for t in tensors]`` (miner.py:655-656) with the
    encoding done on the device: equally sized tensors that sit back to back in one buffer (the trajectory buffers of
    ``record_trajectory``) are encoded by ONE kernel launch and leave the device as ASCII through one pinned DMA."""
    out: List[str] = []
    i = 0
    while i < len(tensors):
        t0 = tensors[i].contiguous()
        nbytes = t0.numel() * t0.element_size()
        j = i + 1
        while (j < len(tensors) and tensors[j].is_contiguous() and tensors[j].dtype == t0.dtype
               and tensors[j].numel() == t0.numel()
               and tensors[j].untyped_storage().data_ptr() == t0.untyped_storage().data_ptr()
               and tensors[j].data_ptr() == t0.data_ptr() + (j - i) * nbytes):
            j += 1
        rows = j - i
        if rows > 1:
            flat = torch.empty(0, dtype=t0.dtype, device=t0.device).set_(t0.untyped_storage(), t0.storage_offset(),
                                                                         (rows, t0.numel()), (t0.numel(), 1))
        else:
            flat = t0.view(1, -1)
        enc = ops.base64_rows(flat)
        blobs = proof.leaf_bytes_to_host([enc[k] for k in range(rows)])()
        out.extend(b.decode("ascii") for b in blobs)
        i = j
    return out


def record_trajectory(pipe, z0, encoder_hidden_states, num_steps, commit="end") -> Trajectory:
    """miner.py:571-589. ``z0`` [1,C,F,h,w] (the reference draws fp16 noise on the device; parity tests pass it in),
    ``encoder_hidden_states`` [1,77,D]; ``commit``: "end" | "loop" | False (True = "end"). Returns a :class:`Trajectory`."""
    if z0.dim() != 5 or z0.shape[0] != 1:
        raise ValueError("the miner path is batch 1: z0 must be [1,C,F,h,w]")
    if commit is True:
        commit = "end"
    if commit not in ("end", "loop", False, None):
        raise ValueError("commit must be 'end', 'loop' or False")
    sched, unet = pipe.scheduler, pipe.unet
    sched.set_timesteps(num_steps)
    ts = [int(t) for t in sched.host_timesteps()]
    alphas = [float(sched.alphas_cumprod[t]) for t in ts]
    dev = z0.device
    n = len(ts)
    # trajectory buffers: the leaf images of all steps, allocated once (XL, 50 steps: 2 x 88 MB)
    zbuf = torch.empty((n,) + tuple(z0.shape), dtype=torch.float16, device=dev)
    ebuf = torch.empty((n,) + tuple(z0.shape), dtype=torch.float16, device=dev)
    in_loop = commit == "loop"
    digests = torch.empty((n, 32), dtype=torch.uint8, device=dev) if in_loop else None
    main = torch.cuda.current_stream(dev)
    side = torch.cuda.Stream(device=dev) if in_loop else None
    lat = z0.contiguous()
    x = ops.make_unet_input(lat, 1)
    eps = torch.empty_like(x)
    emb = encoder_hidden_states.contiguous()
    for i, t in enumerate(ts):
        unet.forward_into(x, t, emb, eps)
        lat, x = sched.fused_cfg_step(lat, eps, t, None, n_dup=1, leaf_z=zbuf[i], leaf_eps=ebuf[i], xnext=x)
        if in_loop:
            # leaf i is complete once the step kernel has run: hash it under the next UNet forward
            side.wait_stream(main)
            with torch.cuda.stream(side):
                digests[i:i + 1].copy_(ops.sha256_leaves([zbuf[i]], [ebuf[i]], prefixes=[t], mode=3))
    zs, es = [zbuf[i] for i in range(n)], [ebuf[i] for i in range(n)]
    if in_loop:
        main.wait_stream(side)
    elif commit == "end":
        digests = ops.sha256_leaves(zs, es, prefixes=ts)  # n chains side by side, one CTA each
    return Trajectory(zs, es, ts, alphas, lat, digests)


# ---------------------------------------------------------------------------------------------------------------------
# Validator side: spot-check by RE-EXECUTION (BASELINE config 5 "10 % randomised spot-check
# re-execution"). The reference's own check (proof.py:307-392, `run_unet_step`) never runs a UNet despite its name: it
# replays one scheduler step on frame 0 and tests the result for finiteness. With a forward that costs ~0.1 s and is
# bit-reproducible, the validator can afford the real thing for the k revealed leaves.
# ---------------------------------------------------------------------------------------------------------------------
def reexecute_spot_check(pipe, encoder_hidden_states, revealed: Dict[int, Tuple[bytes, bytes, List[bytes]]],
                         merkle_root: bytes, timesteps: List[int], latent_shape, num_steps: int,
                         rel_tol: float = 2e-2, rank: int = 0, world: int = 1) -> Dict:
    """Check revealed leaves {t: (z_bytes, eps_bytes, path)} of a committed trajectory three ways:
      1. membership: SHA256(be16(t) ‖ z ‖ eps), hashed on the device, folds up the sibling path to `merkle_root`;
      2. re-execution: eps' = UNet(z_t, t, text) must reproduce the revealed eps_t (rel-l2 <= rel_tol; on the same build
         and GPU generation the forward is bit-reproducible and `eps_exact` reports equality);
      3. chaining: when the next timestep is revealed too, DDIM(z_t, eps_t, t) must be its z.
    `timesteps` is the committed schedule (order matters for 3); `rank`/`world` shard the revealed leaves over ranks
    (independent units, no communication). Returns {"ok": bool, "leaves": {t: {...}}}; never raises on bad proofs."""
    sched, unet = pipe.scheduler, pipe.unet
    sched.set_timesteps(num_steps)
    dev = encoder_hidden_states.device
    emb = encoder_hidden_states.contiguous()
    n_el = 1
    for d in latent_shape:
        n_el *= int(d)
    mine = sorted(revealed)[rank::world]
    order = {int(t): i for i, t in enumerate(timesteps)}
    report = {}
    for t in mine:
        entry = {"member": False, "eps_rel_l2": None, "eps_exact": False, "reexec": False, "chain": None}
        report[int(t)] = entry
        try:
            zb, eb, path = revealed[t]
            if len(zb) != 2 * n_el or len(eb) != 2 * n_el or int(t) not in order:
                continue
            z = torch.frombuffer(bytearray(zb), dtype=torch.float16).to(dev).view(tuple(latent_shape))
            e = torch.frombuffer(bytearray(eb), dtype=torch.float16).to(dev).view(tuple(latent_shape))
            digest = ops.sha256_leaves([z], [e], prefixes=[int(t)])
            leaf_hash = digest[0].cpu().numpy().tobytes()
            entry["member"] = bool(proof.verify_merkle_leaf(leaf_hash, list(path), merkle_root))
            x = ops.make_unet_input(z, 1)
            eps = torch.empty_like(x)
            unet.forward_into(x, int(t), emb.to(z.dtype), eps)
            diff = (eps.float() - e.float()).norm() / e.float().norm().clamp_min(1e-12)
            entry["eps_rel_l2"] = float(diff)
            entry["eps_exact"] = bool(torch.equal(eps, e))
            entry["reexec"] = bool(torch.isfinite(eps).all()) and entry["eps_rel_l2"] <= rel_tol
            i = order[int(t)]
            if i + 1 < len(timesteps) and int(timesteps[i + 1]) in revealed:
                zn_b = revealed[int(timesteps[i + 1])][0]
                if len(zn_b) == 2 * n_el:
                    zn = torch.frombuffer(bytearray(zn_b), dtype=torch.float16).to(dev).view(tuple(latent_shape))
                    nxt, _ = sched.fused_cfg_step(z, e, int(t), None, n_dup=0)
                    cd = (nxt.float() - zn.float()).norm() / zn.float().norm().clamp_min(1e-12)
                    entry["chain"] = bool(float(cd) <= 1e-3)  # the same fp16 step on the revealed bytes: rounding only
                else:
                    entry["chain"] = False
        except Exception as ex:  # a malformed proof is a failed proof, like proof.py:390-392
            entry["error"] = repr(ex)
    ok = bool(report) and all(v["member"] and v["reexec"] and v["chain"] is not False for v in report.values())
    return {"ok": ok, "leaves": report}
