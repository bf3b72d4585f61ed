"""-m gpu: reveal of ONLY the requested leaves (InferNet/neurons/miner.py:352-456 spot-check fast path,
813-888 ``open_leaves`` / ``forward_reveal_leaves``; template/protocol.py:74-95 ``RevealLeavesSynapse``).

The store keeps committed trajectories on the device per (caller_hotkey, request_id); a reveal base64-encodes and ships
just the k requested leaves. Checked against the reference's construction on the host (hashlib + base64 on the same
bytes): identical strings, identical paths, the reference's None / delete-after-reveal behaviour, payload k/n of the full
wire form, and the verdicts of the validator round unchanged."""
import base64
import random
import types

import pytest
import torch

pytestmark = pytest.mark.gpu

N_STEPS = 12
SHAPE = (1, 4, 3, 8, 8)


def _trajectory(cuda, seed=0, shape=SHAPE, n=N_STEPS):
    """An honest fp16 trajectory under the validator's library-default DDIM (as tests/test_gpu_validator.py builds one)."""
    from b2v import ops
    from b2v.miner import Trajectory
    from oracle.ddim_ref import default_scheduler
    g = torch.Generator().manual_seed(seed)
    sch = default_scheduler()
    sch.set_timesteps(n)
    ts = [int(1000 - 80 * i - seed % 7) for i in range(n)]
    z = (0.5 * torch.randn(*shape, generator=g)).half()
    zs, es = [], []
    for i in range(n):
        e = (0.05 * torch.randn(*shape, generator=g)).half()
        zs.append(z)
        es.append(e)
        z = sch.step(e, sch.timesteps[i], z).prev_sample.half()
    zb = torch.stack(zs).to(cuda)
    eb = torch.stack(es).to(cuda)
    zl, el = [zb[i] for i in range(n)], [eb[i] for i in range(n)]
    digests = ops.sha256_leaves(zl, el, prefixes=ts)
    return Trajectory(zl, el, ts, [0.5] * n, z.to(cuda), digests), zs, es


def _reference_open_leaves(zs, es, ts, indices):
    """miner.py:180-264 + 813-842 on the host: leaf_data keyed by timestep, base64 of the requested entries."""
    from oracle import merkle_ref as M
    root, leaf_data = M.commit_trajectory(zs, es, ts)
    out = {}
    for i in indices:
        if i in leaf_data:
            zb, eb, path = leaf_data[i]
            out[i] = (base64.b64encode(zb).decode("ascii"), base64.b64encode(eb).decode("ascii"),
                      [base64.b64encode(p).decode("ascii") for p in path])
        else:
            out[i] = None
    return root, out


def test_open_leaves_equals_reference_and_deletes(cuda):
    from b2v.miner import LeafStore
    traj, zs, es = _trajectory(cuda, seed=3)
    ts = traj.timesteps
    store = LeafStore()
    store.put("hk", "req1", traj)
    ask = [ts[2], ts[3], ts[7], 12345, ts[2]]      # an unknown index and a repeat
    root, want = _reference_open_leaves(zs, es, ts, ask)
    got = store.open_leaves(ask, request_id="req1", caller_hotkey="hk")
    assert got == want and got[12345] is None
    assert len(store) == 0                                           # deleted after one reveal (miner.py:841)
    assert store.open_leaves(ask, request_id="req1", caller_hotkey="hk") == {i: None for i in ask}
    assert store.open_leaves([1, 2], request_id=None, caller_hotkey="hk") == {1: None, 2: None}
    store.put("unknown", "r", traj)                                  # caller_hotkey None -> 'unknown' (miner.py:820)
    assert store.open_leaves([ts[0]], request_id="r") == _reference_open_leaves(zs, es, ts, [ts[0]])[1]
    # the commitment the validator already holds is the same tree
    assert traj.merkle(with_bytes=False)[0] == root


def test_forward_reveal_leaves_synapse_and_spot_check_fast_path(cuda):
    from b2v.miner import LeafStore
    traj, zs, es = _trajectory(cuda, seed=5)
    ts = traj.timesteps
    store = LeafStore()
    store.put("hk", "a", traj)
    syn = types.SimpleNamespace(indices=[str(ts[1]), ts[4]], request_id="a", caller_hotkey="hk", leaves=None)
    out = store.forward_reveal_leaves(syn)
    _, want = _reference_open_leaves(zs, es, ts, [ts[1], ts[4]])
    assert out is syn and syn.leaves == want and len(store) == 0
    syn2 = types.SimpleNamespace(indices=[ts[1]], request_id="a", caller_hotkey="hk", leaves=None)
    assert store.forward_reveal_leaves(syn2).leaves == {ts[1]: None}
    bad = types.SimpleNamespace(indices=["x"], request_id="a", caller_hotkey="hk", leaves=None)
    assert store.forward_reveal_leaves(bad).leaves == {"x": None}    # never raises (miner.py:884-887)
    # fast path of forward (miner.py:352-456): all leaves, proof-shaped dict, deleted afterwards
    store.put("hk", "b", traj)
    resp = store.spot_check_all("hk", "b")
    _, allw = _reference_open_leaves(zs, es, ts, ts)
    assert resp["leaf_data"] == allw and resp["seed"] == 0 and resp["merkle_root"] == "" and len(store) == 0
    assert store.spot_check_all("hk", "b")["leaf_data"] == {} and store.spot_check_all(None, "b")["leaf_data"] == {}


def test_reveal_payload_is_k_over_n_and_verdicts_unchanged(cuda):
    """k = 4 of 12 leaves: a third of the text of ``wire()``; the validator round gives the same verdicts on the partial
    reveal as on the full one, and catches a flipped byte in it."""
    from b2v.miner import LeafStore
    from b2v.validator import Submission, check_miners, check_revealed_leaves, select_spot_check_pairs
    cfg = {"latent_channels": SHAPE[1], "latent_height": SHAPE[3], "latent_width": SHAPE[4]}
    subs, sizes = [], []
    for seed in range(4):
        traj, zs, es = _trajectory(cuda, seed=seed)
        ts = traj.timesteps
        pick = select_spot_check_pairs(ts, 2, random.Random(seed))
        store = LeafStore()
        store.put("v", "r%d" % seed, traj)
        full = traj.wire()
        leaves = store.open_leaves(pick, request_id="r%d" % seed, caller_hotkey="v")
        part = sum(len(v[0]) + len(v[1]) for v in leaves.values())
        whole = sum(len(v[0]) + len(v[1]) for v in full["leaf_data"].values())
        sizes.append((part, whole, len(pick)))
        assert part * len(ts) == whole * len(pick)
        assert all(leaves[t] == tuple(full["leaf_data"][t]) or list(leaves[t]) == list(full["leaf_data"][t]) for t in pick)
        reveal = {str(t): v for t, v in leaves.items()}
        if seed == 2:   # tamper with one revealed byte
            t = str(pick[1])
            raw = bytearray(base64.b64decode(reveal[t][0]))
            raw[9] ^= 0x10
            reveal[t] = (base64.b64encode(bytes(raw)).decode("ascii"), reveal[t][1], reveal[t][2])
        root = base64.b64decode(full["merkle_root"])
        subs.append(Submission("uid%d" % seed, root, ts, traj.alphas, pick, reveal, dict(cfg)))
    verdicts = check_miners(subs)
    assert verdicts == [check_revealed_leaves(s) for s in subs] == [True, True, False, True]
    print("reveal payload (chars) part / whole / k:", sizes)


def test_reveal_at_zeroscope_xl_leaf_size(cuda):
    """k = 5 of 50 XL leaves (1 x 4 x 24 x 72 x 128 fp16 = 1 769 472 B per image): 10 images of 2 359 296 base64
    characters = 23.6 MB of text instead of 236 MB, each identical to the stdlib encoding of the same bytes."""
    from b2v import ops
    from b2v.miner import Trajectory
    n, shape = 50, (1, 4, 24, 72, 128)
    g = torch.Generator(device=cuda).manual_seed(0)
    zb = torch.randn((n,) + shape, device=cuda, generator=g).half()
    eb = torch.randn((n,) + shape, device=cuda, generator=g).half()
    zl, el = [zb[i] for i in range(n)], [eb[i] for i in range(n)]
    ts = [981 - 20 * i for i in range(n)]
    traj = Trajectory(zl, el, ts, [0.5] * n, zl[-1], ops.sha256_leaves(zl, el, prefixes=ts))
    ask = [ts[i] for i in (0, 7, 8, 31, 49)]
    got = traj.reveal(ask)
    total = sum(len(v[0]) + len(v[1]) for v in got.values())
    assert total == 10 * 2359296
    for t in ask:
        i = ts.index(t)
        assert got[t][0] == base64.b64encode(zl[i].cpu().numpy().tobytes()).decode("ascii")
        assert got[t][1] == base64.b64encode(el[i].cpu().numpy().tobytes()).decode("ascii")
        assert len(got[t][2]) == 6   # depth of a 50-leaf tree
