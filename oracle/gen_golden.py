"""Generates tests/golden/*.json by EXECUTING THE REFERENCE (run here, where /root/reference exists; the fixtures
travel, the reference does not):

  proof_golden.json     InferNet/template/validator/proof.py imported by file path with `bittensor` stubbed to a logger.
                        run_unet_step / verify_temporal_coherence_spot_check are recorded for BOTH branches:
                          "fallback"  - diffusers absent (the state of this image): manual formula / False
                          "scheduler" - a stub `diffusers` module whose DDIMScheduler is oracle/ddim_ref.DDIMSchedulerRef
                                        (the reference's control flow, shapes, tolerances are real; the scheduler
                                        arithmetic is the restatement — diffusers itself is not installable here)
                        Miner._build_merkle_tree is lifted from neurons/miner.py:180-264 by source slice (the module
                        imports bittensor/diffusers at top level and cannot be imported).
  strategy_golden.json  the planner (chunk_only.py:80-105, fsdp_chunked_coherent.py:149-177) and blend
                        (chunk_only.py:136-150) source lines, sliced from the files and exec'd on seeded inputs.

usage: python oracle/gen_golden.py   (writes under tests/golden/)
"""
import base64
import hashlib
import importlib.util
import json
import logging
import os
import signal
import sys
import textwrap
import types

import numpy as np
import torch

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tests", "golden")


def b64(b):
    return base64.b64encode(b).decode()


def load_reference_proof(with_scheduler_stub):
    bt = types.ModuleType("bittensor")
    bt.logging = logging.getLogger("ref.bt")
    bt.logging.setLevel(logging.CRITICAL)
    sys.modules["bittensor"] = bt
    sys.modules.pop("diffusers", None)
    spec = importlib.util.spec_from_file_location("ref_proof_%d" % with_scheduler_stub,
                                                  os.path.join(REF, "InferNet/template/validator/proof.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


class branch:
    """Select which branch of the reference's lazy `from diffusers import DDIMScheduler` runs."""

    def __init__(self, name):
        self.name = name

    def __enter__(self):
        if self.name == "scheduler":
            from oracle.ddim_ref import DDIMSchedulerRef
            d = types.ModuleType("diffusers")
            d.DDIMScheduler = DDIMSchedulerRef
            sys.modules["diffusers"] = d
        else:
            sys.modules["diffusers"] = None  # import raises ImportError: the state of this image

    def __exit__(self, *a):
        sys.modules.pop("diffusers", None)


def both(fn):
    out = {}
    for name in ("fallback", "scheduler"):
        with branch(name):
            out[name] = bool(fn())
    return out


def lift_miner_merkle():
    src = open(os.path.join(REF, "InferNet/neurons/miner.py")).read().split("\n")[179:264]
    code = textwrap.dedent("\n".join(src))
    ns = {"hashlib": hashlib, "torch": torch, "time": __import__("time"), "logger": logging.getLogger("ref.miner"),
          "datetime": __import__("datetime").datetime, "List": list, "Tuple": tuple, "Dict": dict}
    ns["logger"].setLevel(logging.CRITICAL)
    from typing import Dict, List, Tuple
    ns.update(List=List, Tuple=Tuple, Dict=Dict)
    exec(code, ns)
    return ns["_build_merkle_tree"]


def trajectory(seed, n, shape, kind="honest"):
    """A seeded default-scheduler DDIM trajectory in fp16 (what an honest miner's leaves look like to the verifier)."""
    from oracle.ddim_ref import DDIMSchedulerRef
    g = torch.Generator().manual_seed(seed)
    s = DDIMSchedulerRef()
    s.set_timesteps(n)
    z = torch.randn(shape, generator=g).half()
    zs, es = [], []
    for t in s.timesteps:
        e = (0.5 * torch.randn(shape, generator=g)).half()
        zs.append(z)
        es.append(e)
        # frame 0 follows the scheduler exactly; other frames too (same op on the whole tensor)
        z = s.step(e, t, z).prev_sample
    if kind == "nan":
        zs[1] = zs[1].clone(); zs[1].view(-1)[3] = float("nan")
    elif kind == "scaled":
        zs[2] = zs[2] * 20
    elif kind == "scaled200":
        zs[2] = zs[2] * 200
    elif kind == "swapped":
        zs[1], zs[2] = zs[2], zs[1]
    elif kind == "inf_eps":
        es[0] = es[0].clone(); es[0].view(-1)[0] = float("inf")
    return zs, es, [int(t) for t in s.timesteps]


def proof_goldens():
    out = {"generated_from": "InferNet/template/validator/proof.py + neurons/miner.py:180-264 @ reference 1e43e8b8"}
    ref_fb = load_reference_proof(False)
    ref_sc = ref_fb  # one module; the branch is chosen per call through sys.modules (see `branch`)
    # --- Merkle ---
    cases = []
    rng = np.random.RandomState(7)
    leafsets = [[bytes([i]) * 32 for i in range(8)], [bytes([i]) * 32 for i in range(5)], [b""], [b"a", b"a"],
                [rng.bytes(int(rng.randint(0, 300))) for _ in range(13)], [rng.bytes(70) for _ in range(50)],
                [rng.bytes(5) for _ in range(3)]]
    for leaves in leafsets:
        root, proofs = ref_fb.build_merkle_root(leaves)
        ok = [ref_fb.verify_merkle_leaf(hashlib.sha256(l).digest(), proofs[i], root) for i, l in enumerate(leaves)]
        bad = [ref_fb.verify_merkle_leaf(hashlib.sha256(l + b"x").digest(), proofs[i], root) for i, l in enumerate(leaves)]
        cases.append({"leaves": [b64(l) for l in leaves], "root": root.hex(),
                      "proofs": [[p.hex() for p in pr] for pr in proofs], "verify": ok, "verify_tampered": bad})
    out["merkle"] = cases
    out["merkle_empty"] = {"root": ref_fb.build_merkle_root([])[0].hex(), "proofs": ref_fb.build_merkle_root([])[1]}
    out["derive_seed"] = [{"challenge": b64(c), "wallet": w, "seed": ref_fb.derive_seed(c, w)}
                          for c, w in [(b"\x01" * 32, "5Fwallet"), (b"", "x"), (bytes(range(32)), "5GrwvaEF5zXb26Fz9rcQpDWS57CtERHpNehXCPcNoHGKutQY")]]
    out["spotcheck"] = [{"n": n, "k": k, "seed": s, "indices": ref_fb.commit_then_reveal_merkle_spotcheck(b"", n, k, s)}
                        for n, k, s in [(50, 5, 1234), (8, 3, 0), (2, 5, 9), (1, 3, 1), (0, 2, 1), (25, 30, 77)]]
    # --- miner commit ---
    build = lift_miner_merkle()
    dummy = types.SimpleNamespace(metrics={"errors": []})
    commits = []
    for seed, n, shape in [(1, 3, (1, 4, 2, 4, 4)), (2, 8, (1, 4, 3, 8, 8)), (3, 5, (1, 4, 1, 2, 2)), (4, 1, (1, 4, 2, 2, 2))]:
        g = torch.Generator().manual_seed(seed)
        zs = [torch.randn(shape, generator=g).half() for _ in range(n)]
        es = [torch.randn(shape, generator=g).half() for _ in range(n)]
        ts = [1000 - 37 * i - seed for i in range(n)]
        root, leaf_data = build(dummy, zs, es, ts)
        commits.append({"seed": seed, "n": n, "shape": list(shape), "timesteps": ts, "root": root.hex(),
                        "paths": {str(t): [p.hex() for p in leaf_data[t][2]] for t in ts},
                        "z0_sha": hashlib.sha256(leaf_data[ts[0]][0]).hexdigest()})
    out["miner_commit"] = commits
    # --- run_unet_step / temporal coherence, both branches ---
    cfg = {"latent_channels": 4, "latent_height": 8, "latent_width": 8}
    steps = []
    for kind in ["honest", "nan", "scaled", "scaled200", "swapped", "inf_eps"]:
        for shape in [(1, 4, 3, 8, 8), (4, 8, 8)]:
            zs, es, ts = trajectory(11, 8, shape, kind)
            alphas = [0.5] * 8
            c = dict(cfg, alphas=alphas)
            zb = [z.numpy().tobytes() for z in zs]
            eb = [e.numpy().tobytes() for e in es]
            rec = {"kind": kind, "shape": list(shape), "z": [b64(x) for x in zb], "eps": [b64(x) for x in eb],
                   "alphas": alphas,
                   "run_unet_step": {}}
            per = [both(lambda i=i: ref_fb.run_unet_step(zb[i], eb[i], c, i)) for i in range(8)]
            rec["run_unet_step"] = {k: [p[k] for p in per] for k in ("fallback", "scheduler")}
            t_to_step = {t: i for i, t in enumerate(ts)}
            leaves = {str(t): (b64(zb[i]), b64(eb[i]), []) for i, t in enumerate(ts)}
            coh = {}
            for name, idx in [("all", ts), ("pair01", ts[0:2]), ("pair23", ts[2:4]), ("single", ts[:1]),
                              ("pairs_0_1_4_5", [ts[0], ts[1], ts[4], ts[5]])]:
                coh[name] = dict(both(lambda idx=idx: ref_fb.verify_temporal_coherence_spot_check(
                    idx, leaves, t_to_step, alphas, cfg, "g", QUIET)), indices=idx)
            rec["coherence"] = coh
            steps.append(rec)
    # malformed inputs
    zs, es, ts = trajectory(5, 4, (4, 8, 8))
    c = dict(cfg, alphas=[0.9] * 4)
    mal = []
    for name, z, e, i in [("step_out_of_range", zs[0].numpy().tobytes(), es[0].numpy().tobytes(), 4),
                          ("short_z", zs[0].numpy().tobytes()[:100], es[0].numpy().tobytes(), 0),
                          ("odd_bytes", zs[0].numpy().tobytes()[:-1], es[0].numpy().tobytes()[:-1], 0),
                          ("eps_shorter", zs[0].numpy().tobytes() * 2, es[0].numpy().tobytes(), 1),
                          ("not_multiple", zs[0].numpy().tobytes() + b"\0" * 10, es[0].numpy().tobytes() + b"\0" * 10, 1)]:
        mal.append(dict(both(lambda: ref_fb.run_unet_step(z, e, c, i)), name=name, z=b64(z), eps=b64(e), step=i,
                        alphas=c["alphas"]))
    out["ddim_checks"] = steps
    out["ddim_malformed"] = mal
    out["unet_config"] = cfg
    return out


QUIET = logging.getLogger("ref.quiet")
QUIET.setLevel(logging.CRITICAL)
QUIET.propagate = False


class _Timeout(Exception):
    pass


def _alarm(*_):
    raise _Timeout()


def lift_planner(path, first, last):
    lines = open(os.path.join(REF, path)).read().split("\n")[first - 1:last]
    body = textwrap.dedent("\n".join(lines))
    src = "def _plan(cfg, self, T):\n" + textwrap.indent(body, "    ") + "\n    return cs, ov, ranges\n"
    ns = {}
    exec(src, ns)
    return ns["_plan"]


def lift_blend():
    lines = open(os.path.join(REF, "Distribution/strategies/chunk_only.py")).read().split("\n")[135:150]
    body = textwrap.dedent("\n".join(lines))
    src = "def _blend(base, T, ov, gathered):\n" + textwrap.indent(body, "    ") + "\n    return lat, full, weight\n"
    ns = {"torch": torch}
    exec(src, ns)
    return ns["_blend"]


def strategy_goldens():
    out = {"generated_from": "Distribution/strategies/chunk_only.py:80-105,136-150; fsdp_chunked_coherent.py:149-177"}
    plan_a = lift_planner("Distribution/strategies/chunk_only.py", 80, 105)
    plan_b = lift_planner("Distribution/strategies/fsdp_chunked_coherent.py", 149, 177)
    plans = []
    signal.signal(signal.SIGALRM, _alarm)
    for variant, fn in (("chunk_only", plan_a), ("coherent", plan_b)):
        for T in (8, 16, 24, 25, 32, 96):
            for world in (1, 2, 3, 4, 6, 8):
                for cs, ov in ((0, 4), (0, 0), (0, 2), (4, 1), (16, 4), (6, 4), (5, 8), (10, 3)):
                    cfg = types.SimpleNamespace(chunk_size=cs, overlap=ov, no_chunking=False)
                    me = types.SimpleNamespace(world=world)
                    rec = {"variant": variant, "T": T, "world": world, "chunk_size": cs, "overlap": ov}
                    signal.alarm(1)
                    try:
                        a, b, r = fn(cfg, me, T)
                        rec.update(cs=a, ov=b, ranges=[list(x) for x in r])
                    except _Timeout:
                        rec["hangs"] = True
                    except Exception as e:  # e.g. ZeroDivision / range errors in the reference
                        rec["raises"] = type(e).__name__
                    finally:
                        signal.alarm(0)
                    plans.append(rec)
    out["plans"] = plans
    blend = lift_blend()
    blends = []
    for T, world, ranges, ov, dt in [(8, 2, [(0, 5), (4, 8)], 1, "float16"), (24, 2, [(0, 16), (12, 24)], 4, "float16"),
                                     (24, 4, [(0, 8), (6, 14), (12, 20), (18, 24)], 2, "float16"),
                                     (24, 4, [(0, 6), (2, 8), (4, 10), (6, 12), (8, 14), (10, 16), (12, 18), (14, 20),
                                              (16, 22), (18, 24), (20, 24), (22, 24)], 4, "float16"),
                                     (10, 1, [(0, 10)], 0, "float16"), (9, 3, [(0, 3), (2, 5), (4, 7), (6, 9), (8, 9), (8, 9)], 1, "bfloat16"),
                                     (12, 2, [(0, 8), (4, 12)], 4, "float32")]:
        dtype = getattr(torch, dt)
        g = torch.Generator().manual_seed(T * 100 + world)
        base = torch.randn(1, 4, T, 5, 7, generator=g).to(dtype)
        gathered = []
        for r in range(world):
            mine = [(s, e, (3 * torch.randn(1, 4, e - s, 5, 7, generator=g)).to(dtype))
                    for i, (s, e) in enumerate(ranges) if i % world == r]
            gathered.append(mine)
        lat, full, weight = blend(base, T, ov, gathered)
        blends.append({"T": T, "world": world, "ranges": [list(r) for r in ranges], "ov": ov, "dtype": dt,
                       "seed": T * 100 + world, "lat_sha256": hashlib.sha256(lat.numpy().tobytes()).hexdigest(),
                       "lat_dtype": str(lat.dtype), "weight": weight.flatten().tolist(),
                       "lat_sample": lat.flatten()[::97][:40].tolist()})
    out["blends"] = blends
    return out


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    with open(os.path.join(OUT, "proof_golden.json"), "w") as f:
        json.dump(proof_goldens(), f)
    with open(os.path.join(OUT, "strategy_golden.json"), "w") as f:
        json.dump(strategy_goldens(), f)
    for n in ("proof_golden.json", "strategy_golden.json"):
        print(n, os.path.getsize(os.path.join(OUT, n)))
